#!/bin/bash
# N ranks: timeline of the data-parallel step (rank 0) with the default exchange for that N
NP=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1"
ACT_DP_P2P=1 timeout 240 $TR --master-port 29700 tools/timeline.py > gpurun_out/timeline_dp.log 2>&1; tail -2 gpurun_out/timeline_dp.log
cp gpurun_out/timeline.json gpurun_out/timeline_n$NP.json
python tools/timeline_tail.py 900 > gpurun_out/timeline_tail_n$NP.txt; grep -E "dp_|span" gpurun_out/timeline_tail_n$NP.txt | head -40
python tools/timeline_summary.py > gpurun_out/timeline_summary_n$NP.md
