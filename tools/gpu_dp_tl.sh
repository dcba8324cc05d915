#!/bin/bash
NP=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1"
ACT_DP_P2P=1 timeout 300 $TR --master-port 29700 tools/timeline.py > gpurun_out/timeline_dp.log 2>&1; tail -2 gpurun_out/timeline_dp.log
python tools/timeline_tail.py 600 > gpurun_out/timeline_tail_n$NP.txt; head -80 gpurun_out/timeline_tail_n$NP.txt
python tools/timeline_summary.py > gpurun_out/timeline_summary_n$NP.md
