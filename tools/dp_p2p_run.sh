#!/bin/bash
# N-GPU check + bench of the peer-memory data-parallel step against the NCCL path on the same box
# (run under gpurun --gpus N): tools/dp_p2p_run.sh N
NP=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1"
echo "== dp_check P2P"; ACT_DP_P2P=1 timeout 200 $TR --master-port 29512 tools/dp_check.py > gpurun_out/dp_check_p2p_n$NP.log 2>&1; echo "rc=$?"; grep -E "dp_check" gpurun_out/dp_check_p2p_n$NP.log | grep -v "rank [1-9]" | head -20
echo "== bench NCCL"; timeout 300 $TR --master-port 29513 bench.py --gpus $NP --steps 20 --warmup 5 > gpurun_out/bench_n${NP}_nccl.json 2> gpurun_out/bench_nccl.err; echo "rc=$?"; cut -c1-160 gpurun_out/bench_n${NP}_nccl.json
echo "== bench P2P"; ACT_DP_P2P=1 timeout 300 $TR --master-port 29514 bench.py --gpus $NP --steps 20 --warmup 5 > gpurun_out/bench_n${NP}_p2p.json 2> gpurun_out/bench_p2p.err; echo "rc=$?"; cut -c1-160 gpurun_out/bench_n${NP}_p2p.json; tail -3 gpurun_out/bench_p2p.err
