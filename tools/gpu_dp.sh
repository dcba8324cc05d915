#!/bin/bash
# N-GPU visit (run under gpurun --gpus N): parity of both exchange paths, the default (weak) bench line with inference on
# every rank, the fixed-global-batch (strong) line of BASELINE configs[2], and optionally the 3-camera configs[3] line.
NP=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1"
ACT_DP_P2P=1 timeout 300 $TR --master-port 29512 tools/dp_check.py > gpurun_out/dp_check_p2p_n$NP.log 2>&1; echo "dp_check p2p rc=$?"; grep -E "dp_check" gpurun_out/dp_check_p2p_n$NP.log | grep -v "rank [1-9]" | tail -8
ACT_DP_P2P=0 timeout 300 $TR --master-port 29513 tools/dp_check.py > gpurun_out/dp_check_nccl_n$NP.log 2>&1; echo "dp_check nccl rc=$?"; grep -E "dp_check" gpurun_out/dp_check_nccl_n$NP.log | tail -4
timeout 600 $TR --master-port 29514 bench.py --gpus $NP --steps 50 --warmup 5 > gpurun_out/bench_n${NP}.json 2> gpurun_out/bench_n${NP}.err; echo "bench weak rc=$?"; cut -c1-200 gpurun_out/bench_n${NP}.json; tail -2 gpurun_out/bench_n${NP}.err
timeout 600 $TR --master-port 29515 bench.py --gpus $NP --steps 50 --warmup 5 --global-batch 64 --no-infer > gpurun_out/bench_n${NP}_gb64.json 2> gpurun_out/bench_n${NP}_gb64.err; echo "bench strong rc=$?"; cut -c1-200 gpurun_out/bench_n${NP}_gb64.json
if [ "$2" = "k3" ]; then
  timeout 600 $TR --master-port 29516 bench.py --gpus $NP --steps 30 --warmup 5 --cameras 3 --batch 16 --no-infer > gpurun_out/bench_n${NP}_k3_b16.json 2> gpurun_out/bench_n${NP}_k3_b16.err; echo "bench k3 rc=$?"; cut -c1-200 gpurun_out/bench_n${NP}_k3_b16.json
fi
