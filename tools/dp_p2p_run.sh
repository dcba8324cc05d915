#!/bin/bash
# 2..N-GPU check + bench of the peer-memory data-parallel step (run under gpurun --gpus N): tools/dp_p2p_run.sh N
NP=${1:-2}
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=${CUDA_DEVICE_MAX_CONNECTIONS:-32}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
echo "== single-GPU ghost-peer test"; timeout 300 python -m pytest tests/test_dp_p2p.py -q -m gpu --no-header -p no:cacheprovider 2>&1 | tail -5
echo "== dp_check NCCL"; timeout 200 $TR --master-port 29511 tools/dp_check.py 2>&1 | grep dp_check
echo "== dp_check P2P"; ACT_DP_P2P=1 timeout 200 $TR --master-port 29512 tools/dp_check.py > gpurun_out/dp_check_p2p.log 2>&1; echo "rc=$?"; grep -E "dp_check|Error|error" gpurun_out/dp_check_p2p.log | head -20
echo "== bench NCCL"; timeout 300 $TR --master-port 29513 bench.py --gpus $NP --steps 20 --warmup 5 > gpurun_out/bench_n${NP}_nccl.json 2> gpurun_out/bench_nccl.err; echo "rc=$?"; cut -c1-160 gpurun_out/bench_n${NP}_nccl.json
echo "== bench P2P"; ACT_DP_P2P=1 timeout 300 $TR --master-port 29514 bench.py --gpus $NP --steps 20 --warmup 5 > gpurun_out/bench_n${NP}_p2p.json 2> gpurun_out/bench_p2p.err; echo "rc=$?"; cut -c1-160 gpurun_out/bench_n${NP}_p2p.json; tail -5 gpurun_out/bench_p2p.err
echo "== bench P2P no graph"; ACT_DP_P2P=1 timeout 300 $TR --master-port 29515 bench.py --gpus $NP --steps 20 --warmup 5 --no-graph > gpurun_out/bench_n${NP}_p2p_nograph.json 2> gpurun_out/bench_p2p_ng.err; echo "rc=$?"; cut -c1-160 gpurun_out/bench_n${NP}_p2p_nograph.json
