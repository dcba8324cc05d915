# N-GPU data-parallel check + bench variants (development helper; run under gpurun --gpus N): tools/dp_run.sh N "VAR=1" ...
NP=${1:-2}; shift
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29511 tools/dp_check.py 2>&1 | grep dp_check
port=29530
for v in "A=0" "$@"; do
  port=$((port+1))
  echo "--- $v"; env $v timeout 200 $TR --master-port $port bench.py --gpus $NP --steps 20 --warmup 5 2>gpurun_out/dp_v.err | tail -1 | cut -c1-140
done
