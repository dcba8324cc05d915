#!/bin/bash
# A/B of environment switches on the N-GPU weak-scaling bench line: tools/gpu_dp_ab.sh N "VAR=a" "VAR=b" ...
NP=${1:-2}; shift
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1"
port=29600
for v in "$@"; do
  port=$((port+1))
  env $v timeout 400 $TR --master-port $port bench.py --gpus $NP --steps 50 --warmup 5 --no-infer --no-cpu-baseline > gpurun_out/bench_ab.json 2> gpurun_out/bench_ab.err
  echo "$v rc=$? $(python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_ab.json') if l.startswith('{')][0])
print('N',d['n_gpus'],'ms/step',round(d['ms_per_step'],3),'samples/s',round(d['value'],1),'e2e',round(d['e2e']['value'],1))
")"
done
