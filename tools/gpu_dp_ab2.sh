#!/bin/bash
# N-rank A/B of exchange variants (bench only): each argument after N is an "ENV=VAL[,ENV=VAL]" set; "-" = defaults
NP=${1:-8}; shift
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1"
ACT_DP_P2P=1 timeout 150 $TR --master-port 29521 tools/dp_check.py > gpurun_out/dp_check_ab_n$NP.log 2>&1; echo "dp_check rc=$?"; grep -E "dp_check (exchange|replicas|p2p forward)" gpurun_out/dp_check_ab_n$NP.log | tail -3
port=29530
for cfg in "$@"; do
  port=$((port+1))
  envs=""; [ "$cfg" != "-" ] && envs=$(echo "$cfg" | tr ',' ' ')
  tag=$(echo "$cfg" | tr -c 'A-Za-z0-9=_\n' '_')
  env $envs timeout 240 $TR --master-port $port bench.py --gpus $NP --steps 50 --warmup 5 --no-infer --no-policy-api --no-cpu-baseline > gpurun_out/bench_n${NP}_$tag.json 2> gpurun_out/bench_n${NP}_$tag.err
  echo "[$cfg] rc=$? $(python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_n${NP}_$tag.json') if l.startswith('{')][0])
print('ms/step',round(d['ms_per_step'],3),'samples/s',round(d['value'],1),'e2e',round(d['e2e']['value'],1),d['notes']['grad_exchange'][:40])
")"
done
