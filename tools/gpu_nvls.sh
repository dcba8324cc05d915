#!/bin/bash
# NVLS (multicast) exchange: probe of the mapping, parity, A/B of the weak-scaling step with and without multicast
NP=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1"
timeout 120 $TR --master-port 29520 tools/nvls_probe.py > gpurun_out/nvls_probe_n$NP.log 2>&1; echo "probe rc=$?"; grep -E "rank 0|nvls_probe" gpurun_out/nvls_probe_n$NP.log | tail -8
ACT_DP_P2P=1 timeout 150 $TR --master-port 29521 tools/dp_check.py > gpurun_out/dp_check_nvls_n$NP.log 2>&1; echo "dp_check rc=$?"; grep -E "dp_check" gpurun_out/dp_check_nvls_n$NP.log | grep -v "rank [1-9]" | tail -9
for nv in 1 0; do
  ACT_DP_NVLS=$nv timeout 240 $TR --master-port 2952$((2+nv)) bench.py --gpus $NP --steps 50 --warmup 5 --no-infer --no-policy-api --no-cpu-baseline > gpurun_out/bench_n${NP}_nvls$nv.json 2> gpurun_out/bench_n${NP}_nvls$nv.err
  echo "bench NVLS=$nv rc=$? $(python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_n${NP}_nvls$nv.json') if l.startswith('{')][0])
print('ms/step',round(d['ms_per_step'],3),'samples/s',round(d['value'],1),'e2e',round(d['e2e']['value'],1),d['notes']['grad_exchange'][:60])
")"
done
tail -3 gpurun_out/bench_n${NP}_nvls1.err
