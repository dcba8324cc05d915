#!/bin/bash
mkdir -p gpurun_out
tools/gpu_tests.sh tests/test_engine_golden.py tests/test_dp_p2p.py tests/test_policy_api.py
for mode in "ACT_BB_EARLY=1" "ACT_BB_EARLY=0"; do
  env $mode timeout 600 python bench.py --steps 30 --no-cpu-baseline --no-infer --no-policy-api > gpurun_out/bench_$mode.json 2> gpurun_out/bench_$mode.err
  echo "$mode rc=$? $(python -c "
import json,sys
d=json.loads([l for l in open('gpurun_out/bench_$mode.json') if l.startswith('{')][0])
print('ms/step',round(d['ms_per_step'],3),'samples/s',round(d['value'],1),'e2e',round(d['e2e']['value'],1))
")"
done
