#!/bin/bash
mkdir -p gpurun_out
for mode in "ACT_TC_SHARE=0" "ACT_TC_SHARE=1"; do
  env $mode timeout 600 python bench.py --steps 30 --no-cpu-baseline --no-infer --no-policy-api > gpurun_out/bench_$mode.json 2> gpurun_out/bench_$mode.err
  echo "$mode rc=$? $(python -c "
import json,sys
d=json.loads([l for l in open('gpurun_out/bench_$mode.json') if l.startswith('{')][0])
r=d['roofline']
print('ms/step',round(d['ms_per_step'],3),'samples/s',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'serial',round(r.get('serial_step_ms',0),3),'tensor_ms',round(r.get('tensor_ms_per_step',0),3),'TF',round(r.get('achieved',0),1), {k:round(v['ms'],3) for k,v in r.get('by_kernel_class',{}).items()})
")"
done
ACT_TC_SHARE=1 tools/gpu_tests.sh tests/test_engine_golden.py tests/test_tc_gemm.py
