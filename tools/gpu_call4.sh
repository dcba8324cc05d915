#!/bin/bash
mkdir -p gpurun_out
tools/gpu_tests.sh tests/test_engine_golden.py tests/test_tc_gemm.py tests/test_conv.py
for mode in "ACT_PDL_LATE=1" "ACT_PDL_LATE=0" "ACT_PDL=0"; do
  env $mode timeout 600 python bench.py --steps 30 --no-cpu-baseline --no-infer --no-policy-api > gpurun_out/bench_$mode.json 2> gpurun_out/bench_$mode.err
  echo "$mode rc=$? $(python -c "
import json,sys
d=json.loads([l for l in open('gpurun_out/bench_$mode.json') if l.startswith('{')][0])
r=d['roofline']
print('ms/step',round(d['ms_per_step'],3),'samples/s',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'serial',round(r.get('serial_step_ms',0),3),'tensor_ms',round(r.get('tensor_ms_per_step',0),3),'TF',round(r.get('achieved',0),1), {k:round(v['ms'],3) for k,v in r.get('by_kernel_class',{}).items()})
")"
done
ACT_PDL_LATE=1 timeout 300 python tools/tc_stages.py 816 512 512 > gpurun_out/tc_stages_late.log 2>&1; tail -12 gpurun_out/tc_stages_late.log
ACT_PDL_LATE=1 timeout 600 python bench.py --steps 10 --no-cpu-baseline --no-policy-api --no-graph > /dev/null 2>&1
ACT_PDL_LATE=1 timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-policy-api --infer-envs 128 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][0]); i=d['inference']
print('infer late', i['chunks_per_s'], i['e2e']['value'], i['roofline'].get('achieved'), {k:round(v['ms'],3) for k,v in i['roofline'].get('by_kernel_class',{}).items()})"
ACT_PDL_LATE=0 timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-policy-api --infer-envs 128 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][0]); i=d['inference']
print('infer early', i['chunks_per_s'], i['e2e']['value'], i['roofline'].get('achieved'), {k:round(v['ms'],3) for k,v in i['roofline'].get('by_kernel_class',{}).items()})"
