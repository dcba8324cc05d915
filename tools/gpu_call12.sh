#!/bin/bash
# N=1: tests, default bench, 3-camera / batch-64 lines
mkdir -p gpurun_out
tools/gpu_tests.sh tests/test_fused_kernels.py tests/test_engine_golden.py tests/test_policy_api.py tests/test_dropout.py
timeout 900 python bench.py > gpurun_out/bench_n1c.json 2> gpurun_out/bench_n1c.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_n1c.json') if l.startswith('{')][0])
print('ms/step',round(d['ms_per_step'],3),'samples/s',round(d['value'],1),'e2e',round(d['e2e']['value'],1))
for k in ('e2e_policy_api','e2e_policy_api_unfused'):
    if k in d: print(k, round(d[k]['value'],1), round(d[k]['ms_per_step'],3))
print('roofline',round(d['roofline']['achieved'],1),round(d['roofline']['frac'],3), {k:round(v['ms'],3) for k,v in d['roofline']['by_kernel_class'].items()})
inf=d.get('inference',{})
print('infer',round(inf.get('chunks_per_s',0)),'e2e',round(inf.get('e2e',{}).get('value',0)),'p50',inf.get('p50_latency_ms_b1'),inf.get('p50_latency_ms_b1_direct_launch'), 'roof', inf.get('roofline',{}).get('frac'))
print('cpu', d.get('cpu_baseline',{}).get('value'))
PY
timeout 600 python bench.py --cameras 3 --batch 16 --steps 30 --no-infer --no-policy-api --no-cpu-baseline > gpurun_out/bench_n1_k3_b16.json 2> gpurun_out/bench_n1_k3_b16.err; echo "k3 rc=$? $(cut -c1-160 gpurun_out/bench_n1_k3_b16.json | grep '^{')"
timeout 600 python bench.py --batch 64 --steps 30 --no-infer --no-policy-api --no-cpu-baseline > gpurun_out/bench_n1_b64.json 2> gpurun_out/bench_n1_b64.err; echo "b64 rc=$? $(cut -c1-160 gpurun_out/bench_n1_b64.json | grep '^{')"
