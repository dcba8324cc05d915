#!/bin/bash
# final single-GPU visit: the driver's own sequence (pytest -x in one process, smoke, bench) + the 3-camera A/B
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu --no-header -p no:cacheprovider > gpurun_out/tests_all.log 2>&1; echo "all tests rc=$? $(tail -1 gpurun_out/tests_all.log)"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$? $(tail -1 gpurun_out/smoke.log)"
timeout 900 python bench.py > gpurun_out/bench_n1_final.json 2> gpurun_out/bench_n1_final.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_n1_final.json') if l.startswith('{')][0])
print('ms/step',round(d['ms_per_step'],3),'samples/s',round(d['value'],1),'e2e',round(d['e2e']['value'],1), 'launches', d['gpu_launches'], 'clocks', d['clocks'])
for k in ('e2e_policy_api','e2e_policy_api_unfused'):
    if k in d: print(k, round(d[k]['value'],1), round(d[k]['ms_per_step'],3))
print('roofline',round(d['roofline']['achieved'],1),round(d['roofline']['frac'],3), d['roofline']['traffic'], {k:round(v['ms'],3) for k,v in d['roofline']['by_kernel_class'].items()})
inf=d.get('inference',{})
print('infer',round(inf.get('chunks_per_s',0)),'e2e',round(inf.get('e2e',{}).get('value',0)),'p50',inf.get('p50_latency_ms_b1'),'roof', inf.get('roofline',{}).get('frac'))
print('cpu', d.get('cpu_baseline',{}).get('value'))
PY
for m in 0 1; do
ACT_ATTN_CHUNKED=$m timeout 600 python bench.py --cameras 3 --batch 16 --steps 30 --no-infer --no-policy-api --no-cpu-baseline > gpurun_out/bench_n1_k3_b16_c$m.json 2> gpurun_out/bench_n1_k3_b16_c$m.err; echo "k3 chunked=$m rc=$? $(grep '^{' gpurun_out/bench_n1_k3_b16_c$m.json | cut -c1-150)"
done
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference rc=$? $(cut -c1-200 gpurun_out/bench_reference.json)"
