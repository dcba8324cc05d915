#!/bin/bash
# maxpool backward rewrite + pipelined loss read-back of the policy-API leg: tests, bench, timeline
mkdir -p gpurun_out
tools/gpu_tests.sh tests/test_fused_kernels.py tests/test_engine_golden.py tests/test_policy_api.py
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_n1b.json 2> gpurun_out/bench_n1b.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_n1b.json') if l.startswith('{')][0])
print('ms/step',round(d['ms_per_step'],3),'samples/s',round(d['value'],1),'e2e',round(d['e2e']['value'],1))
for k in ('e2e_policy_api','e2e_policy_api_unfused'):
    if k in d: print(k, {a:(round(b,2) if isinstance(b,float) else b) for a,b in d[k].items()})
print('roofline',round(d['roofline']['achieved'],1),round(d['roofline']['frac'],3), {k:round(v['ms'],3) for k,v in d['roofline']['by_kernel_class'].items()})
inf=d.get('inference',{})
print('infer',round(inf.get('chunks_per_s',0)),'e2e',round(inf.get('e2e',{}).get('value',0)),'p50',inf.get('p50_latency_ms_b1'),inf.get('p50_latency_ms_b1_direct_launch'))
PY
timeout 300 python tools/timeline.py 8 > gpurun_out/timeline_n1.log 2>&1; tail -1 gpurun_out/timeline_n1.log
python tools/timeline_summary.py > gpurun_out/timeline_summary_b8.md 2>&1; head -4 gpurun_out/timeline_summary_b8.md
python - <<'PY'
import json,re
d=json.load(open('gpurun_out/timeline.json'))
for k in d:
    if 'maxpool' in k['name'] or 'wgrad_fold' in k['name']: print(re.sub(r'\(.*','',k['name']), round(k['ts'],1), round(k['dur'],1))
PY
