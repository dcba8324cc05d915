#!/bin/bash
# round 2: no-alloc entry point test (alone, then after the whole suite in one process), N=1 bench + timeline, ncu evidence
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu --no-header -p no:cacheprovider > gpurun_out/tests_zz_alone.log 2>&1
echo "zz alone rc=$? $(tail -1 gpurun_out/tests_zz_alone.log)"
timeout 1500 python -m pytest tests -x -q -m gpu --no-header -p no:cacheprovider > gpurun_out/tests_all.log 2>&1
echo "all rc=$? $(tail -1 gpurun_out/tests_all.log)"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$? $(tail -1 gpurun_out/smoke.log)"
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; grep '^{' gpurun_out/bench_n1.json | cut -c1-600
timeout 300 python tools/timeline.py 8 > gpurun_out/timeline_n1.log 2>&1; tail -2 gpurun_out/timeline_n1.log
python tools/timeline_summary.py > gpurun_out/timeline_summary_b8.md 2>&1; head -30 gpurun_out/timeline_summary_b8.md
tools/gpu_profile_r02.sh
