#!/bin/bash
mkdir -p gpurun_out
tools/gpu_tests.sh tests/test_engine_golden.py tests/test_policy_api.py tests/test_fused_kernels.py tests/test_dropout.py
timeout 900 python bench.py --eager-gpu > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_n1.json; tail -3 gpurun_out/bench_n1.err
timeout 300 python tools/timeline.py > gpurun_out/timeline.log 2>&1; echo "timeline rc=$?"; tail -5 gpurun_out/timeline.log
python tools/timeline_summary.py > gpurun_out/timeline_summary.md 2>&1; head -5 gpurun_out/timeline_summary.md
