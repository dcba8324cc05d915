#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/tc_stages.py > gpurun_out/tc_stages.log 2>&1; echo "stages rc=$?"; cat gpurun_out/tc_stages.log | tail -50
timeout 600 python tools/bench_kernels.py dw > gpurun_out/kernels_dw.log 2>&1; echo "dw rc=$?"; grep dw_sweep gpurun_out/kernels_dw.log | cut -c1-260
timeout 400 python tools/bench_kernels.py gemm gpurun_out/kernels_gemm.json > gpurun_out/kernels_gemm.log 2>&1; echo "gemm bench rc=$?"
timeout 300 python -m pytest tests/test_policy_api.py -q -m gpu -k "open_loop or rollout" --no-header -p no:cacheprovider 2>&1 | tail -3
