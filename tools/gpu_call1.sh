#!/bin/bash
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > gpurun_out/gpu.txt 2>&1
tools/gpu_tests.sh
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$? $(tail -1 gpurun_out/smoke.log)"
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_n1.json
timeout 400 python tools/bench_kernels.py gemm gpurun_out/kernels_gemm.json > gpurun_out/kernels_gemm.log 2>&1; echo "gemm bench rc=$?"
ACT_TC_DEBUG=3 timeout 300 python tools/bench_kernels.py floor > gpurun_out/kernels_floor.log 2>&1; echo "floor rc=$?"; grep floor gpurun_out/kernels_floor.log | head -12
timeout 300 python tools/bench_kernels.py conv gpurun_out/kernels_conv.json > gpurun_out/kernels_conv.log 2>&1; echo "conv bench rc=$?"
