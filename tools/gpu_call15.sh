#!/bin/bash
mkdir -p gpurun_out
tools/gpu_tests.sh tests/test_engine_golden.py tests/test_dropout.py tests/test_fused_kernels.py
# --set full captures of the GEMM template: FFN up-projection inside the 128-env inference step, one training-step GEMM
python tools/profile_infer.py 128 2 > gpurun_out/profile_infer.log 2>&1 && timeout 300 ncu --kernel-name-base demangled -k regex:"tc_kernel<.int.0, .int.0, .int.0, .int.256>" -s 6 -c 3 --set full --import-source on --clock-control none -f -o gpurun_out/ncu_full_gemm_infer python tools/profile_infer.py 128 2 > gpurun_out/ncu_full_gemm.log 2>&1; echo "ncu full gemm infer rc=$?"
python tools/profile_step.py 8 2 gpurun_out/tc_train_b8.csv > gpurun_out/profile_train.log 2>&1 && timeout 300 ncu --kernel-name-base demangled -k regex:"tc_kernel<.int.0, .int.0, .int.0, .int.128>" -s 14 -c 4 --set full --import-source on --clock-control none -f -o gpurun_out/ncu_full_gemm_train python tools/profile_step.py 8 2 > gpurun_out/ncu_full_gemm_train.log 2>&1; echo "ncu full gemm train rc=$?"
ls -la gpurun_out/*.ncu-rep | tail -4
