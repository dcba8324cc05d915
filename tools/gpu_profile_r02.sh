#!/bin/bash
# ncu evidence for round 2 (one GPU; every profiled command first runs WITHOUT ncu and must exit 0).  Outputs: gpurun_out/.
set -u
mkdir -p gpurun_out
python tools/profile_step.py 8 3 gpurun_out/tc_train_b8.csv > gpurun_out/profile_train.log 2>&1 || { echo "train plain run failed"; tail -5 gpurun_out/profile_train.log; exit 1; }
python tools/profile_infer.py 128 3 gpurun_out/tc_infer_e128.csv > gpurun_out/profile_infer.log 2>&1 || { echo "infer plain run failed"; tail -5 gpurun_out/profile_infer.log; exit 1; }
tail -2 gpurun_out/profile_train.log; tail -2 gpurun_out/profile_infer.log
# 1. launch lists (per-launch durations, cold cache + serialised: shares, not absolutes)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/ncu_train_b8.csv python tools/profile_step.py 8 2 > gpurun_out/ncu_train.log 2>&1; echo "ncu train list rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/ncu_infer_e128.csv python tools/profile_infer.py 128 2 > gpurun_out/ncu_infer.log 2>&1; echo "ncu infer list rc=$?"
# 2. DRAM traffic of the tensor-core kernels inside a WARM step (no cache flush between launches)
timeout 900 ncu --cache-control none --clock-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
    -k regex:"tc_kernel|conv3_kernel|attn_scores" --csv --log-file gpurun_out/ncu_tc_dram_warm_b8.csv python tools/profile_step.py 8 3 > gpurun_out/ncu_dram.log 2>&1; echo "ncu dram warm rc=$?"
# 3. --set full + source: the halo-resident conv of the inference step (layer1 pair conv and a 128-channel conv), the
#    FFN up-projection GEMM (N = 256 tiles), and one training-step GEMM at batch 8
timeout 900 ncu -k regex:conv3_kernel -s 13 -c 6 --set full --import-source on --clock-control none -f -o gpurun_out/ncu_full_conv3_infer python tools/profile_infer.py 128 2 > gpurun_out/ncu_full_conv3.log 2>&1; echo "ncu full conv3 rc=$?"
timeout 900 ncu --kernel-name-base demangled -k regex:"tc_kernel<.int.0, .int.0, .int.0, .int.256>" -s 6 -c 3 --set full --import-source on --clock-control none -f -o gpurun_out/ncu_full_gemm_infer python tools/profile_infer.py 128 2 > gpurun_out/ncu_full_gemm.log 2>&1; echo "ncu full gemm rc=$?"
timeout 900 ncu --kernel-name-base demangled -k regex:"tc_kernel<.int.0, .int.0, .int.0, .int.128>" -s 14 -c 4 --set full --import-source on --clock-control none -f -o gpurun_out/ncu_full_gemm_train python tools/profile_step.py 8 2 > gpurun_out/ncu_full_gemm_train.log 2>&1; echo "ncu full gemm train rc=$?"
ls -la gpurun_out/*.ncu-rep
