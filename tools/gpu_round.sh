#!/bin/bash
# One GPU-box visit: full GPU test suite, smoke, the default bench line, the configs[2]/[3]-shaped lines, fresh ncu launch
# lists (training step at B=8, inference step at E=128).  Outputs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > gpurun_out/gpu.txt 2>&1
tools/gpu_tests.sh
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$? $(tail -1 gpurun_out/smoke.log)"
timeout 600 python bench.py --eager-gpu > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_n1.json
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --batch 64 --no-infer --no-cpu-baseline --eager-gpu > gpurun_out/bench_n1_b64.json 2> gpurun_out/bench_n1_b64.err; echo "b64 rc=$?"; cut -c1-300 gpurun_out/bench_n1_b64.json
timeout 600 python bench.py --cameras 3 --batch 16 --no-cpu-baseline --infer-envs 32 > gpurun_out/bench_n1_k3_b16.json 2> gpurun_out/bench_n1_k3_b16.err; echo "k3 rc=$?"; cut -c1-300 gpurun_out/bench_n1_k3_b16.json
timeout 300 python tools/profile_step.py 8 3 gpurun_out/tc_train_b8.csv > gpurun_out/profile_train.log 2>&1; tail -2 gpurun_out/profile_train.log
timeout 300 python tools/profile_infer.py 128 3 gpurun_out/tc_infer_e128.csv > gpurun_out/profile_infer.log 2>&1; tail -3 gpurun_out/profile_infer.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/ncu_train_b8.csv python tools/profile_step.py 8 2 > gpurun_out/ncu_train.log 2>&1; echo "ncu train rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/ncu_infer_e128.csv python tools/profile_infer.py 128 2 > gpurun_out/ncu_infer.log 2>&1; echo "ncu infer rc=$?"
