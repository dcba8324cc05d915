#!/bin/bash
mkdir -p gpurun_out
tools/gpu_tests.sh tests/test_engine_golden.py tests/test_dropout.py tests/test_policy_api.py
timeout 600 python bench.py --cameras 3 --batch 16 --steps 30 --no-infer --no-policy-api --no-cpu-baseline > gpurun_out/bench_n1_k3_b16.json 2> gpurun_out/bench_n1_k3_b16.err; echo "k3 rc=$? $(cut -c1-160 gpurun_out/bench_n1_k3_b16.json | grep '^{')"
ACT_ATTN_CHUNKED=1 timeout 600 python bench.py --cameras 3 --batch 16 --steps 30 --no-infer --no-policy-api --no-cpu-baseline > gpurun_out/bench_n1_k3_b16_chunked.json 2> gpurun_out/bench_n1_k3_b16_chunked.err; echo "k3 chunked rc=$? $(cut -c1-160 gpurun_out/bench_n1_k3_b16_chunked.json | grep '^{')"
timeout 600 python bench.py --cameras 3 --batch 16 --steps 30 --dropout 0.1 --no-infer --no-policy-api --no-cpu-baseline > gpurun_out/bench_n1_k3_b16_drop.json 2> gpurun_out/bench_n1_k3_b16_drop.err; echo "k3 dropout rc=$? $(cut -c1-160 gpurun_out/bench_n1_k3_b16_drop.json | grep '^{')"
