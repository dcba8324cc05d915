#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fused_kernels.py tests/test_dropout.py -q -m gpu --no-header -p no:cacheprovider -k "attn or dropout or scores" > gpurun_out/tests_attn.log 2>&1; echo "attn tests rc=$? $(tail -1 gpurun_out/tests_attn.log)"
timeout 300 python -m pytest tests/test_engine_golden.py -q -m gpu --no-header -p no:cacheprovider -k "chunked or full_k3" > gpurun_out/tests_k3.log 2>&1; echo "k3 golden rc=$? $(tail -1 gpurun_out/tests_k3.log)"
timeout 300 python tools/bench_kernels.py attn > gpurun_out/attn_bench2.txt 2>&1; tail -3 gpurun_out/attn_bench2.txt
