#!/bin/bash
NP=${1:-2}; shift
mkdir -p gpurun_out
tools/gpu_tests.sh tests/test_fused_kernels.py tests/test_engine_golden.py tests/test_dp_p2p.py
tools/gpu_dp_ab2.sh $NP "$@"
