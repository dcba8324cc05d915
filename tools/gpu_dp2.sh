#!/bin/bash
# N-GPU visit with the unit tests of the exchange kernel first
NP=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dp_p2p.py -q -m gpu --no-header -p no:cacheprovider > gpurun_out/tests_dp_p2p_n$NP.log 2>&1; echo "test_dp_p2p rc=$? $(tail -1 gpurun_out/tests_dp_p2p_n$NP.log)"
tools/gpu_dp.sh "$@"
