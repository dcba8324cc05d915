timeout 300 python -m pytest tests/test_engine_golden.py tests/test_policy_api.py tests/test_conv.py -q -m gpu --no-header -p no:cacheprovider -x 2>&1 | tail -5
timeout 300 python tools/profile_infer.py 128 3 gpurun_out/tc_infer_e128.csv 2>&1 | tail -4
grep -E "conv3 128x120" gpurun_out/tc_infer_e128.csv | head -5
timeout 300 python bench.py --no-cpu-baseline --steps 20 --warmup 5 2>&1 | tail -1 | cut -c1-200
ACT_CONV_PAIR=0 timeout 300 python bench.py --no-cpu-baseline --no-infer --steps 20 --warmup 5 2>&1 | tail -1 | cut -c1-200
