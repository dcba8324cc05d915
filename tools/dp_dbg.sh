ACT_STEM_GROUP=4 timeout 300 python -m pytest tests/test_engine_golden.py tests/test_policy_api.py -q -m gpu --no-header -p no:cacheprovider -x 2>&1 | tail -3
for g in 4; do
ACT_STEM_GROUP=$g timeout 300 python tools/profile_infer.py 128 3 gpurun_out/tc_infer_g$g.csv 2>&1 | grep -E "step 2|tc ms"
grep -E "^0,conv" gpurun_out/tc_infer_g$g.csv
done
ACT_STEM_GROUP=4 timeout 300 python tools/profile_step.py 8 3 gpurun_out/tc_train_b8.csv 2>&1 | tail -1; grep -E "conv 8x240" gpurun_out/tc_train_b8.csv | head -2
ACT_STEM_GROUP=4 timeout 300 python bench.py --no-cpu-baseline --steps 20 --warmup 5 2>&1 | tail -1 | cut -c1-200
