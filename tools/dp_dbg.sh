ncu -k regex:tc_kernel -s 0 -c 1 --set full --import-source on --clock-control none -f -o gpurun_out/ncu_full_stem python tools/profile_infer.py 128 1 > gpurun_out/ncu_full_stem.log 2>&1; tail -2 gpurun_out/ncu_full_stem.log
ncu -k regex:conv3_kernel -s 0 -c 1 --set full --import-source on --clock-control none -f -o gpurun_out/ncu_full_conv3pair python tools/profile_infer.py 128 1 > gpurun_out/ncu_full_conv3pair.log 2>&1; tail -2 gpurun_out/ncu_full_conv3pair.log
ls -la gpurun_out/*.ncu-rep
