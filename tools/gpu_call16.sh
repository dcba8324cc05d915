#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
ACT_DP_NVLS=1 timeout 120 $TR --master-port 29601 tools/dp_policy_check.py > gpurun_out/dp_policy_nvls1.log 2>&1; echo "nvls=1 unfused rc=$?"; grep dp_policy_check gpurun_out/dp_policy_nvls1.log | tail -3
ACT_DP_NVLS=1 timeout 120 $TR --master-port 29602 tools/dp_policy_check.py fused > gpurun_out/dp_policy_nvls1_fused.log 2>&1; echo "nvls=1 fused rc=$?"; grep dp_policy_check gpurun_out/dp_policy_nvls1_fused.log | tail -3
ACT_DP_NVLS=0 timeout 120 $TR --master-port 29603 tools/dp_policy_check.py fused > gpurun_out/dp_policy_nvls0_fused.log 2>&1; echo "nvls=0 fused rc=$?"; grep dp_policy_check gpurun_out/dp_policy_nvls0_fused.log | tail -3
