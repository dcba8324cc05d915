#!/bin/bash
# ncu captures of the bench workload's training step (run on the GPU box; outputs under gpurun_out/).
#   1. section-level metrics for every tcgen05 launch of the 2nd step   -> gpurun_out/ncu_tc_sections.ncu-rep
#   2. --set full + source for selected launches (indices into the per-step tc launch list)
set -u
mkdir -p gpurun_out
TC_PER_STEP=${TC_PER_STEP:-277}
python tools/profile_step.py 8 2 > gpurun_out/profile_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/profile_plain.log; exit 1; }
ncu -k regex:tc_kernel -s $TC_PER_STEP -c $TC_PER_STEP --clock-control none \
    --section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section LaunchStats --section Occupancy --section SchedulerStats \
    -f -o gpurun_out/ncu_tc_sections python tools/profile_step.py 8 2 > gpurun_out/ncu_sections.log 2>&1
tail -1 gpurun_out/ncu_sections.log
for idx in "$@"; do
  ncu -k regex:tc_kernel -s $((TC_PER_STEP + idx)) -c 1 --set full --import-source on --clock-control none \
      -f -o gpurun_out/ncu_full_tc${idx} python tools/profile_step.py 8 2 > gpurun_out/ncu_full_${idx}.log 2>&1
  tail -1 gpurun_out/ncu_full_${idx}.log
done
ls -la gpurun_out/*.ncu-rep
