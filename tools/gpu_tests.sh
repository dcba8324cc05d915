#!/bin/bash
# Run every GPU test file in its own process (a CUDA fault in one file must not poison the others).
# Usage: tools/gpu_tests.sh [files...]   -> logs in gpurun_out/tests_*.log, summary in gpurun_out/tests_summary.txt
mkdir -p gpurun_out
files=("$@")
if [ ${#files[@]} -eq 0 ]; then files=(tests/test_*.py); fi
: > gpurun_out/tests_summary.txt
for f in "${files[@]}"; do
  name=$(basename "$f" .py)
  timeout 600 python -m pytest "$f" -q -m gpu --no-header -p no:cacheprovider > "gpurun_out/tests_${name}.log" 2>&1
  rc=$?
  echo "$name rc=$rc $(tail -1 gpurun_out/tests_${name}.log)" | tee -a gpurun_out/tests_summary.txt
done
