#!/bin/bash
# compute-sanitizer over the small parity cases (run under gpurun): memcheck on everything quick, racecheck and
# synccheck on the team kernel's shared-memory phases.  Reports land in gpurun_out/sanitize_*.log.
K='test_one_step or test_repair_path or test_tiles_agree'
for tool in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 \
    python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "$K" > gpurun_out/sanitize_$tool.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/sanitize_$tool.log | tail -3
done
