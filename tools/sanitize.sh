#!/bin/bash
# compute-sanitizer over the shared-memory team kernels (members of an env's team hand work to each other through
# shared memory across three CTA barriers per substep): memcheck, racecheck and synccheck on the parity tests of the
# pair-heavy / jointed / mixed-shape worlds.  Run under gpurun from the repo root; logs land in gpurun_out/sanitizer/.
#   gpurun --timeout 2400 -- 'bash tools/sanitize.sh'
OUT=gpurun_out/sanitizer
mkdir -p $OUT
SEL='(football or joint_passage or pollock or navigation) and (one_step or tiles_agree)'
for tool in memcheck racecheck synccheck; do
  VMASB200_JIT=${JIT:-1} timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --log-file $OUT/$tool.log \
    python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "$SEL" > $OUT/${tool}_pytest.log 2>&1
  echo "== $tool: pytest $(tail -1 $OUT/${tool}_pytest.log)"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|error" $OUT/$tool.log | sort | uniq -c | head -10
done
