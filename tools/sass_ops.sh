#!/bin/bash
# SASS evidence (no GPU needed): mnemonic counts of the library's kernels and of the world-compiled cubins of the BASELINE
# worlds - UBLKCP (TMA bulk copies), SYNCS (mbarriers), REDUX (warp reductions), FFMA2 / FADD2 / FMUL2 (packed fp32),
# ACQBULK / griddepcontrol (PDL), LDS / STS, local-memory traffic.   bash tools/sass_ops.sh > profiles/r2_sass_ops.txt
SO=vectorizedmultiagentsimulator_b200/libvmasb200.so
OPS="UBLKCP SYNCS REDUX FFMA2 FADD2 FMUL2 FFMA FADD FMUL MUFU LDS STS LDG STG LDL STL ATOM RED BAR ACQBULK PREEXIT WARPSYNC SHFL"
count() {  # stdin: sass of one function
  awk -v ops="$OPS" 'BEGIN{n=split(ops,o," ")} /^ +\/\*[0-9a-f]+\*\/ / {t++; for(i=1;i<=n;i++){ if ($0 ~ ("[ .@!P0-9]" o[i] "[ .;]")) c[o[i]]++ }} END{printf "  instructions %d:", t; for(i=1;i<=n;i++) if (c[o[i]]) printf " %s %d", o[i], c[o[i]]; printf "\n"}'
}
echo "# $SO ($(date -u +%F)), cuobjdump -sass, per kernel"
cuobjdump -sass $SO | awk '/Function :/ {name=$3} {print name "\t" $0}' > /tmp/sass_all.txt
for k in $(cut -f1 /tmp/sass_all.txt | sort -u | grep -v '^$'); do
  echo "$(echo $k | c++filt | cut -c1-110)"; grep -P "^\Q$k\E\t" /tmp/sass_all.txt | cut -f2- | count
done
CACHE=${1:-/tmp/sass_jit}; rm -rf $CACHE; mkdir -p $CACHE
for w in navigation@32768 navigation@1048576 transport@32768 simple_spread@1048576 football@1048576 joint_passage@32768; do
  sc=${w%@*}; n=${w#*@}; d=$CACHE/$sc-$n; mkdir -p $d
  VMASB200_JIT_CACHE=$d python - "$sc" "$n" <<'PY'
import sys
sys.path.insert(0, '.')
from vectorizedmultiagentsimulator_b200.plan import StepPlan
from vectorizedmultiagentsimulator_b200.workload import Workload
StepPlan(Workload(sys.argv[1], 64).desc).jit(int(sys.argv[2]))
PY
  for f in $d/*.cubin; do
    echo "vmasb200_step_world  $sc, $n envs  ($(cuobjdump --dump-resource-usage $f | grep -A1 step_world | grep -o 'REG:[0-9]* STACK:[0-9]* SHARED:[0-9]*'))"
    cuobjdump -sass $f | awk '/Function : vmasb200_step_world/ {on=1} on' | count
  done
done
