#!/bin/bash
# A/B of world-compiled kernel variants on one box: every variant is a set of -D options for NVRTC (VMASB200_JIT_DEFINES,
# csrc/jit.h), each measured with the same bench command.
#   gpurun -- 'bash tools/ab_jit.sh "navigation@32768 navigation@1048576" "" "-DVB_AB_X" "-DVB_AB_X -DVB_AB_Y" "NT=64"'
WLS=$1; shift
mkdir -p gpurun_out
i=0
for v in "$@"; do
  i=$((i+1))
  # a token NT=<threads> sets VMASB200_JIT_NT (threads per CTA of a LIDAR world), the rest are -D options
  nt=$(echo "$v" | tr ' ' '\n' | sed -n 's/^NT=//p'); defs=$(echo "$v" | tr ' ' '\n' | grep -v '^NT=' | tr '\n' ' ')
  if [ -n "$nt" ]; then export VMASB200_JIT_NT=$nt; else unset VMASB200_JIT_NT; fi
  for wl in $WLS; do
    sc=${wl%@*}; n=${wl#*@}
    for rep in 1 2; do
      r=$(VMASB200_JIT_CACHE=/tmp/ab_$i VMASB200_JIT_DEFINES="$defs" python bench.py --scenario $sc --num-envs $n --steps 200 --warmup 5 --no-cpu-baseline --no-e2e 2>/dev/null | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("%.2f us  %.3f G" % (d["ms_per_step"]*1000, d["value"]/1e9))')
      echo "[$v] $wl run$rep: $r" | tee -a gpurun_out/ab_jit.txt
    done
  done
done
