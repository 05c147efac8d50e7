#!/bin/bash
# One `ncu --set full` capture of the step kernel per workload, reduced ON THE BOX to the text summaries kept under
# profiles/ (the .ncu-rep files are ~13 MB each and gpurun_out/ is capped at 64 MiB).
#   gpurun --timeout 900 -- 'bash tools/profile_lines.sh r1h navigation navigation@1048576 transport@1048576'
TAG=$1; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
SO=vectorizedmultiagentsimulator_b200/libvmasb200.so
for wl in "$@"; do
  sc=${wl%@*}; n=32768; name=$sc
  if [[ $wl == *@* ]]; then n=${wl#*@}; name=${sc}_$n; fi
  rep=/tmp/step_kernel_$name
  export VMASB200_JIT_CACHE=/tmp/jit_$name   # this run's world-compiled cubin alone (ncu_lines needs its line table)
  rm -rf $VMASB200_JIT_CACHE
  ncu --set full --metrics smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum --import-source on --clock-control none -k 'regex:step_kernel|vmasb200_step_world' -s 6 -c 1 -f -o $rep \
    python bench.py --scenario $sc --num-envs $n --steps 4 --warmup 3 --no-cpu-baseline --no-e2e --no-1m --no-torch-cuda > $OUT/ncu_$name.log 2>&1
  python tools/ncu_summary.py $rep.ncu-rep > $OUT/step_kernel_$name.txt 2>&1
  kname=$(head -1 $OUT/step_kernel_$name.txt)
  if echo "$kname" | grep -q vmasb200_step_world; then
    python tools/ncu_lines.py $rep.ncu-rep $(ls $VMASB200_JIT_CACHE/*.cubin | head -1) vmasb200_step_world > $OUT/${name}_lines.txt 2>&1
  else
    te=$(echo "$kname" | sed -n 's/.*step_kernel<[0-9], \([0-9]*\)>.*/\1/p')
    python tools/ncu_lines.py $rep.ncu-rep $SO step_kernelILb1ELi$te > $OUT/${name}_lines.txt 2>&1
  fi
  ncu -i $rep.ncu-rep --page source --csv 2>/dev/null | gzip > $OUT/${name}_source.csv.gz  # per-instruction stall reasons
  rm -f $rep.ncu-rep
done
ls -la $OUT
