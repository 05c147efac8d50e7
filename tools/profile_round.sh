#!/bin/bash
# Round artifacts on one B200 (run under gpurun from the repo root): bench lines of every BASELINE.json config, the
# reference arms, the ncu launch list and one `--set full` capture of the step kernel per workload.  Everything lands
# in gpurun_out/; tools/profile_collect.py (run afterwards, no GPU needed) turns it into profiles/.
#   gpurun --timeout 2400 -- 'bash tools/profile_round.sh r1'
TAG=${1:-r1}
OUT=gpurun_out/$TAG
mkdir -p $OUT
PY=python

bench() {  # name, args...
  local name=$1; shift
  timeout 900 $PY bench.py "$@" 2> $OUT/bench_$name.err | tail -1 > $OUT/bench_$name.json
  echo "bench_$name: $(cut -c1-200 $OUT/bench_$name.json)"
}

# 1. the driver's command (navigation, 32768 envs, e2e + CPU baseline) and the other configs at the same size
bench navigation
for sc in transport football joint_passage simple_spread; do bench $sc --scenario $sc; done
# 2. the large-batch points the roofline fraction is quoted on
for sc in navigation transport football; do bench ${sc}_1m --scenario $sc --num-envs 1048576 --steps 50 --no-cpu-baseline --no-e2e; done
for n in 4096 16384 65536 262144; do bench football_$n --scenario football --num-envs $n --steps 100 --no-cpu-baseline --no-e2e; done
# 3. reference arms: the reference's torch CPU step, its torch CUDA step, and the drop-in inside its Environment
bench queries --impl queries --num-envs 1048576 --steps 100
bench reference_navigation --impl reference --steps 3 --warmup 1
for sc in navigation transport; do
  bench reference-cuda_$sc --impl reference-cuda --scenario $sc --steps 50
  bench dropin_$sc --impl dropin --scenario $sc --steps 50
done
# 4. ncu: launch list of the driver's command, then one full capture per workload (only after the plain runs above)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_navigation.csv \
  $PY bench.py --steps 20 --warmup 3 --no-cpu-baseline > $OUT/launches_navigation.log 2>&1
for sc in navigation transport football joint_passage; do
  ncu --set full --import-source on --clock-control none -k regex:step_kernel -s 6 -c 1 -f -o $OUT/step_kernel_$sc \
    $PY bench.py --scenario $sc --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/ncu_$sc.log 2>&1
done
ncu --set full --import-source on --clock-control none -k regex:step_kernel -s 6 -c 1 -f -o $OUT/step_kernel_navigation_1m \
  $PY bench.py --scenario navigation --num-envs 1048576 --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/ncu_navigation_1m.log 2>&1
ls -la $OUT
