#!/bin/bash
# Round artifacts on one B200 (run under gpurun from the repo root): bench lines of every BASELINE.json config, the
# reference arms, the ncu launch list and one `--set full` capture of the step kernel per workload, reduced ON THE BOX to
# text (tools/profile_lines.sh: the .ncu-rep files would exceed gpurun_out's 64 MiB).  Everything lands in
# gpurun_out/<tag>/; tools/profile_collect.py (run afterwards, no GPU needed) turns it into profiles/.
#   gpurun --timeout 2400 -- 'bash tools/profile_round.sh r1w'
TAG=${1:-r2}
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
for sc in transport football joint_passage simple_spread transport_s5; do bench $sc --scenario $sc; done
# (KERNEL_ONLY=1: only what depends on the step kernel - skips the generic-kernel runs and the reference arms)
if [ -z "$KERNEL_ONLY" ]; then
# the same launches through the generic table-walking kernel (world-compiled kernels off)
for sc in navigation transport simple_spread; do VMASB200_JIT=0 bench generic_$sc --scenario $sc --no-cpu-baseline --no-e2e; done
fi
# 2. the large-batch points the roofline fraction is quoted on
for sc in navigation transport football simple_spread transport_s5 joint_passage; do bench ${sc}_1m --scenario $sc --num-envs 1048576 --steps 50 --no-cpu-baseline --no-e2e --no-1m --no-torch-cuda; done
if [ -z "$KERNEL_ONLY" ]; then
for sc in navigation transport simple_spread; do VMASB200_JIT=0 bench generic_${sc}_1m --scenario $sc --num-envs 1048576 --steps 50 --no-cpu-baseline --no-e2e; done
fi
for n in 4096 16384 65536 262144; do bench football_$n --scenario football --num-envs $n --steps 100 --no-cpu-baseline --no-e2e; done
# 3. reference arms: the reference's torch CPU step, its torch CUDA step, and the drop-in inside its Environment
if [ -z "$KERNEL_ONLY" ]; then
bench reference_navigation --impl reference --steps 3 --warmup 1
bench reset_navigation --impl reset
bench reset_transport --impl reset --scenario transport
for sc in navigation transport simple_spread joint_passage football; do
  bench reference-cuda_$sc --impl reference-cuda --scenario $sc --steps 30
  bench dropin-hooks_$sc --impl dropin-hooks --scenario $sc --steps 50
done
for sc in navigation transport simple_spread joint_passage; do bench dropin-graph_$sc --impl dropin-graph --scenario $sc --steps 50; done
fi
if [ -n "$SKIP_NCU" ]; then ls $OUT | wc -l; exit 0; fi
# 4. ncu: launch list of the driver's command, then one full capture per workload (only after the plain runs above)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_navigation.csv \
  $PY bench.py --steps 20 --warmup 3 --no-cpu-baseline > $OUT/launches_navigation.log 2>&1
bash tools/profile_lines.sh $TAG navigation transport simple_spread football joint_passage transport_s5 \
  navigation@1048576 transport@1048576 simple_spread@1048576 football@1048576
ls -la $OUT
