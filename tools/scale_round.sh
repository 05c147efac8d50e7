#!/bin/bash
# Multi-GPU lines of one box (run under `gpurun --gpus 8`): topology, the driver's navigation command at 2 / 4 / 8 GPUs (e2e
# scaling), and BASELINE config 5: football at 2 / 4 / 8 GPUs, weak (N per GPU) and strong ("N split evenly").
TAG=${1:-r2}; OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi topo -m > $OUT/topo.txt 2>&1
run() { local name=$1 g=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 200)) \
    bench.py --gpus $g "$@" 2> $OUT/bench_${name}.err | grep '^{' | tail -1 > $OUT/bench_${name}.json
  echo "$name: $(cut -c1-150 $OUT/bench_${name}.json)"; }
GS=${GS:-2 4 8}   # GPU counts to run (each must fit the box: `gpurun --gpus N`)
for g in $GS; do run navigation_n$g $g --steps 100 --warmup 5 --no-cpu-baseline --no-torch-cuda --no-1m; done
for g in $GS; do
  run football_weak_262144_n$g $g --scenario football --num-envs 262144 --steps 50 --no-e2e --no-cpu-baseline --no-torch-cuda --no-1m
  run football_strong_1048576_n$g $g --scenario football --num-envs 1048576 --scaling strong --steps 50 --no-e2e --no-cpu-baseline --no-torch-cuda --no-1m
done
if [ -z "$SHORT" ]; then
run football_weak_1048576_n8 8 --scenario football --num-envs 1048576 --steps 30 --no-e2e --no-cpu-baseline --no-torch-cuda --no-1m
run football_strong_65536_n8 8 --scenario football --num-envs 65536 --scaling strong --steps 100 --no-e2e --no-cpu-baseline --no-torch-cuda --no-1m
fi
