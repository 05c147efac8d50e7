#!/bin/bash
# generic team kernel (VMASB200_JIT=0) against the default selection (world-compiled where jit_eligible says so) per world
mkdir -p gpurun_out
for sc in ${WORLDS:-waterfall passage balance give_way ball_passage asym_joint reverse_transport wheel football joint_passage transport}; do
  for n in ${SIZES:-32768 262144}; do
    for mode in 0 1; do
      r=$(VMASB200_JIT=$mode python bench.py --scenario $sc --num-envs $n --steps 60 --warmup 5 --no-cpu-baseline --no-e2e --no-1m --no-torch-cuda --fixture-inputs 2>/dev/null | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("%.2f us  %.3f G  %s te=%s rep=%s" % (d["ms_per_step"]*1000, d["value"]/1e9, d["roofline"]["kernel"][:19], d["config"]["envs_per_cta"], d["config"]["speculation_repairs"]))')
      echo "[$sc@$n jit=$mode] $r" | tee -a gpurun_out/ab_worlds.txt
    done
  done
done
