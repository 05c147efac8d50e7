#!/bin/bash
# BASELINE configs 3 and 5 on one B200: transport with substeps = 5 at 32768 envs (and 1 M), football sweep 4 K .. 1 M.
#   gpurun --timeout 1500 -- 'bash tools/sweep_c3_c5.sh r2'
TAG=${1:-r2}; OUT=gpurun_out/$TAG; mkdir -p $OUT
b() { local name=$1; shift; timeout 600 python bench.py "$@" 2> $OUT/bench_$name.err | tail -1 > $OUT/bench_$name.json; echo "$name: $(cut -c1-160 $OUT/bench_$name.json)"; }
b transport_s5 --scenario transport_s5 --steps 100
b transport_s5_1m --scenario transport_s5 --num-envs 1048576 --steps 50 --no-cpu-baseline --no-e2e --no-torch-cuda
for n in 4096 16384 65536 262144 1048576; do b football_$n --scenario football --num-envs $n --steps 100 --no-cpu-baseline --no-e2e --no-torch-cuda --no-1m; done
