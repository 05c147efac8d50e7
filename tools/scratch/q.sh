timeout 900 python -m pytest tests/test_gpu_hooks.py tests/test_gpu_graph.py -x -q 2>&1 | grep -v "^    \|^E    " | tail -12
for impl in reference-cuda dropin dropin-hooks dropin-graph; do
  timeout 300 python bench.py --impl $impl --scenario simple_spread --num-envs 32768 --steps 200 > gpurun_out/r1_${impl}_simple_spread.json 2> gpurun_out/r1_${impl}_simple_spread.err
  tail -1 gpurun_out/r1_${impl}_simple_spread.json | cut -c1-600; tail -2 gpurun_out/r1_${impl}_simple_spread.err
done
