timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for sc in navigation transport; do
  for impl in dropin dropin-hooks; do
    timeout 300 python bench.py --impl $impl --scenario $sc --steps 200 > gpurun_out/r1_${impl}_${sc}.json 2> gpurun_out/r1_${impl}_${sc}.err
    tail -1 gpurun_out/r1_${impl}_${sc}.json | cut -c1-700
  done
  timeout 300 python bench.py --impl dropin-hooks --static-io 1 --scenario $sc --steps 200 > gpurun_out/r1_dropin-hooks-static_${sc}.json 2>> gpurun_out/r1_dropin-hooks_${sc}.err
  tail -1 gpurun_out/r1_dropin-hooks-static_${sc}.json | cut -c1-700
done
