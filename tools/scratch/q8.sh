for sc in navigation football; do
  ncu --set full --import-source on --clock-control none -k regex:step_kernel -s 6 -c 1 -f -o gpurun_out/r1s_$sc python bench.py --scenario $sc --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > /dev/null 2>&1
done
