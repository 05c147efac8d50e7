timeout 600 python -m pytest tests/test_gpu_graph.py -x -q 2>&1 | grep -v "^    \|^E    " | tail -30
for sc in navigation transport; do
  timeout 300 python bench.py --impl dropin-graph --scenario $sc --steps 300 > gpurun_out/r1_dropin-graph_${sc}.json 2> gpurun_out/r1_dropin-graph_${sc}.err
  tail -1 gpurun_out/r1_dropin-graph_${sc}.json | cut -c1-700; tail -3 gpurun_out/r1_dropin-graph_${sc}.err
done
