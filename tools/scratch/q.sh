timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k zero_copy 2>&1 | grep -v "^    " | tail -8
for sc in navigation transport football; do for n in 32768 262144; do for d in 0 1; do
  VMASB200_HOST_DIRECT=$d timeout 300 python bench.py --scenario $sc --num-envs $n --steps 100 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print('$sc n=$n direct=$d value %.0fM e2e %.1fM' % (d['value']/1e6, d['e2e']['value']/1e6))"
done; done; done
