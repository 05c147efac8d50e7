set -x
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -8
for te in 32 64 128; do
for sc in navigation transport football joint_passage simple_spread; do
  timeout 300 python bench.py --scenario $sc --steps 30 --warmup 5 --envs-per-cta $te --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.readline()
try:
    d = json.loads(l); print('$sc te=$te', d['value'], d['ms_per_step'], d.get('roofline',{}).get('frac'), d.get('speculation_repairs'))
except Exception as e: print('$sc', 'ERR', l[:300])
"
done; done
