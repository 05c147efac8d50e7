run() { timeout 300 python bench.py "$@" --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.readline()
try:
    d = json.loads(l); print('$*', '%.3g' % d['value'], '%.4f' % d['ms_per_step'], '%.4f' % d.get('roofline',{}).get('frac'))
except Exception as e: print('$*', 'ERR', l[:300])
"; }
for sc in navigation football joint_passage; do
  run --scenario $sc --steps 10 --warmup 3 --num-envs 1048576 --envs-per-cta 128
done
for sc in navigation transport football simple_spread; do
for te in 32 64 128; do
  run --scenario $sc --steps 10 --warmup 3 --num-envs 262144 --envs-per-cta $te
done; done
python - <<'PY'
import sys
sys.path.insert(0, '.')
from vectorizedmultiagentsimulator_b200 import workload
from vectorizedmultiagentsimulator_b200.plan import Plan
PY
