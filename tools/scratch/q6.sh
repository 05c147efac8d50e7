run() { timeout 600 python bench.py "$@" --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.readline()
try:
    d = json.loads(l); c = d['config']; print('$*', '%.3g' % d['value'], '%.4f' % d['ms_per_step'], 'hot %.4f' % c['l2_resident_ms_per_step'], 'repairs', c['speculation_repairs'], 'te', c['envs_per_cta'])
except Exception as e: print('$*', 'ERR', l[:300])
"; }
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
for sc in joint_passage football navigation transport; do
  run --scenario $sc
done
run --scenario joint_passage --num-envs 1048576 --steps 20
run --scenario football --num-envs 1048576 --steps 20
run --scenario football --num-envs 262144 --steps 100
