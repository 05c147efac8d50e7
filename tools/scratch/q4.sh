run() { timeout 300 python bench.py "$@" --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.readline()
try:
    d = json.loads(l); print('$*', '%.3g' % d['value'], '%.4f' % d['ms_per_step'], '%.4f' % d.get('roofline',{}).get('frac'))
except Exception as e: print('$*', 'ERR', l[:300])
"; }
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
sweep() {
for sc in navigation transport football joint_passage simple_spread; do
  run --scenario $sc --steps 30 --warmup 5
  run --scenario $sc --steps 10 --warmup 3 --num-envs 1048576
done
}
echo "== noinline"; sweep
run --scenario transport --steps 10 --warmup 3 --num-envs 1048576 --envs-per-cta 128
run --scenario simple_spread --steps 10 --warmup 3 --num-envs 1048576 --envs-per-cta 128
run --scenario navigation --steps 10 --warmup 3 --num-envs 131072 --envs-per-cta 32
run --scenario navigation --steps 10 --warmup 3 --num-envs 131072 --envs-per-cta 64
run --scenario football --steps 10 --warmup 3 --num-envs 65536 --envs-per-cta 32
run --scenario football --steps 10 --warmup 3 --num-envs 65536 --envs-per-cta 64
cp tools/scratch/lib_inline.so vectorizedmultiagentsimulator_b200/libvmasb200.so
echo "== inline"; sweep
