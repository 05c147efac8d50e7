run() { timeout 600 python bench.py "$@" --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys, json, os
l = sys.stdin.readline()
try:
    d = json.loads(l); e = d['e2e']; print(os.environ.get('VMASB200_HOST_CHUNK_ENVS'), '$*', 'e2e %.4g' % e['value'])
except Exception as e: print('$*', 'ERR', l[:300])
"; }
for ce in 8192 16384 32768 65536 1000000; do
export VMASB200_HOST_CHUNK_ENVS=$ce
for sc in navigation simple_spread football; do
  run --scenario $sc --steps 100
done
run --scenario navigation --steps 30 --num-envs 262144
done
