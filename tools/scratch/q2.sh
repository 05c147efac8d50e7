run() { timeout 300 python bench.py "$@" --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.readline()
try:
    d = json.loads(l); print('$*', d['value'], d['ms_per_step'], d.get('roofline',{}).get('frac'))
except Exception as e: print('$*', 'ERR', l[:300])
"; }
for te in 32 64; do
for sc in navigation transport football joint_passage simple_spread; do
  run --scenario $sc --steps 10 --warmup 3 --envs-per-cta $te --num-envs 1048576
done; done
for sc in navigation transport football joint_passage simple_spread; do
  ncu --metrics gpu__time_duration.sum,launch__registers_per_thread,launch__occupancy_limit_shared_mem,launch__occupancy_limit_registers,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -c 12 --csv --log-file gpurun_out/ncu_$sc.csv python bench.py --scenario $sc --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > /dev/null 2>&1
done
