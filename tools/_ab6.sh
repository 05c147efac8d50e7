B="--steps 60 --warmup 5 --no-cpu-baseline --no-e2e --no-1m --no-torch-cuda"
for sc in football joint_passage; do for n in 32768 1048576; do for te in 0 32 64; do
  if [ $te = 0 ]; then unset VMASB200_JIT_TE; else export VMASB200_JIT_TE=$te; fi
  if [ $n = 32768 ] && [ $te != 0 ]; then continue; fi
  r=$(python bench.py --scenario $sc --num-envs $n $B 2>/dev/null | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("%.2f us  %.3f G" % (d["ms_per_step"]*1000, d["value"]/1e9))')
  echo "[$sc@$n te=$te] $r"
done; done; done
