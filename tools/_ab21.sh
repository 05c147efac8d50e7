B="--steps 200 --warmup 10 --no-cpu-baseline --no-1m --no-torch-cuda"
run() { python bench.py --scenario $1 $B $2 2>/dev/null | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); e=d["e2e"]; print("e2e %.1f M  %.3f ms (floor %.3f)" % (e["value"]/1e6, e["ms_per_step"], e["pcie_floor_ms"]))'; }
for sc in navigation transport simple_spread football; do
echo "[$sc copy] $(run $sc)"
echo "[$sc direct] $(VMASB200_HOST_ACTIONS_DIRECT=1 run $sc)"
echo "[$sc direct 1chunk] $(VMASB200_HOST_ACTIONS_DIRECT=1 VMASB200_HOST_CHUNK_ENVS=131072 run $sc)"
done
