set -x
timeout 900 python -m pytest tests/test_gpu_hooks.py -x -q 2>&1 | tail -25

timeout 300 python tools/scratch/phases.py hooks 2>&1 | tail -4
