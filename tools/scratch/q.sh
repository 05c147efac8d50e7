set -x
timeout 600 python tools/scratch/dbg_static.py joint_passage 1 2>&1 | tail -30
timeout 600 python tools/scratch/dbg_static.py joint_passage 0 2>&1 | tail -30
timeout 900 python -m pytest tests -q -m gpu 2>&1 | tail -15
