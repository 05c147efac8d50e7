set -x
nvidia-smi -L
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 200 --warmup 10 2>gpurun_out/r1_n2.err | tail -1 > gpurun_out/r1_n2.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 2>>gpurun_out/r1_n2.err | tail -1 > gpurun_out/r1_n2_ref.json
timeout 900 python -m pytest tests -q -m gpu 2>&1 | tail -5
