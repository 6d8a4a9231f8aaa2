#!/bin/bash
# Round-1g (8), 2 GPUs: the whole GPU suite (multi-device tests included) and the N=1M line on 2 ranks, final build
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi -L | wc -l
timeout 400 python -m pytest tests -m gpu -q > $O/pytest37_2gpu.log 2>&1; echo "pytest rc=$?"
tail -n 6 $O/pytest37_2gpu.log | cut -c1-300
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29537 bench.py --gpus 2 --steps 5 --warmup 3 > $O/scale2_37.json 2> $O/scale2_37.err; echo "scale2 rc=$?"
cut -c1-300 $O/scale2_37.json
