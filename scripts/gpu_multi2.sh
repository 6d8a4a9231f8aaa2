#!/bin/bash
# 8-GPU pass: multi-GPU tests, config 5 (cold collapse N=4,194,304) and the N=1M line, equal-mass kernel.
mkdir -p gpurun_out
nvidia-smi -L | wc -l > gpurun_out/multi2_probe.log
timeout 600 python -m pytest tests -m gpu -q -x -k "multi" > gpurun_out/pytest_multi2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/multi2_probe.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29512 bench.py --gpus 8 --scene cold_collapse --bodies 4194304 --steps 4 --warmup 3 > gpurun_out/config5_8gpu.log 2> gpurun_out/config5_8gpu.err; echo "config5 rc=$?" >> gpurun_out/multi2_probe.log
timeout 600 $TR --master-port 29513 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/scale8.log 2> gpurun_out/scale8.err; echo "scale8 rc=$?" >> gpurun_out/multi2_probe.log
cat gpurun_out/multi2_probe.log; tail -3 gpurun_out/pytest_multi2.log; grep -h '^{' gpurun_out/config5_8gpu.log gpurun_out/scale8.log | cut -c1-330; tail -3 gpurun_out/config5_8gpu.err
