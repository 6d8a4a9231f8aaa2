#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" > gpurun_out/probe2.log
timeout 600 python bench.py > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/probe2.log
timeout 300 python bench.py --no-cpu --no-e2e --steps 2 --warmup 3 --n 131072 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:force_f32 -s 3 -c 2 -o gpurun_out/prof_r1_force_f32 -f python bench.py --no-cpu --no-e2e --steps 2 --warmup 3 --n 131072 > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?" >> gpurun_out/probe2.log
cat gpurun_out/probe2.log; tail -30 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.log
