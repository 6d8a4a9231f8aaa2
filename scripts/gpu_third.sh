#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_addon.py -m gpu -q > gpurun_out/pytest_addon.log 2>&1; echo "addon rc=$?" > gpurun_out/probe3.log
timeout 900 ./n-body-sim_b200/force_lab 262144 3 > gpurun_out/force_lab.log 2>&1; echo "lab rc=$?" >> gpurun_out/probe3.log
timeout 300 python bench.py --no-cpu --no-e2e --steps 1 --warmup 3 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:force_f32 -s 3 -c 1 -o gpurun_out/prof_r1_force_f32_n1m -f python bench.py --no-cpu --no-e2e --steps 1 --warmup 3 > gpurun_out/ncu_full_1m.log 2>&1
echo "ncu rc=$?" >> gpurun_out/probe3.log
cat gpurun_out/probe3.log; tail -5 gpurun_out/pytest_addon.log; cat gpurun_out/force_lab.log
