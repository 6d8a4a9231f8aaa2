#!/bin/bash
mkdir -p gpurun_out
timeout 600 ./n-body-sim_b200/force_lab 262144 3 > gpurun_out/force_lab4.log 2>&1; echo "lab rc=$?" > gpurun_out/probe4.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/probe4.log
timeout 600 python bench.py > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/probe4.log
for w in 2000 21312 60000; do NBODY_WANT_UNITS=$w timeout 300 python bench.py --no-cpu --no-e2e --steps 3 --warmup 3 >> gpurun_out/bench_units.log 2>&1; done
NBODY_R=2 timeout 300 python bench.py --no-cpu --no-e2e --steps 3 --warmup 3 >> gpurun_out/bench_units.log 2>&1
cat gpurun_out/probe4.log; tail -5 gpurun_out/pytest_gpu.log; grep PRODUCT gpurun_out/force_lab4.log; cat gpurun_out/bench.log
