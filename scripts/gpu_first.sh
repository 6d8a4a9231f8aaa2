#!/bin/bash
# First GPU pass: probes, parity tests, smoke, bench, pipe microbench, ncu launch list.
mkdir -p gpurun_out
{ nvidia-smi -L; nproc; (node --version || echo "node: absent"); } > gpurun_out/probe.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/probe.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/probe.log
timeout 120 ./n-body-sim_b200/pipe_bench > gpurun_out/pipe_bench.log 2>&1; echo "pipe rc=$?" >> gpurun_out/probe.log
timeout 600 python bench.py > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/probe.log
for R in 1 2 4; do NBODY_R=$R timeout 300 python bench.py --no-cpu --no-e2e --steps 3 --warmup 3 >> gpurun_out/bench_R.log 2>&1; done
timeout 300 python bench.py --no-cpu --no-e2e --steps 2 --warmup 3 --n 131072 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv python bench.py --no-cpu --no-e2e --steps 2 --warmup 3 --n 131072 > gpurun_out/ncu_launch.log 2>&1
echo "ncu rc=$?" >> gpurun_out/probe.log
cat gpurun_out/probe.log; tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log; cat gpurun_out/bench.log
