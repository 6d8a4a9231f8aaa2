#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" > gpurun_out/probe6.log
timeout 600 python bench.py > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/probe6.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>gpurun_out/bench_ref.err; echo "ref rc=$?" >> gpurun_out/probe6.log
timeout 600 python scripts/size_sweep.py > gpurun_out/size_sweep.jsonl 2> gpurun_out/size_sweep.err; echo "sweep rc=$?" >> gpurun_out/probe6.log
cat gpurun_out/probe6.log; tail -3 gpurun_out/pytest_gpu.log; cut -c1-330 gpurun_out/bench.log; cut -c1-300 gpurun_out/bench_ref.log; cat gpurun_out/size_sweep.jsonl
