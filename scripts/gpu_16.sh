#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_edges.py tests/test_gpu_parity.py -m gpu -q > gpurun_out/pytest16.log 2>&1; echo "pytest rc=$?"
timeout 300 python scripts/small_n_sweep.py > gpurun_out/small_n_cluster.jsonl 2>&1
tail -5 gpurun_out/pytest16.log | cut -c1-200; cat gpurun_out/small_n_cluster.jsonl
