#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest15.log 2>&1; echo "pytest rc=$?"
timeout 300 python scripts/small_n_sweep.py > gpurun_out/small_n_cluster.jsonl 2>&1
tail -15 gpurun_out/pytest15.log | cut -c1-200; cat gpurun_out/small_n_cluster.jsonl
