#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest18.log 2>&1; echo "pytest rc=$?"
timeout 300 python scripts/f64_sweep.py --math 2 --units 0 --sizes 4096,16384,32768,65536 --tag "auto R, equal-mass" --out gpurun_out/f64_final.jsonl > gpurun_out/f64_final.log 2>&1
tail -6 gpurun_out/pytest18.log | cut -c1-220; cut -c1-240 gpurun_out/f64_final.jsonl
