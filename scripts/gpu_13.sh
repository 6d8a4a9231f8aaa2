#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest13.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest13.log
