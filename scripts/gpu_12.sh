#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_contacts.py tests/test_gpu_edges.py -m gpu -q -x > gpurun_out/pytest12.log 2>&1; echo "pytest rc=$?"
timeout 300 python scripts/e2e_breakdown.py > gpurun_out/e2e_breakdown.log 2>&1
timeout 300 python scripts/e2e_breakdown.py --n 131072 >> gpurun_out/e2e_breakdown.log 2>&1
tail -2 gpurun_out/pytest12.log; cat gpurun_out/e2e_breakdown.log
