#!/bin/bash
# fp64 kernel variants: parity, then the math x work-unit sweep; outlier parity rerun.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py tests/test_gpu_addon.py -m gpu -q -x -s > gpurun_out/pytest9.log 2>&1; echo "pytest rc=$?" > gpurun_out/probe9.log
timeout 900 python scripts/f64_sweep.py > gpurun_out/f64_sweep.log 2>&1; echo "sweep rc=$?" >> gpurun_out/probe9.log
cat gpurun_out/probe9.log; grep -h "f64 math\|outliers k\|passed\|failed\|Error" gpurun_out/pytest9.log | tail -20; cat gpurun_out/f64_sweep.log | cut -c1-200
