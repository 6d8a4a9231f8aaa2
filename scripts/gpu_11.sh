#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/e2e_breakdown.py > gpurun_out/e2e_breakdown.log 2>&1
NBODY_UNIFORM=0 timeout 300 python scripts/e2e_breakdown.py >> gpurun_out/e2e_breakdown.log 2>&1
cat gpurun_out/e2e_breakdown.log
