#!/bin/bash
# Outlier-corrected equal-mass sweep: parity tests, then config 3's force pass with and without it.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_contacts.py -m gpu -q -x > gpurun_out/pytest8.log 2>&1; echo "pytest rc=$?" > gpurun_out/probe8.log
CONTACTS_QUICK=1 timeout 300 python scripts/contacts_bench.py --reps 5 --out gpurun_out/cfg3_outlier.json > gpurun_out/cfg3_a.log 2>&1; echo "a rc=$?" >> gpurun_out/probe8.log
NBODY_UNIFORM=0 CONTACTS_QUICK=1 timeout 300 python scripts/contacts_bench.py --reps 5 --out gpurun_out/cfg3_general.json > gpurun_out/cfg3_b.log 2>&1; echo "b rc=$?" >> gpurun_out/probe8.log
cat gpurun_out/probe8.log; tail -5 gpurun_out/pytest8.log; cat gpurun_out/cfg3_outlier.json gpurun_out/cfg3_general.json
