#!/bin/bash
# round 2, first GPU session: the pairs-once kernel's tests, the whole GPU suite, kernel-variant timings
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2a_smi.txt
timeout 600 python -m pytest tests/test_gpu_sym32.py -x -q > gpurun_out/r2a_sym_tests.txt 2>&1; echo "sym tests rc=$?"; tail -15 gpurun_out/r2a_sym_tests.txt
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/r2a_pytest_gpu.txt 2>&1; echo "all gpu tests rc=$?"; tail -8 gpurun_out/r2a_pytest_gpu.txt
timeout 600 python scripts/sym_lab.py all > gpurun_out/r2a_sym_lab.jsonl 2> gpurun_out/r2a_sym_lab.err; echo "lab rc=$?"
cut -c1-330 gpurun_out/r2a_sym_lab.jsonl; tail -5 gpurun_out/r2a_sym_lab.err
