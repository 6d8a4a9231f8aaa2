#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_contacts.py tests/test_gpu_fullsize.py tests/test_gpu_addon.py -m gpu -q -x > gpurun_out/pytest19.log 2>&1; echo "pytest rc=$?"
timeout 600 python scripts/contacts_bench.py --big --out gpurun_out/contacts_compact.jsonl > gpurun_out/contacts19.log 2>&1; echo "bench rc=$?"
tail -5 gpurun_out/pytest19.log | cut -c1-200; cut -c1-600 gpurun_out/contacts_compact.jsonl
