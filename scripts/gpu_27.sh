#!/bin/bash
# where the one-cluster kernel hands over to the all-SM one-launch kernel
mkdir -p gpurun_out
O=gpurun_out
( for m in 512 0; do echo "# NBODY_SMALL_MAX=$m"; NBODY_SMALL_MAX=$m timeout 300 python scripts/small_n_sweep.py; done ) > $O/small_n27.jsonl 2> $O/small_n27.err
python - <<'PY'
import json
for l in open('gpurun_out/small_n27.jsonl'):
    if l.startswith('#'): print(l.strip()); continue
    r = json.loads(l); print(r['precision'], r['n'], f"{r['us_per_step_k2000']:.2f} us/step (k=2000)  {r['us_per_step_k2_with_sync']:.2f} (k=2 + sync)")
PY
timeout 600 python -m pytest tests/test_gpu_edges.py -m gpu -q -x -k "medium or small" > $O/pytest27.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest27.log
