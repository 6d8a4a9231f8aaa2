#!/bin/bash
# one-launch all-SM kernel for 513..4096 bodies (medium.cuh): parity, then latency against the launch-per-step path
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_edges.py -m gpu -q -x -k "medium or small" > $O/pytest26a.log 2>&1; echo "pytest medium rc=$?"
tail -6 $O/pytest26a.log | cut -c1-300
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest26.log 2>&1; echo "pytest rc=$?"
tail -6 $O/pytest26.log | cut -c1-300
( echo "# NBODY_MEDIUM=1 (default)"; timeout 300 python scripts/small_n_sweep.py; echo "# NBODY_MEDIUM=0"; NBODY_MEDIUM=0 timeout 300 python scripts/small_n_sweep.py ) > $O/small_n26.jsonl 2> $O/small_n26.err
python - <<'PY'
import json
for l in open('gpurun_out/small_n26.jsonl'):
    if l.startswith('#'): print(l.strip()); continue
    r = json.loads(l); print(r['precision'], r['n'], f"{r['us_per_step_k2000']:.2f} us/step (k=2000)  {r['us_per_step_k2_with_sync']:.2f} (k=2 + sync)")
PY
