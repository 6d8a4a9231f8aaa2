#!/bin/bash
# medium kernel up to 8,192 bodies in fp32; crossovers in place; full suite
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest28.log 2>&1; echo "pytest rc=$?"
tail -6 $O/pytest28.log | cut -c1-300
( echo "# default"; timeout 300 python scripts/small_n_sweep.py; echo "# NBODY_MEDIUM=0"; NBODY_MEDIUM=0 timeout 300 python scripts/small_n_sweep.py ) > $O/small_n28.jsonl 2> $O/small_n28.err
python - <<'PY'
import json
for l in open('gpurun_out/small_n28.jsonl'):
    if l.startswith('#'): print(l.strip()); continue
    r = json.loads(l); print(r['precision'], r['n'], f"{r['us_per_step_k2000']:.2f} us/step (k=2000)  {r['us_per_step_k2_with_sync']:.2f} (k=2 + sync)")
PY
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke28.log 2>&1; echo "smoke rc=$?"
