#!/bin/bash
# fp64 pairs-once pass (force_f64_sym.cuh): parity, then timing against the one-sided kernel
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py -m gpu -q -x -k "f64 or small or tiny or ragged or golden or fused or thermal or energy or zero" > $O/pytest21.log 2>&1; echo "pytest rc=$?"
tail -15 $O/pytest21.log | cut -c1-250
rm -f $O/f64_sym.jsonl
for cfg in "1 0" "1 1" "1 2" "1 4" "0 0"; do
  set -- $cfg
  NBODY_F64_SYM=$1 NBODY_F64_SYM_K=$2 timeout 300 python scripts/f64_sweep.py --math 2 --units 0 --sizes 4096,16384,32768,65536 --tag "sym=$1 k=$2" --out $O/x.jsonl > $O/f64_sym.log 2>&1 || tail -5 $O/f64_sym.log
  cat $O/x.jsonl >> $O/f64_sym.jsonl
done
rm -f $O/x.jsonl
python - <<'PY'
import json
for l in open('gpurun_out/f64_sym.jsonl'):
    r = json.loads(l)
    print(r['tag'], r['n'], 'k', r['pairs_once_k'], 'uni', r['uniform_kernel'], f"{r['us_per_step']:.1f} us  {r['g_int_per_s']:.0f} G  err {r['force_err_n4096']:.2e}")
PY
