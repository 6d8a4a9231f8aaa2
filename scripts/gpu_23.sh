#!/bin/bash
# fp64 pairs-once pass: i-bodies per thread by system size
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "f64" > $O/pytest23.log 2>&1; echo "pytest rc=$?"
tail -5 $O/pytest23.log | cut -c1-250
rm -f $O/f64_sym_r.jsonl
run() {  # tag, env...
  tag=$1; shift
  env "$@" timeout 300 python scripts/f64_sweep.py --math 2 --units 0 --sizes 1024,2048,4096,8192,12288,16384,24576,32768 --tag "$tag" --out $O/x.jsonl > $O/f64_symr.log 2>&1 || tail -5 $O/f64_symr.log
  cat $O/x.jsonl >> $O/f64_sym_r.jsonl
}
run "auto" X=1
run "R=1" NBODY_F64_SYM_R=1
run "R=2" NBODY_F64_SYM_R=2
run "R=4" NBODY_F64_SYM_R=4
run "one-sided" NBODY_F64_SYM=0
rm -f $O/x.jsonl
python - <<'PY'
import json, collections
t = collections.defaultdict(dict)
for l in open('gpurun_out/f64_sym_r.jsonl'):
    r = json.loads(l); t[r['n']][r['tag']] = (r['us_per_step'], r['pairs_once_k'])
for n in sorted(t):
    print(n, '  '.join(f"{k}: {v[0]:.1f}us(k{v[1]})" for k, v in t[n].items()))
PY
