#!/bin/bash
# fp64 pairs-once pass: shape variants (unroll, i-bodies per thread), general masses, full ncu capture
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/f64_sym_variants.jsonl
run() {  # tag, env...
  tag=$1; shift
  env "$@" timeout 300 python scripts/f64_sweep.py --math 2 --units 0 --sizes 8192,16384,65536 --tag "$tag" --out $O/x.jsonl > $O/f64_symv.log 2>&1 || tail -5 $O/f64_symv.log
  cat $O/x.jsonl >> $O/f64_sym_variants.jsonl
}
run "default (R=4, unroll 2)" X=1
run "unroll 1" NBODY_LIB=$PWD/n-body-sim_b200/variants/libnbody_sym_u1.so
run "unroll 4" NBODY_LIB=$PWD/n-body-sim_b200/variants/libnbody_sym_u4.so
run "R=2" NBODY_LIB=$PWD/n-body-sim_b200/variants/libnbody_sym_r2.so
run "general masses" NBODY_UNIFORM=0
run "general masses, one-sided" NBODY_UNIFORM=0 NBODY_F64_SYM=0
rm -f $O/x.jsonl
python - <<'PY'
import json
for l in open('gpurun_out/f64_sym_variants.jsonl'):
    r = json.loads(l)
    print(r['tag'], '|', r['n'], 'k', r['pairs_once_k'], 'uni', r['uniform_kernel'], f"{r['us_per_step']:.1f} us  {r['g_int_per_s']:.0f} G")
PY
timeout 600 ncu --clock-control none --set full --import-source on -k regex:force_f64_sym_kernel -s 4 -c 1 -o $O/full_force_f64_sym -f python scripts/f64_sweep.py --math 2 --units 0 --sizes 65536 --out $O/x.jsonl > $O/ncu_f64_sym.log 2>&1; echo "ncu rc=$?"
timeout 600 ncu --clock-control none --set full --import-source on -k regex:f64_sym -s 8 -c 2 -o $O/full_force_f64_sym_n16384 -f python scripts/f64_sweep.py --math 2 --units 0 --sizes 16384 --out $O/x.jsonl > $O/ncu_f64_sym2.log 2>&1; echo "ncu rc=$?"
rm -f $O/x.jsonl
