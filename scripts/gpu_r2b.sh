#!/bin/bash
# round 2, second GPU session: sym tests again, kernel-variant builds of both fp32 force kernels at N=1M
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sym32.py -x -q > gpurun_out/r2b_sym_tests.txt 2>&1; echo "sym tests rc=$?"; tail -4 gpurun_out/r2b_sym_tests.txt
: > gpurun_out/r2b_variants.jsonl
for v in "" u8c3 u12c3 u16c2 u4c3s3; do
  if [ -n "$v" ]; then export NBODY_LIB=$PWD/n-body-sim_b200/variants/libnbody_$v.so; else unset NBODY_LIB; fi
  LAB_TAG=${v:-default} timeout 300 python scripts/sym_lab.py onesided >> gpurun_out/r2b_variants.jsonl 2>> gpurun_out/r2b_variants.err
  LAB_TAG=${v:-default} timeout 300 python scripts/sym_lab.py sym >> gpurun_out/r2b_variants.jsonl 2>> gpurun_out/r2b_variants.err
done
unset NBODY_LIB
python - <<'PY'
import json
for ln in open("gpurun_out/r2b_variants.jsonl"):
    d = json.loads(ln)
    print(f"{d['variant']:28s} general={int(d['general_masses'])} {d['force_ms']:8.2f} ms {d['G_int_per_s']:8.1f} G/s symR={d['sym_R']}")
PY
tail -3 gpurun_out/r2b_variants.err
