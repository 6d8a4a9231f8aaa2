mkdir -p gpurun_out; : > gpurun_out/r2k_variants.jsonl
for v in "" c3 s6 s2 sg; do
  if [ -n "$v" ]; then export NBODY_LIB=$PWD/n-body-sim_b200/variants/libnbody_$v.so; else unset NBODY_LIB; fi
  LAB_TAG=${v:-default} timeout 300 python scripts/sym_lab.py sym >> gpurun_out/r2k_variants.jsonl 2>> gpurun_out/r2k_variants.err
done
unset NBODY_LIB
python - <<'PY'
import json
for ln in open("gpurun_out/r2k_variants.jsonl"):
    d = json.loads(ln)
    print(f"{d['variant']:28s} general={int(d['general_masses'])} {d['force_ms']:8.2f} ms {d['G_int_per_s']:8.1f} G/s symR={d['sym_R']}")
PY
tail -3 gpurun_out/r2k_variants.err
python -c "import __graft_entry__ as g; g.smoke()"
