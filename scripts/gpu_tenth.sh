#!/bin/bash
# fp64 kernel shape variants (unroll, ring depth) at the new default work-unit count, then one full ncu capture.
mkdir -p gpurun_out
rm -f gpurun_out/f64_variants.jsonl
timeout 300 python scripts/f64_sweep.py --math 2 --units 0 --sizes 4096,16384,65536 --out gpurun_out/x.jsonl > gpurun_out/f64v_default.log 2>&1; cat gpurun_out/x.jsonl >> gpurun_out/f64_variants.jsonl
for v in f64u1 f64u4 f64s2 f64s2t8 f64u4s2; do
  NBODY_LIB=$PWD/n-body-sim_b200/variants/libnbody_$v.so timeout 300 python scripts/f64_sweep.py --math 2 --units 0 --sizes 4096,16384,65536 --out gpurun_out/x.jsonl > gpurun_out/f64v_$v.log 2>&1
  cat gpurun_out/x.jsonl >> gpurun_out/f64_variants.jsonl
done
rm -f gpurun_out/x.jsonl
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest10.log 2>&1; echo "pytest rc=$?"
timeout 900 ncu --clock-control none --set full --import-source on -k regex:force_f64_kernel -s 4 -c 1 -o gpurun_out/full_force_f64 -f python scripts/f64_sweep.py --math 2 --units 0 --sizes 65536 --out gpurun_out/x.jsonl > gpurun_out/ncu_f64.log 2>&1; echo "ncu rc=$?"
rm -f gpurun_out/x.jsonl
tail -2 gpurun_out/pytest10.log; cut -c1-170 gpurun_out/f64_variants.jsonl
