#!/bin/bash
# fp64 kernel: equal-mass variant and two bodies per thread, on config 2's scene (Plummer, masses 1/N)
mkdir -p gpurun_out
rm -f gpurun_out/f64_r_uniform.jsonl
for cfg in "1 0" "1 1" "2 0" "2 1"; do
  set -- $cfg
  NBODY_F64_R=$1 NBODY_UNIFORM=$2 timeout 300 python scripts/f64_sweep.py --math 2 --units 0 --sizes 4096,16384,65536 --tag "R=$1 uniform=$2" --out gpurun_out/x.jsonl > gpurun_out/f64_ru.log 2>&1 || tail -5 gpurun_out/f64_ru.log
  cat gpurun_out/x.jsonl >> gpurun_out/f64_r_uniform.jsonl
done
rm -f gpurun_out/x.jsonl
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/pytest17.log 2>&1; echo "pytest rc=$?"
NBODY_F64_R=2 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py -m gpu -q -x -k "f64 or small or tiny" > gpurun_out/pytest17_r2.log 2>&1; echo "pytest R=2 rc=$?"
tail -3 gpurun_out/pytest17.log; tail -3 gpurun_out/pytest17_r2.log; cut -c1-260 gpurun_out/f64_r_uniform.jsonl
