#!/bin/bash
# The GPU-side command lists of this repository, one section per purpose (replaces the per-session scripts of round 1).
#   gpurun [--gpus N] -- 'bash scripts/gpu_suite.sh <section> [tag]'
# Every section writes under gpurun_out/<tag>_*; copy what should be judged into profiles/.
#   tests      pytest -m gpu (all of it; multi-GPU tests run when the lease has the devices)
#   multitests the multi-GPU tests only (bit-identity across GPU counts, rank mode under torchrun)
#   bench      bench.py on one GPU: the default line, the general-mass line, the reference arm
#   scale      bench.py under torchrun on every GPU count the lease allows (2, 4, 8), plus the single-process model
#   config1    configs[0]: the page's default scene four ways + the complete step with collisions (scripts/config1_bench.py)
#   energy     configs[2]: 10k steps of the two-galaxy scene with the energy-drift record
#   launches   ncu launch list of a short bench run (after the same command ran plain)
#   ncu        ncu --set full of the pairs-once force kernel, equal and general masses, N=262,144
#   checked    scripts/sanitize_cases.py under the CHECKED build (device-side asserts, make variant NAME=checked EXTRA=-DNB_DEBUG_CHECKS;
#              build it here first): what stands in for compute-sanitizer, which is closed on this pool
#   sanitize   compute-sanitizer <tool> (second argument: memcheck | racecheck | synccheck | initcheck) over small-N runs of every kernel family
set -u
section=${1:-tests}
tag=${2:-r2}
mkdir -p gpurun_out
ngpu=$(nvidia-smi -L | wc -l)
port=29511
export NBODY_EXCHANGE_TIMEOUT_S=${NBODY_EXCHANGE_TIMEOUT_S:-30}
torchrun_n() { n=$1; shift; port=$((port + 1)); timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port $port "$@"; }

case "$section" in
multitests)   # only the tests that need several devices (an 8-GPU lease is charged eightfold)
  timeout 1200 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_sym32.py tests/test_gpu_rank_mode.py tests/test_gpu_edges.py -m gpu -q -rs -k "multi or rank_mode" > gpurun_out/${tag}_pytest_multi_${ngpu}gpu.txt 2>&1; echo "pytest rc=$?"
  tail -8 gpurun_out/${tag}_pytest_multi_${ngpu}gpu.txt ;;
tests)
  timeout 1500 python -m pytest tests -m gpu -q -rs > gpurun_out/${tag}_pytest_gpu_${ngpu}gpu.txt 2>&1; echo "pytest rc=$?"
  tail -15 gpurun_out/${tag}_pytest_gpu_${ngpu}gpu.txt ;;
bench)
  timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${tag}_bench_reference_arm.json 2> gpurun_out/${tag}_bench_reference_arm.err; echo "reference arm rc=$?"
  timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench_n1.err; echo "bench rc=$?"
  timeout 900 python bench.py --steps 10 --warmup 3 --general-masses --no-cpu > gpurun_out/${tag}_bench_n1_general_masses.json 2> gpurun_out/${tag}_bench_n1_general.err; echo "bench general rc=$?"
  cut -c1-1500 gpurun_out/${tag}_bench_n1.json; tail -3 gpurun_out/${tag}_bench_n1.err ;;
scale)
  counts=${3:-"2 4 8"}
  top=0
  for n in $counts; do
    [ "$n" -le "$ngpu" ] || continue
    top=$n
    torchrun_n $n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/${tag}_scale_${n}.json 2> gpurun_out/${tag}_scale_${n}.err; echo "torchrun x$n rc=$?"
  done
  if [ "$top" -ge 2 ]; then   # north_star's host model: ONE process driving all devices
    timeout 900 python bench.py --gpus $top --single-process --steps 20 --warmup 5 > gpurun_out/${tag}_scale_${top}_single_process.json 2> gpurun_out/${tag}_scale_${top}_single_process.err; echo "single process x$top rc=$?"
  fi
  if [ "$ngpu" -ge 8 ]; then
    torchrun_n 8 bench.py --gpus 8 --steps 5 --warmup 3 --scene cold_collapse --bodies 4194304 > gpurun_out/${tag}_config5_8gpu.json 2> gpurun_out/${tag}_config5_8gpu.err; echo "config 5 rc=$?"
  fi
  for f in gpurun_out/${tag}_scale_*.json gpurun_out/${tag}_config5_8gpu.json; do [ -s "$f" ] && python - "$f" <<'PY'
import json, sys
d = json.loads([ln for ln in open(sys.argv[1]) if ln.startswith("{")][-1])
print(sys.argv[1], "gpus", d["n_gpus"], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "fused", round(d["fused"]["value"], 1),
      "check", d["check"]["ok"], d["check"]["state_hash"], "err", d["check"]["sampled_acc_err"], "h2d", d["e2e"]["h2d_bytes_per_step"])
PY
  done ;;
config1)
  timeout 600 python scripts/config1_bench.py > gpurun_out/${tag}_config1.json 2> gpurun_out/${tag}_config1.err; echo "config1 rc=$?"
  python -m json.tool gpurun_out/${tag}_config1.json | head -40; tail -3 gpurun_out/${tag}_config1.err ;;
energy)
  timeout 900 python bench.py --energy --scene two_galaxies --bodies 131072 --steps 10000 > gpurun_out/${tag}_energy_drift_n131072_10k.json 2> gpurun_out/${tag}_energy.err; echo "energy rc=$?"
  cut -c1-700 gpurun_out/${tag}_energy_drift_n131072_10k.json ;;
launches)
  python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/${tag}_launches_plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_n1m.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/${tag}_launches_ncu.log 2>&1
  echo "launch list rc=$?"; tail -5 gpurun_out/${tag}_launches_n1m.csv ;;
ncu)
  export LAB_N=${LAB_N:-1048576}
  LAB_GENERAL=0 python scripts/sym_lab.py ncu > gpurun_out/${tag}_ncu_plain_u.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:force_f32_sym -s 1 -c 1 -f -o gpurun_out/${tag}_sym_uniform python scripts/sym_lab.py ncu > gpurun_out/${tag}_ncu_u.log 2>&1
  echo "ncu equal-mass rc=$?"
  LAB_GENERAL=1 python scripts/sym_lab.py ncu > gpurun_out/${tag}_ncu_plain_g.log 2>&1 &&
  LAB_GENERAL=1 ncu --set full --clock-control none --import-source on -k regex:force_f32_sym -s 1 -c 1 -f -o gpurun_out/${tag}_sym_general python scripts/sym_lab.py ncu > gpurun_out/${tag}_ncu_g.log 2>&1
  echo "ncu general rc=$?" ;;
checked)
  NBODY_LIB=$PWD/n-body-sim_b200/variants/libnbody_checked.so CASES_REPEAT=5 timeout 900 python scripts/sanitize_cases.py > gpurun_out/${tag}_checked_build_cases.txt 2>&1; echo "checked build rc=$?"
  tail -32 gpurun_out/${tag}_checked_build_cases.txt | cut -c1-200
  NBODY_LIB=$PWD/n-body-sim_b200/variants/libnbody_checked.so timeout 900 python -m pytest tests/test_gpu_sym32.py tests/test_gpu_world.py tests/test_gpu_edges.py tests/test_zz_no_alloc.py -q -x 2>&1 | tail -3 ;;
sanitize)
  tool=${2:-memcheck}; tag=${3:-r2}
  python scripts/sanitize_cases.py > gpurun_out/${tag}_sanitize_plain.log 2>&1 &&
  timeout 1700 compute-sanitizer --tool "$tool" --error-exitcode 99 --log-file gpurun_out/${tag}_sanitizer_${tool}.txt python scripts/sanitize_cases.py > gpurun_out/${tag}_sanitize_${tool}_run.log 2>&1
  echo "compute-sanitizer $tool rc=$?"; tail -12 gpurun_out/${tag}_sanitizer_${tool}.txt; tail -3 gpurun_out/${tag}_sanitize_${tool}_run.log ;;
*) echo "unknown section $section"; exit 2 ;;
esac
