#!/bin/bash
# round-1f: fp64 kernel with the 6-instruction s^-1.5 (inv_cube_f64); evidence refresh on the final build
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q > $O/pytest20.log 2>&1; echo "pytest rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke20.log 2>&1; echo "smoke rc=$?"
timeout 300 python scripts/f64_sweep.py --math 1,2 --units 0 --sizes 4096,16384,32768,65536 --tag "inv_cube_f64" --out $O/f64_invcube.jsonl > $O/f64_invcube.log 2>&1; echo "f64 rc=$?"
timeout 300 python scripts/small_n_sweep.py > $O/small_n20.jsonl 2> $O/small_n20.err; echo "small rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref20.json 2> $O/bench_ref20.err; echo "ref rc=$?"
timeout 600 python bench.py > $O/bench20.json 2> $O/bench20.err; echo "bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_n1m_20.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > $O/ncu_launches20.log 2>&1; echo "ncu launches rc=$?"
timeout 600 python scripts/energy_drift.py > $O/energy_drift20.json 2> $O/energy_drift20.err; echo "drift rc=$?"
tail -3 $O/pytest20.log | cut -c1-200; tail -1 $O/smoke20.log; cut -c1-230 $O/f64_invcube.jsonl; cut -c1-500 $O/bench20.json; cut -c1-300 $O/energy_drift20.json
