#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/energy_drift.py > gpurun_out/energy_drift.json 2> gpurun_out/energy_drift.err; echo "drift rc=$?" > gpurun_out/probe5.log
timeout 300 python bench.py --no-cpu --no-e2e --steps 2 --warmup 3 --n 131072 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_r1b.csv python bench.py --no-cpu --no-e2e --steps 2 --warmup 3 --n 131072 > gpurun_out/ncu_launch.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:force_f32 -s 3 -c 2 -o gpurun_out/prof_r1b_force_f32_n131072 -f python bench.py --no-cpu --no-e2e --steps 2 --warmup 3 --n 131072 > gpurun_out/ncu_full.log 2>&1
echo "ncu small rc=$?" >> gpurun_out/probe5.log
timeout 300 python bench.py --no-cpu --no-e2e --steps 1 --warmup 3 > gpurun_out/plain1m.log 2>&1 &&
timeout 900 ncu --set full --clock-control none -k regex:force_f32 -s 3 -c 1 -o gpurun_out/prof_r1b_force_f32_n1m -f python bench.py --no-cpu --no-e2e --steps 1 --warmup 3 > gpurun_out/ncu_full_1m.log 2>&1
echo "ncu 1m rc=$?" >> gpurun_out/probe5.log
cat gpurun_out/probe5.log; cut -c1-700 gpurun_out/energy_drift.json; tail -3 gpurun_out/energy_drift.err
