#!/bin/bash
# Round-1d evidence pass: full GPU test suite, bench lines, then ncu (each only after its command ran clean).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" > gpurun_out/probe7.log
timeout 600 python bench.py > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/probe7.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>gpurun_out/bench_ref.err; echo "ref rc=$?" >> gpurun_out/probe7.log
timeout 600 python scripts/contacts_bench.py --big > gpurun_out/contacts.log 2>&1; echo "contacts rc=$?" >> gpurun_out/probe7.log
timeout 600 python scripts/trails_bench.py > gpurun_out/trails.log 2>&1; echo "trails rc=$?" >> gpurun_out/probe7.log
timeout 600 python scripts/render_bench.py > gpurun_out/render.log 2>&1; echo "render rc=$?" >> gpurun_out/probe7.log
# launch lists
NCU="ncu --clock-control none"
timeout 900 $NCU --metrics gpu__time_duration.sum -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --no-cpu --no-e2e --steps 2 --warmup 3 --n 131072 > gpurun_out/ncu_bench.log 2>&1
CONTACTS_QUICK= timeout 900 $NCU --metrics gpu__time_duration.sum -c 200 --csv --log-file gpurun_out/launches_contacts.csv python scripts/contacts_bench.py --reps 1 --out gpurun_out/contacts_under_ncu.json > gpurun_out/ncu_contacts.log 2>&1
timeout 900 $NCU --metrics gpu__time_duration.sum -k regex:"render|trail|history" -c 200 --csv --log-file gpurun_out/launches_render.csv python scripts/render_bench.py --reps 1 --out gpurun_out/render_under_ncu.json > gpurun_out/ncu_render.log 2>&1
# full-set captures, one launch each
timeout 900 $NCU --set full --import-source on -k regex:force_f32_kernel -s 3 -c 1 -o gpurun_out/full_force_uniform_n1m -f python bench.py --no-cpu --no-e2e --steps 1 --warmup 3 > gpurun_out/ncu_full1.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:force_f32_near_kernel -s 1 -c 1 -o gpurun_out/full_force_near -f python scripts/contacts_bench.py --reps 1 --out gpurun_out/x.json > gpurun_out/ncu_full2.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:contact_pairs_kernel -s 1 -c 1 -o gpurun_out/full_contact_pairs -f python scripts/contacts_bench.py --reps 1 --out gpurun_out/x.json > gpurun_out/ncu_full3.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:trail_ribbons_kernel -s 2 -c 1 -o gpurun_out/full_ribbons -f python scripts/trails_bench.py --reps 2 --out gpurun_out/x.json > gpurun_out/ncu_full4.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:render_tile_kernel -s 40 -c 1 -o gpurun_out/full_render_tile -f python scripts/render_bench.py --reps 1 --out gpurun_out/x.json > gpurun_out/ncu_full5.log 2>&1
rm -f gpurun_out/x.json
cat gpurun_out/probe7.log; tail -3 gpurun_out/pytest_gpu.log; cut -c1-400 gpurun_out/bench.log; cut -c1-250 gpurun_out/bench_ref.log; ls -la gpurun_out/*.ncu-rep
