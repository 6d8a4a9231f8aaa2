#!/bin/bash
# Round-1g (3): one-record small-system I/O + nbody_step_host: the whole suite, config 1 timed four ways (and with NBODY_SMALL_IO=0)
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > $O/pytest32.log 2>&1; echo "pytest rc=$?"
tail -n 15 $O/pytest32.log | cut -c1-300
timeout 200 python scripts/config1_bench.py > $O/config1_32.json 2> $O/config1_32.err; echo "config1 rc=$?"
cut -c1-1200 $O/config1_32.json; tail -n 5 $O/config1_32.err
NBODY_SMALL_IO=0 timeout 200 python scripts/config1_bench.py > $O/config1_32_old_io.json 2> $O/config1_32_old_io.err; echo "config1 old rc=$?"
cut -c1-1200 $O/config1_32_old_io.json
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > $O/bench32.json 2> $O/bench32.err; echo "bench rc=$?"
cut -c1-250 $O/bench32.json
