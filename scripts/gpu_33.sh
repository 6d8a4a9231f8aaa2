#!/bin/bash
# Round-1g (4): zero-copy small-system I/O (n <= 1,024): the whole suite, config 1 timed four ways (and with NBODY_SMALL_IO=0)
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > $O/pytest33.log 2>&1; echo "pytest rc=$?"
tail -n 15 $O/pytest33.log | cut -c1-300
timeout 200 python scripts/config1_bench.py > $O/config1_33.json 2> $O/config1_33.err; echo "config1 rc=$?"
cut -c1-1200 $O/config1_33.json; tail -n 5 $O/config1_33.err
NBODY_SMALL_IO=0 timeout 200 python scripts/config1_bench.py > $O/config1_33_old_io.json 2> $O/config1_33_old_io.err; echo "config1 old rc=$?"
cut -c1-1200 $O/config1_33_old_io.json
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > $O/bench33.json 2> $O/bench33.err; echo "bench rc=$?"
cut -c1-250 $O/bench33.json
