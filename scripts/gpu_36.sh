#!/bin/bash
# Round-1g (7): cluster kernel writes the caller's record itself (nbody_step_host); suite, config 1, both bench arms, smoke
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -q > $O/pytest36.log 2>&1; echo "pytest rc=$?"
tail -n 12 $O/pytest36.log | cut -c1-400
timeout 200 python scripts/config1_bench.py > $O/config1_36.json 2> $O/config1_36.err; echo "config1 rc=$?"
cut -c1-1200 $O/config1_36.json; tail -n 5 $O/config1_36.err
timeout 300 python bench.py > $O/bench36.json 2> $O/bench36.err; echo "bench rc=$?"
cut -c1-200 $O/bench36.json
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 > $O/ref36.json 2> $O/ref36.err; echo "ref rc=$?"
cut -c1-200 $O/ref36.json
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke36.log 2>&1; echo "smoke rc=$?"
