#!/bin/bash
# Round-1g (2): the tests against the reference-source vectors on the GPU, config 1 timed three ways, ncu of the cluster kernel
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -q > $O/pytest31.log 2>&1; echo "pytest rc=$?"
tail -n 4 $O/pytest31.log | cut -c1-300
timeout 200 python scripts/config1_bench.py > $O/config1_31.json 2> $O/config1_31.err; echo "config1 rc=$?"
cut -c1-900 $O/config1_31.json
cat > /tmp/small.py <<'PY'
import numpy as np, nbody_b200 as nb
s = nb.scenes.galaxy_default()
with nb.Engine(s.n, precision="f64") as e:
    e.upload(s.pos_mass, s.vel, None, s.temp); e.step(s.dt, s.G, s.softening, 2); e.sync(); e.step(s.dt, s.G, s.softening, 100); e.sync(); print(e.info())
PY
PYTHONPATH=. timeout 150 ncu --set full --clock-control none --import-source on -k regex:small_step -c 2 -o $O/small31 -f python /tmp/small.py > $O/ncu_small31.log 2>&1; echo "ncu rc=$?"
tail -n 3 $O/ncu_small31.log | cut -c1-200
