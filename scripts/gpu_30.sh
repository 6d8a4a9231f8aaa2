#!/bin/bash
# Round-1g: final-build confirmation on one GPU: suite, smoke, both bench arms, ncu of the all-SM kernel
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -q > $O/pytest30.log 2>&1; echo "pytest rc=$?"
tail -n 4 $O/pytest30.log | cut -c1-300
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke30.log 2>&1; echo "smoke rc=$?"
timeout 300 python bench.py > $O/bench30.json 2> $O/bench30.err; echo "bench rc=$?"
cut -c1-300 $O/bench30.json
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 > $O/ref30.json 2> $O/ref30.err; echo "ref rc=$?"
cut -c1-200 $O/ref30.json
cat > /tmp/med.py <<'PY'
import numpy as np, nbody_b200 as nb
s = nb.scenes.plummer(2048)
pm, vel, _ = s.astype(np.float32)
with nb.Engine(2048, precision="f32") as e:
    e.upload(pm, vel); e.step(s.dt, s.G, s.softening, 4); e.sync(); print(e.info())
PY
PYTHONPATH=. timeout 150 ncu --set full --clock-control none --import-source on -k regex:medium -c 1 -o $O/medium30 -f python /tmp/med.py > $O/ncu_medium30.log 2>&1; echo "ncu rc=$?"
tail -n 3 $O/ncu_medium30.log | cut -c1-200
