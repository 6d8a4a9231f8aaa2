#!/bin/bash
# full GPU suite on the pairs-once build; fp32 mid-size shapes (R=2 kernel) for reference
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest24.log 2>&1; echo "pytest rc=$?"
tail -5 $O/pytest24.log | cut -c1-250
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke24.log 2>&1; echo "smoke rc=$?"
for r in 4 2; do
NBODY_R=$r timeout 300 python - <<'PY'
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.getcwd())
import nbody_b200 as nb
for uni in ("1", "0"):
    os.environ["NBODY_UNIFORM"] = uni
    for n in (1024, 2048, 4096, 8192, 16384, 32768):
        s = nb.scenes.plummer(n)
        pm, vel, _ = s.astype(np.float32)
        with nb.Engine(n, precision="f32") as e:
            e.upload(pm, vel); e.step(s.dt, s.G, s.softening, 10); e.sync()
            k = 2000 if n <= 8192 else 400
            t0 = time.perf_counter(); e.step(s.dt, s.G, s.softening, k); e.sync()
            us = (time.perf_counter() - t0) / k * 1e6
            i = e.info()
        print(json.dumps({"R": os.environ["NBODY_R"], "uniform": i["uniform_mass"], "n": n, "us_per_step": round(us, 2), "j_chunks": i["j_chunks"], "i_tile": i["i_tile"]}), flush=True)
PY
done | tee $O/f32_midsize_r.jsonl
