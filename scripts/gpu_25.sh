#!/bin/bash
# fp32 j-tile size (512 vs 128 bodies) by system size; GPU suite on the build
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > $O/pytest25.log 2>&1; echo "pytest rc=$?"
tail -5 $O/pytest25.log | cut -c1-250
for tile in 512 128; do
NBODY_F32_TILE=$tile timeout 300 python - <<'PY'
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.getcwd())
import nbody_b200 as nb
for uni in ("1", "0"):
    os.environ["NBODY_UNIFORM"] = uni
    for n in (600, 1024, 2048, 4096, 8192, 16384, 32768, 65536, 131072):
        s = nb.scenes.plummer(n)
        pm, vel, _ = s.astype(np.float32)
        with nb.Engine(n, precision="f32") as e:
            e.upload(pm, vel); e.step(s.dt, s.G, s.softening, 10); e.sync()
            k = 2000 if n <= 8192 else (400 if n <= 32768 else 40)
            t0 = time.perf_counter(); e.step(s.dt, s.G, s.softening, k); e.sync()
            us = (time.perf_counter() - t0) / k * 1e6
            i = e.info()
        print(json.dumps({"tile": int(os.environ["NBODY_F32_TILE"]), "uniform": i["uniform_mass"], "n": n, "us_per_step": round(us, 2),
                          "g_int_per_s": round(n * n / us * 1e-3, 1), "j_chunks": i["j_chunks"], "i_tile": i["i_tile"]}), flush=True)
PY
done > $O/f32_tile_sweep.jsonl
python - <<'PY'
import json, collections
t = collections.defaultdict(dict)
for l in open('gpurun_out/f32_tile_sweep.jsonl'):
    r = json.loads(l); t[(r['uniform'], r['n'])][r['tile']] = (r['us_per_step'], r['j_chunks'])
for k in sorted(t):
    print(k, '  '.join(f"tile{a}: {v[0]}us(c{v[1]})" for a, v in t[k].items()))
PY
