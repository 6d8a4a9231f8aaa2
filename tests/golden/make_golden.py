"""Generates the committed fixtures under tests/golden/ from the CPU oracle.

The reference (/root/reference/index.html) is a browser page with no tests and
no runnable engine in this image, so these vectors are produced by the oracle
restatement itself: they pin the oracle against regressions and give the GPU
tests fixed bytes to compare with; they do NOT pin the oracle against the
reference (parity unpinned).   Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402
from oracle import oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def emit(name, scene, nsteps):
    p, v, a, t = oracle.integrate(scene.pos_mass, scene.vel, np.zeros_like(scene.vel), scene.temp, scene.dt,
                                  scene.G, scene.softening, nsteps=nsteps)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), pos_mass0=scene.pos_mass, vel0=scene.vel,
                        temp0=scene.temp, pos_mass=p, vel=v, acc=a, temp=t, dt=scene.dt, G=scene.G,
                        softening=scene.softening, nsteps=nsteps)
    print(name, scene.n, nsteps)


if __name__ == "__main__":
    emit("galaxy151_f64", nb.scenes.galaxy_default(), 100)     # config 1 scene, 100 x integrate(0.008)
    emit("plummer512_f64", nb.scenes.plummer(512), 20)         # config 2 recipe, reduced n
