"""Generates the committed fixtures under tests/golden/ from the CPU oracle.

The reference (/root/reference/index.html) is a browser page with no tests and
no runnable engine in this image, so these vectors are produced by the oracle
restatement itself: they pin the oracle against regressions and give the GPU
tests fixed bytes to compare with; they do NOT pin the oracle against the
reference (make_reference_golden.py's vectors do).   Run:  python tests/golden/make_golden.py
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


def render_inputs(scene):
    """The default scene as the page would show it (radii idx:221 x3 so bodies cover pixels at 96x54,
    per-body seeds standing in for Math.random()*100, idx:492; the centre is metal at 1000 K, idx:750)."""
    n = scene.n
    dens = np.where(np.arange(n) == 0, 3.0, 1.0)
    radius = 3.0 * (3 * scene.pos_mass[:, 3] / (4 * np.pi * dens)) ** (1 / 3)
    types = np.where(np.arange(n) == 0, 1, np.where(np.arange(n) % 5 == 0, 2, 0)).astype(np.int32)
    seed = (nb.scenes.splitmix64_uniform(0x5EED0F00, np.arange(n)) * 100).astype(np.float32)
    temp = scene.temp.copy()
    temp[1::7] = 30.0      # a few warm bodies: magma cracks (idx:188-191)
    temp[2::11] = 120.0    # and a few glowing ones (idx:194)
    return radius, temp, types, seed


def emit_render(name, scene, width, height):
    radius, temp, types, seed = render_inputs(scene)
    rgba, ids = oracle.render(scene.pos_mass, radius, temp, types, seed, width, height)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), rgba=rgba, ids=ids, width=width, height=height)
    print(name, width, height, int((ids >= 0).sum()), "covered pixels")


if __name__ == "__main__":
    emit_render("render_galaxy_96x54", nb.scenes.galaxy_default(), 96, 54)
    emit("galaxy151_f64", nb.scenes.galaxy_default(), 100)     # config 1 scene, 100 x integrate(0.008)
    emit("plummer512_f64", nb.scenes.plummer(512), 20)         # config 2 recipe, reduced n
