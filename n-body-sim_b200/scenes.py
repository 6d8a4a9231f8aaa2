"""Seeded scene generators for the five BASELINE.json configs.

Only scene 1 comes from the reference (resetScenario('galaxy'),
/root/reference/index.html:749-759, whose unseeded Math.random is replaced by a
seeded stream so every implementation consumes the same bytes).  Scenes 2-5 are
this build's definitions of the names in BASELINE.json.  Everything is
generated in float64 and rounded to float32 by the caller for fp32 runs.

RNG: splitmix64 used as a counter-based generator.  Draw k (k >= 0) of a stream
with seed s is mix(s + (k+1)*GAMMA) -> top 53 bits -> [0,1).  Scene 1 uses one
sequential stream in the reference's draw order; scenes 2-5 give each body its
own sub-stream (draw index = body*STRIDE + t) so rejection loops vectorise and
the result does not depend on evaluation order.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

_GAMMA = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)
SEED_BASE = 0x5EED0000
STRIDE = 256  # draws reserved per body in scenes 2-5


def splitmix64_uniform(seed: int, index) -> np.ndarray:
    """Uniform doubles in [0,1): draw number `index` (array, >= 0) of stream `seed`."""
    with np.errstate(over="ignore"):
        k = np.asarray(index, dtype=np.uint64)
        z = np.uint64(seed) + (k + np.uint64(1)) * _GAMMA
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        z = z ^ (z >> np.uint64(31))
    return (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def _body_draw(seed: int, bodies: np.ndarray, t: int) -> np.ndarray:
    return splitmix64_uniform(seed, bodies.astype(np.uint64) * np.uint64(STRIDE) + np.uint64(t))


@dataclass
class Scene:
    name: str
    pos_mass: np.ndarray          # [n,4] float64: x,y,z,m
    vel: np.ndarray               # [n,4] float64
    temp: np.ndarray              # [n]   float64
    G: float
    softening: float
    dt: float
    meta: dict = field(default_factory=dict)

    @property
    def n(self) -> int:
        return self.pos_mass.shape[0]

    def astype(self, dtype):
        return (np.ascontiguousarray(self.pos_mass, dtype=dtype), np.ascontiguousarray(self.vel, dtype=dtype),
                np.ascontiguousarray(self.temp, dtype=dtype))


def _zeros(n):
    return np.zeros((n, 4)), np.zeros((n, 4)), np.zeros(n)


# -- config 1: the reference's default scene ------------------------------------
def galaxy_default(n_disc: int = 150, seed: int = SEED_BASE + 1, G: float = 0.5) -> Scene:
    """resetScenario('galaxy'), idx:749-759: a mass-500 centre and n_disc bodies
    on circular orbits.  Draw order per disc body: angle, dist, y, type, mass."""
    n = n_disc + 1
    pm, vel, temp = _zeros(n)
    pm[0] = (0.0, 0.0, 0.0, 500.0)          # idx:750: Body(500, V3(0,0,0), V3(0,0,0), 'metal', 1000)
    temp[0] = 1000.0
    u = splitmix64_uniform(seed, np.arange(5 * n_disc)).reshape(n_disc, 5)
    angle = u[:, 0] * np.pi * 2              # idx:752
    dist = 30 + u[:, 1] * 60                 # idx:753
    velocity = np.sqrt(G * 500 / dist)       # idx:754
    pm[1:, 0] = np.cos(angle) * dist         # idx:755
    pm[1:, 1] = (u[:, 2] - 0.5) * 2
    pm[1:, 2] = np.sin(angle) * dist
    vel[1:, 0] = -np.sin(angle) * velocity   # idx:756
    vel[1:, 2] = np.cos(angle) * velocity
    types = np.where(u[:, 3] > 0.8, "ice", "rock")  # idx:757
    pm[1:, 3] = u[:, 4] * 2 + 0.5            # idx:758
    return Scene("galaxy_default", pm, vel, temp, G=G, softening=0.5, dt=0.008,
                 meta={"types": ["metal"] + types.tolist(), "config": 1})


# -- config 2: Plummer sphere ----------------------------------------------------
def plummer(n: int = 16384, seed: int = SEED_BASE + 2) -> Scene:
    """Equal masses 1/n, M = a = G = 1, Aarseth-Henon-Wielen sampling, COM frame."""
    pm, vel, temp = _zeros(n)
    b = np.arange(n)
    # radius: r = (X^(-2/3) - 1)^(-1/2), redrawn while r > 25 (draws t = 0..7)
    r = np.zeros(n)
    pending = np.ones(n, dtype=bool)
    for t in range(8):
        idx = b[pending]
        x = np.maximum(_body_draw(seed, idx, t), 1e-300)
        rr = 1.0 / np.sqrt(x ** (-2.0 / 3.0) - 1.0 + 1e-300)
        ok = rr <= 25.0
        r[idx[ok]] = rr[ok]
        pending[idx[ok]] = False
        if not pending.any():
            break
    r[pending] = 1.0
    cz = 1.0 - 2.0 * _body_draw(seed, b, 8)
    ph = 2.0 * np.pi * _body_draw(seed, b, 9)
    s = np.sqrt(np.maximum(0.0, 1.0 - cz * cz))
    pm[:, 0], pm[:, 1], pm[:, 2] = r * s * np.cos(ph), r * s * np.sin(ph), r * cz
    pm[:, 3] = 1.0 / n
    # speed: q by rejection on g(q) = q^2 (1-q^2)^3.5 < 0.1 (draw pairs t = 10+2a, 11+2a)
    q = np.zeros(n)
    pending[:] = True
    for a in range(110):
        idx = b[pending]
        x4 = _body_draw(seed, idx, 10 + 2 * a)
        x5 = _body_draw(seed, idx, 11 + 2 * a)
        ok = 0.1 * x5 < x4 * x4 * (1.0 - x4 * x4) ** 3.5
        q[idx[ok]] = x4[ok]
        pending[idx[ok]] = False
        if not pending.any():
            break
    v = q * np.sqrt(2.0) * (1.0 + r * r) ** (-0.25)
    cz = 1.0 - 2.0 * _body_draw(seed, b, 250)
    ph = 2.0 * np.pi * _body_draw(seed, b, 251)
    s = np.sqrt(np.maximum(0.0, 1.0 - cz * cz))
    vel[:, 0], vel[:, 1], vel[:, 2] = v * s * np.cos(ph), v * s * np.sin(ph), v * cz
    pm[:, :3] -= pm[:, :3].mean(axis=0)   # equal masses: COM = mean
    vel[:, :3] -= vel[:, :3].mean(axis=0)
    return Scene("plummer", pm, vel, temp, G=1.0, softening=0.01, dt=1.0 / 512, meta={"config": 2})


# -- config 3: two-galaxy collision -------------------------------------------------
def two_galaxies(n: int = 131072, seed: int = SEED_BASE + 3, G: float = 0.5) -> Scene:
    """Two copies of the reference's disc recipe (idx:750-758) scaled up: each a
    mass-500 centre plus n/2-1 disc bodies of total mass 50, on the approach
    geometry of the reference's 'collision' scenario (idx:761-762)."""
    assert n % 2 == 0 and n >= 4
    half = n // 2
    pm, vel, temp = _zeros(n)
    centres = np.array([[-120.0, 0.0, 10.0], [120.0, 0.0, -10.0]])
    bulk = np.array([[1.0, 0.0, 0.0], [-1.0, 0.0, 0.0]])
    for g in range(2):
        o = g * half
        pm[o] = (*centres[g], 500.0)
        vel[o, :3] = bulk[g]
        b = np.arange(o + 1, o + half)
        angle = _body_draw(seed, b, 0) * np.pi * 2
        dist = 30 + _body_draw(seed, b, 1) * 60
        speed = np.sqrt(G * 500 / dist)
        pm[b, 0] = centres[g, 0] + np.cos(angle) * dist
        pm[b, 1] = centres[g, 1] + (_body_draw(seed, b, 2) - 0.5) * 2
        pm[b, 2] = centres[g, 2] + np.sin(angle) * dist
        pm[b, 3] = 50.0 / (half - 1)
        vel[b, 0] = bulk[g, 0] - np.sin(angle) * speed
        vel[b, 2] = bulk[g, 2] + np.cos(angle) * speed
    return Scene("two_galaxies", pm, vel, temp, G=G, softening=0.5, dt=0.008, meta={"config": 3})


# -- config 4: uniform cube ------------------------------------------------------------
def uniform_cube(n: int = 1048576, seed: int = SEED_BASE + 4) -> Scene:
    pm, vel, temp = _zeros(n)
    b = np.arange(n)
    for c in range(3):
        pm[:, c] = _body_draw(seed, b, c) * 100.0 - 50.0
    pm[:, 3] = 1.0 / n
    return Scene("uniform_cube", pm, vel, temp, G=1.0, softening=1.0, dt=0.01, meta={"config": 4})


# -- config 5: cold collapse -------------------------------------------------------------
def cold_collapse(n: int = 4194304, seed: int = SEED_BASE + 5) -> Scene:
    pm, vel, temp = _zeros(n)
    b = np.arange(n)
    pending = np.ones(n, dtype=bool)
    for a in range(40):
        idx = b[pending]
        p = np.stack([_body_draw(seed, idx, 3 * a + c) * 2.0 - 1.0 for c in range(3)], axis=1)
        ok = (p * p).sum(axis=1) <= 1.0
        pm[idx[ok], :3] = p[ok]
        pending[idx[ok]] = False
        if not pending.any():
            break
    pm[:, 3] = 1.0 / n
    return Scene("cold_collapse", pm, vel, temp, G=1.0, softening=0.005, dt=1.0 / 1024, meta={"config": 5})


SCENES = {
    "galaxy_default": galaxy_default,
    "plummer": plummer,
    "two_galaxies": two_galaxies,
    "uniform_cube": uniform_cube,
    "cold_collapse": cold_collapse,
}
