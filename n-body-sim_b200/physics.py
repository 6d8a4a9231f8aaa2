"""Host-side mirror of the reference's physics interface, backed by the GPU engine.

Same names, argument meaning and (absence of) error behaviour as the JS classes
in /root/reference/index.html: CONFIG (idx:52-75), Body (idx:208-231),
PhysicsSystem.addBody / integrate (idx:239-262).  `integrate(dt)` is the
gravity-only step contract (idx:247-262 with resolveCollisions, idx:256, left to
the host); `integrate(dt, collide=fn)` runs the split form so a host can keep
its own collision stage between force evaluation and kick-2 (idx:255-259).

All arithmetic happens in libnbody.so on the GPU; this module only marshals
Body objects to and from the typed arrays, like the JS shim in INTEGRATION.md.
"""
from __future__ import annotations

import math
import random
from types import SimpleNamespace
from typing import Callable, List, Optional

import numpy as np

from .engine import Engine

# CONFIG, idx:52-75 (physics-relevant fields; colours belong to the renderer)
CONFIG = SimpleNamespace(
    G=0.5, softening=0.5, dt=0.016, subSteps=2, maxBodies=300, trailLength=80,
    fractureThreshold=25, minFragmentMass=0.2, mergeTime=0.5,
    heatDecay=0.998, heatToGlow=3.0,
    materials={"rock": SimpleNamespace(density=1.0), "metal": SimpleNamespace(density=3.0),
               "ice": SimpleNamespace(density=0.5)},
)


class Vector3:
    """The four THREE.Vector3 members the hot path touches (idx:251,252,259,266,277,278)."""
    __slots__ = ("x", "y", "z")

    def __init__(self, x=0.0, y=0.0, z=0.0):
        self.x, self.y, self.z = float(x), float(y), float(z)

    def clone(self):
        return Vector3(self.x, self.y, self.z)

    def set(self, x, y, z):
        self.x, self.y, self.z = float(x), float(y), float(z)
        return self

    def __iter__(self):
        return iter((self.x, self.y, self.z))

    def __repr__(self):
        return f"Vector3({self.x!r}, {self.y!r}, {self.z!r})"


def V3(x, y, z):  # idx:77
    return Vector3(x, y, z)


class Body:
    """Body, idx:208-231."""

    def __init__(self, mass, pos, vel, type="rock", temp=0):
        self.id = "".join(random.choice("0123456789abcdefghijklmnopqrstuvwxyz") for _ in range(9))
        self.mass = float(mass)
        self.invMass = 1 / mass if mass > 0 else 0
        self.pos = pos.clone()
        self.prevPos = pos.clone()
        self.vel = vel.clone()
        self.acc = Vector3(0, 0, 0)
        self.type = type
        self.matProps = CONFIG.materials[type]
        self.radius = math.pow((3 * mass) / (4 * math.pi * self.matProps.density), 1 / 3)
        self.temperature = float(temp)
        self.history = []


class PhysicsSystem:
    """PhysicsSystem, idx:233-262, with the two O(N) loops and computeGravity on the GPU."""

    def __init__(self, precision: str = "f64", capacity: Optional[int] = None, n_gpus: int = 1):
        self.bodies: List[Body] = []
        self.contactPairs = {}
        self._cap = int(capacity or max(CONFIG.maxBodies, 1024))
        self._engine = Engine(self._cap, precision=precision, n_gpus=n_gpus)

    @property
    def engine(self) -> Engine:
        return self._engine

    def addBody(self, body: Body) -> Body:  # idx:239-245
        if len(self.bodies) >= CONFIG.maxBodies:
            self.bodies.pop(0)
        self.bodies.append(body)
        return body

    # -- marshalling: Body objects <-> typed arrays --------------------------------
    def _pack(self):
        n = len(self.bodies)
        dt = self._engine.dtype
        pm, vel, acc = np.empty((n, 4), dt), np.zeros((n, 4), dt), np.zeros((n, 4), dt)
        temp = np.empty(n, dt)
        for i, b in enumerate(self.bodies):
            pm[i] = (b.pos.x, b.pos.y, b.pos.z, b.mass)
            vel[i, :3] = (b.vel.x, b.vel.y, b.vel.z)
            acc[i, :3] = (b.acc.x, b.acc.y, b.acc.z)
            temp[i] = b.temperature
        return pm, vel, acc, temp

    def _unpack(self, out, what=("pos_mass", "vel", "acc", "temp")):
        for i, b in enumerate(self.bodies):
            if "pos_mass" in what:
                b.pos.set(*out["pos_mass"][i, :3])
            if "vel" in what:
                b.vel.set(*out["vel"][i, :3])
            if "acc" in what:
                b.acc.set(*out["acc"][i, :3])
            if "temp" in what:
                b.temperature = float(out["temp"][i])

    def integrate(self, dt: float, collide: Optional[Callable[["PhysicsSystem", float], None]] = None,
                  nsteps: int = 1) -> None:
        """integrate(dt), idx:247-262.  Reads CONFIG.G / CONFIG.softening at call
        time like the reference (idx:268-269).  With `collide`, runs
        kick1+drift -> force -> collide(self, dt) on the host -> kick-2."""
        if not self.bodies:
            return
        if len(self.bodies) > self._cap:
            raise ValueError(f"{len(self.bodies)} bodies exceed the engine capacity {self._cap}")
        eng = self._engine
        eng.set_heat_decay(CONFIG.heatDecay)
        eng.upload(*self._pack())
        if collide is None:
            eng.step(dt, CONFIG.G, CONFIG.softening, nsteps)
            self._unpack(eng.download(temp=True))
            return
        for _ in range(nsteps):
            eng.kick1_drift(dt)
            eng.force(CONFIG.G, CONFIG.softening)
            self._unpack(eng.download(temp=True))
            collide(self, dt)                       # the host's own resolveCollisions (idx:256)
            eng.upload(*self._pack())               # collisions may edit pos/vel and the body list
            eng.kick2(dt, CONFIG.heatDecay)
            self._unpack(eng.download(pos_mass=False, acc=False, temp=True), what=("vel", "temp"))

    def computeGravity(self) -> None:
        """computeGravity(), idx:264-291: fills body.acc only."""
        if not self.bodies:
            return
        eng = self._engine
        eng.upload(*self._pack())
        eng.force(CONFIG.G, CONFIG.softening)
        self._unpack(eng.download(pos_mass=False, vel=False), what=("acc",))


def resetScenario(physics: PhysicsSystem, type: str = "galaxy", rng: Optional[random.Random] = None) -> None:
    """resetScenario(type), idx:744-764 (physics part)."""
    rnd = (rng or random).random
    physics.bodies = []
    physics.contactPairs.clear()
    if type == "galaxy":
        physics.addBody(Body(500, V3(0, 0, 0), V3(0, 0, 0), "metal", 1000))
        for _ in range(150):
            angle = rnd() * math.pi * 2
            dist = 30 + rnd() * 60
            velocity = math.sqrt(CONFIG.G * 500 / dist)
            pos = V3(math.cos(angle) * dist, (rnd() - 0.5) * 2, math.sin(angle) * dist)
            vel = V3(-math.sin(angle) * velocity, 0, math.cos(angle) * velocity)
            kind = "ice" if rnd() > 0.8 else "rock"
            physics.addBody(Body(rnd() * 2 + 0.5, pos, vel, kind))
    elif type == "collision":
        physics.addBody(Body(100, V3(-40, 0, 0), V3(1, 0, 0), "rock"))
        physics.addBody(Body(100, V3(40, 0, 10), V3(-1, 0, 0), "ice"))
