"""Host-side mirror of the reference's physics interface, backed by the GPU engine.

Same names, argument meaning and (absence of) error behaviour as the JS classes
in /root/reference/index.html: CONFIG (idx:52-75), Body (idx:208-231),
PhysicsSystem.addBody / integrate (idx:239-262).  `integrate(dt)` is the
gravity-only step contract (idx:247-262 with resolveCollisions, idx:256, left to
the host); `integrate(dt, collide=fn)` runs the split form so a host can keep
its own collision stage between force evaluation and kick-2 (idx:255-259).
`PhysicsSystem(collisions=True)` runs the reference's COMPLETE step: the O(N^2)
overlap search of resolveCollisions (idx:301-313) happens on the GPU
(`Engine.contacts`), and only the order-dependent response (impulse, positional
correction, merge, fracture, contact-time map; idx:314-389, 392-443) runs here,
on the candidate pairs, in the reference's visiting order and association.

All O(N) and O(N^2) arithmetic happens in libnbody.so on the GPU; this module
marshals Body objects to and from the typed arrays, like the JS shim in
INTEGRATION.md, and keeps the sequential host logic the reference keeps in JS.
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

    # members used by the collision stage (idx:309-369, 392-443), three@0.160.0 semantics
    def add(self, v):
        self.x += v.x; self.y += v.y; self.z += v.z
        return self

    def sub(self, v):
        self.x -= v.x; self.y -= v.y; self.z -= v.z
        return self

    def subVectors(self, a, b):
        self.x, self.y, self.z = a.x - b.x, a.y - b.y, a.z - b.z
        return self

    def addScaledVector(self, v, s):
        self.x += v.x * s; self.y += v.y * s; self.z += v.z * s
        return self

    def multiplyScalar(self, s):
        self.x *= s; self.y *= s; self.z *= s
        return self

    def divideScalar(self, s):
        return self.multiplyScalar(1 / s)

    def dot(self, v):
        return self.x * v.x + self.y * v.y + self.z * v.z

    def lengthSq(self):
        return self.x * self.x + self.y * self.y + self.z * self.z

    def length(self):
        return math.sqrt(self.x * self.x + self.y * self.y + self.z * self.z)

    def normalize(self):
        return self.divideScalar(self.length() or 1)

    def __iter__(self):
        return iter((self.x, self.y, self.z))

    def __repr__(self):
        return f"Vector3({self.x!r}, {self.y!r}, {self.z!r})"


def V3(x, y, z):  # idx:77
    return Vector3(x, y, z)


class SplitMix64:
    """Seedable stand-in for Math.random (the reference's is unseeded): sequential
    splitmix64 -> [0,1) with 53 bits.  The same stream drives the test oracle."""
    _M = (1 << 64) - 1

    def __init__(self, seed: int = 0x5EED0001):
        self.state = seed & self._M

    def random(self) -> float:
        self.state = (self.state + 0x9E3779B97F4A7C15) & self._M
        z = self.state
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & self._M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & self._M
        z ^= z >> 31
        return (z >> 11) * (1.0 / 9007199254740992.0)


class _Math:
    """`Math.random` as the classes below call it; rebind with seed_random()."""
    random = staticmethod(random.random)


Math = _Math()


def seed_random(seed: Optional[int]) -> None:
    """Math.random := SplitMix64(seed).random (None restores Python's unseeded generator)."""
    Math.random = random.random if seed is None else SplitMix64(seed).random


def _rng_state(restore=None):
    """Save / restore the Math.random stream (seeded SplitMix64 or Python's generator)."""
    owner = getattr(Math.random, "__self__", None)
    if isinstance(owner, SplitMix64):
        if restore is not None:
            owner.state = restore
        return owner.state
    if restore is not None:
        random.setstate(restore)
    return random.getstate()


class _NoVisuals:
    """The `visuals` global the collision code calls into (idx:358, 384, 408, 441)."""
    def triggerFlash(self, pos, energy): pass
    def garbageCollect(self, bodies): pass
    def spawnExplosion(self, pos, temp): pass


visuals = _NoVisuals()


class Body:
    """Body, idx:208-231."""

    def __init__(self, mass, pos, vel, type="rock", temp=0):
        self.id = Math.random()  # idx:210 spends one draw on a base-36 string; the draw itself is the id here
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
    """PhysicsSystem, idx:233-443.  The two O(N) loops, computeGravity and the O(N^2)
    overlap search of resolveCollisions run on the GPU; the sequential collision
    response stays on the host, as in the reference."""

    def __init__(self, precision: str = "f64", capacity: Optional[int] = None, n_gpus: int = 1,
                 collisions: bool = False, contact_margin: float = 0.5):
        self.bodies: List[Body] = []
        self.contactPairs = {}
        self.collisions = collisions          # False: the gravity-only contract of SURVEY.md 8(b)
        self.contact_margin = float(contact_margin)
        self.stats = {"overlap": 0, "merge": 0, "fracture": 0, "broad_phase_pairs": 0, "margin_retries": 0}
        self._cap = int(capacity or max(CONFIG.maxBodies, 1024))
        self._engine_args = dict(precision=precision, n_gpus=n_gpus)
        self._engine_obj: Optional[Engine] = None

    @property
    def _engine(self) -> Engine:
        # created on first use: the host-side list logic (addBody, resolveCollisions on a
        # given pair list) is usable, and testable, before any device work is requested
        if self._engine_obj is None:
            self._engine_obj = Engine(self._cap, **self._engine_args)
        return self._engine_obj

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

    def _upload(self):
        self._engine.upload(*self._pack())
        if self.collisions:
            self._engine.set_radius(np.array([b.radius for b in self.bodies], dtype=self._engine.dtype))

    def integrate(self, dt: float, collide: Optional[Callable[["PhysicsSystem", float], None]] = None,
                  nsteps: int = 1) -> None:
        """integrate(dt), idx:247-262.  Reads CONFIG.G / CONFIG.softening at call
        time like the reference (idx:268-269).  With `collide`, runs
        kick1+drift -> force -> collide(self, dt) on the host -> kick-2.  With
        `self.collisions`, the collision stage is the reference's own (idx:256)."""
        if not self.bodies:
            return
        if len(self.bodies) > self._cap:
            raise ValueError(f"{len(self.bodies)} bodies exceed the engine capacity {self._cap}")
        eng = self._engine
        eng.set_heat_decay(CONFIG.heatDecay)
        self._upload()
        if collide is None and not self.collisions:
            eng.step(dt, CONFIG.G, CONFIG.softening, nsteps)
            self._unpack(eng.download(temp=True))
            return
        for _ in range(nsteps):
            eng.kick1_drift(dt)
            eng.force(CONFIG.G, CONFIG.softening)
            if collide is not None:
                self._unpack(eng.download(temp=True))
                collide(self, dt)                       # the host's own resolveCollisions (idx:256)
                self._upload()                          # collisions may edit pos/vel and the body list
            elif self._collide_gpu(dt):
                self._upload()
            eng.kick2(dt, CONFIG.heatDecay)
        self._unpack(eng.download(temp=True))

    def computeGravity(self) -> None:
        """computeGravity(), idx:264-291: fills body.acc only."""
        if not self.bodies:
            return
        eng = self._engine
        self._upload()
        eng.force(CONFIG.G, CONFIG.softening)
        self._unpack(eng.download(pos_mass=False, vel=False), what=("acc",))

    # -- collision stage (idx:256, 293-443) ------------------------------------------
    def _collide_gpu(self, dt: float) -> bool:
        """resolveCollisions(dt) with the overlap search on the GPU.  Returns True when
        host state changed (the caller re-uploads).  The candidate list is every pair
        within `margin` of touching; the response may move bodies (idx:351-352), so the
        run is accepted only if no body pair outside the list can have come into contact
        (sum of the two largest displacements < margin), else it is redone from the saved
        state with a larger margin: the result equals the reference's full double loop."""
        eng = self._engine
        margin = self.contact_margin
        pairs = eng.contacts(margin)
        self.stats["broad_phase_pairs"] += len(pairs)
        if len(pairs) == 0:
            if self.contactPairs:
                self.contactPairs.clear()               # idx:376-380: nothing is in contact
            return False
        self._unpack(eng.download(temp=True))
        while True:
            saved = ([(b.pos.clone(), b.vel.clone(), b.temperature) for b in self.bodies],
                     dict(self.contactPairs), list(self.bodies), dict(self.stats),
                     _rng_state())
            moved = self.resolveCollisions(dt, pairs)
            top = sorted(moved.values(), reverse=True)[:2]
            if sum(top) < margin:
                return True
            # redo with a wider net
            for b, (p, v, t) in zip(saved[2], saved[0]):
                b.pos, b.vel, b.temperature = p, v, t
            self.contactPairs, self.bodies, self.stats = saved[1], saved[2], saved[3]
            _rng_state(saved[4])
            self.stats["margin_retries"] += 1
            margin = 2.0 * sum(top) + margin
            pairs = eng.contacts(margin)

    def resolveCollisions(self, dt: float, pairs=None) -> dict:
        """resolveCollisions(dt), idx:293-390.  `pairs`: candidate (i, j) rows, i < j, in
        the reference's visiting order (ascending i, then ascending j); None = every
        pair, i.e. the reference's own double loop (O(N^2) on the host: small N only).
        Returns {body index: distance moved by positional corrections}."""
        newBodies: List[Body] = []
        removedIndices = set()
        N = len(self.bodies)
        vec, relVel = Vector3(), Vector3()
        currentFrameContacts = set()
        moved: dict = {}

        if pairs is None:
            rows = ((i, range(i + 1, N)) for i in range(N))
        else:
            by_row: dict = {}
            for i, j in pairs:
                by_row.setdefault(int(i), []).append(int(j))
            rows = ((i, by_row[i]) for i in sorted(by_row))

        for i, js in rows:
            if i in removedIndices:                                   # idx:302
                continue
            for j in js:
                if j in removedIndices:                               # idx:304
                    continue
                A, B = self.bodies[i], self.bodies[j]
                vec.subVectors(B.pos, A.pos)                          # idx:309
                distSq = vec.lengthSq()
                minDist = A.radius + B.radius
                if not (distSq < minDist * minDist):                  # idx:313
                    continue
                self.stats["overlap"] += 1
                pairId = (A.id, B.id) if A.id < B.id else (B.id, A.id)   # [A.id,B.id].sort().join('-')
                currentFrameContacts.add(pairId)
                prevContactTime = self.contactPairs.get(pairId) or 0
                newContactTime = prevContactTime + dt
                self.contactPairs[pairId] = newContactTime

                dist = math.sqrt(distSq)
                normal = vec.clone().normalize()
                relVel.subVectors(B.vel, A.vel)
                velAlongNormal = relVel.dot(normal)
                if not (velAlongNormal < 0):                          # idx:327
                    continue
                restitution = 0.2
                jImpulse = -(1 + restitution) * velAlongNormal / (A.invMass + B.invMass)
                impulse = normal.clone().multiplyScalar(jImpulse)

                mu = (A.mass * B.mass) / (A.mass + B.mass)
                impactEnergy = 0.5 * mu * velAlongNormal * velAlongNormal
                A.temperature += (impactEnergy / A.mass) * 0.2
                B.temperature += (impactEnergy / B.mass) * 0.2
                Q = impactEnergy / (A.mass + B.mass)

                if newContactTime > CONFIG.mergeTime and Q < CONFIG.fractureThreshold * 2:   # idx:340
                    self.mergeBodies(A, B, newBodies)
                    removedIndices.add(i)
                    removedIndices.add(j)
                    self.contactPairs.pop(pairId, None)
                    continue

                isFracture = Q > CONFIG.fractureThreshold
                cs = (minDist - dist) / (A.invMass + B.invMass) * 0.8
                correction = normal.clone().multiplyScalar(cs)
                A.pos.sub(correction.clone().multiplyScalar(A.invMass))
                B.pos.add(correction.clone().multiplyScalar(B.invMass))
                moved[i] = moved.get(i, 0.0) + abs(cs) * A.invMass * 1.000001
                moved[j] = moved.get(j, 0.0) + abs(cs) * B.invMass * 1.000001

                if isFracture and (A.mass > CONFIG.minFragmentMass or B.mass > CONFIG.minFragmentMass):
                    self.fractureBodies(A, B, impactEnergy, newBodies)
                    removedIndices.add(i)
                    removedIndices.add(j)
                    visuals.triggerFlash(A.pos.clone().add(B.pos).multiplyScalar(0.5), impactEnergy)
                    self.contactPairs.pop(pairId, None)
                else:
                    A.vel.sub(impulse.clone().multiplyScalar(A.invMass))
                    B.vel.add(impulse.clone().multiplyScalar(B.invMass))
                    tangent = relVel.clone().sub(normal.clone().multiplyScalar(velAlongNormal)).normalize()
                    jt = -relVel.dot(tangent) * 0.5 / (A.invMass + B.invMass)
                    tangentImpulse = tangent.multiplyScalar(jt)
                    A.vel.sub(tangentImpulse.clone().multiplyScalar(A.invMass))
                    B.vel.add(tangentImpulse.clone().multiplyScalar(B.invMass))

        for key in list(self.contactPairs.keys()):                    # idx:376-380
            if key not in currentFrameContacts:
                del self.contactPairs[key]

        if removedIndices:                                            # idx:382-385
            self.bodies = [b for idx, b in enumerate(self.bodies) if idx not in removedIndices]
            visuals.garbageCollect(self.bodies)
        for nb in newBodies:                                          # idx:387-389
            self.addBody(nb)
        return moved

    def mergeBodies(self, A: Body, B: Body, newBodiesList: List[Body]) -> None:   # idx:392-409
        newMass = A.mass + B.mass
        momentumA = A.vel.clone().multiplyScalar(A.mass)
        momentumB = B.vel.clone().multiplyScalar(B.mass)
        newVel = momentumA.add(momentumB).multiplyScalar(1 / newMass)
        posA = A.pos.clone().multiplyScalar(A.mass)
        posB = B.pos.clone().multiplyScalar(B.mass)
        newPos = posA.add(posB).multiplyScalar(1 / newMass)
        newTemp = (A.temperature * A.mass + B.temperature * B.mass) / newMass
        type = A.type if A.mass > B.mass else B.type
        newBodiesList.append(Body(newMass, newPos, newVel, type, newTemp))
        visuals.triggerFlash(newPos, newMass * 0.5)
        self.stats["merge"] += 1

    def fractureBodies(self, A: Body, B: Body, energy: float, newBodiesList: List[Body]) -> None:   # idx:411-443
        totalMass = A.mass + B.mass
        centerOfMass = A.pos.clone().multiplyScalar(A.mass).add(B.pos.clone().multiplyScalar(B.mass)).multiplyScalar(1 / totalMass)
        totalMomentum = A.vel.clone().multiplyScalar(A.mass).add(B.vel.clone().multiplyScalar(B.mass))
        baseVel = totalMomentum.multiplyScalar(1 / totalMass)
        severity = min(energy / CONFIG.fractureThreshold, 5.0)
        numFragments = math.floor(3 + Math.random() * 3 * severity)
        remainingMass = totalMass
        for k in range(numFragments):
            if remainingMass < CONFIG.minFragmentMass:
                break
            fragMass = remainingMass * (0.3 + Math.random() * 0.4)
            if k == numFragments - 1:
                fragMass = remainingMass
            if fragMass < CONFIG.minFragmentMass:
                continue
            remainingMass -= fragMass
            x = Math.random() - 0.5                                   # argument order of idx:430
            y = Math.random() - 0.5
            z = Math.random() - 0.5
            scatterDir = Vector3(x, y, z).normalize()
            offset = scatterDir.clone().multiplyScalar((A.radius + B.radius) * 0.5 * Math.random())
            fragPos = centerOfMass.clone().add(offset)
            ejectSpeed = math.sqrt(energy / totalMass) * (0.5 + Math.random())
            fragVel = baseVel.clone().add(scatterDir.multiplyScalar(ejectSpeed))
            fragTemp = max(A.temperature, B.temperature) + (energy / totalMass) * 0.1
            type = A.type if A.mass > B.mass else B.type
            newBodiesList.append(Body(fragMass, fragPos, fragVel, type, fragTemp))
            visuals.spawnExplosion(fragPos, fragTemp)
        self.stats["fracture"] += 1


def resetScenario(physics: PhysicsSystem, type: str = "galaxy", rng: Optional[random.Random] = None) -> None:
    """resetScenario(type), idx:744-764 (physics part)."""
    rnd = rng.random if rng is not None else (lambda: Math.random())
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
