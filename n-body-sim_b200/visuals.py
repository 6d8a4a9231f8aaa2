"""Host-side mirror of the arithmetic half of the reference's VisualSystem
(/root/reference/index.html:450-688), backed by the GPU engine.

`VisualSystem.update(bodies, dt)` in the page does two things per frame that are plain
arithmetic: it keeps each body's `history` (idx:540, 564-565) and rebuilds the ribbon
trail vertices from it (idx:567-592); the planet meshes are shaded by the GLSL of
idx:83-202.  Here the history ring and the ribbon builder run on the device
(`Engine.history_push`, `Engine.trail_ribbons`) and `render()` evaluates the shaders
headless (`Engine.render`).  three.js itself -- meshes, particles, flashes, bloom, tone
mapping -- is not mirrored (no graphics stack on this box).

Body identity: the page keys meshes and trails by `body.id` (idx:505-506), and a Body object
carries its `history` through list changes.  The array contract has no identity, so when the
list changes (merge, fracture, spawn) `update` tells the engine which row each surviving body
had before (`Engine.history_remap`); bodies it has not seen start with trailLength copies of
their position (idx:540).
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np

from . import physics as _ph

TYPE_INDEX = {"rock": 0, "metal": 1, "ice": 2}          # CONFIG.materials, idx:70-74


class VisualSystem:
    """update()/render() over the engine that already holds the bodies of `physics`."""

    def __init__(self, physics: "_ph.PhysicsSystem", camera_position=(0.0, 80.0, 150.0), camera_target=(0.0, 0.0, 0.0),
                 fov_deg: float = 45.0):
        self.physics = physics
        self.camera_position = tuple(float(x) for x in camera_position)   # camera.position.set(0, 80, 150), idx:705
        self.camera_target = tuple(float(x) for x in camera_target)       # OrbitControls target
        self.fov_deg = float(fov_deg)                                     # PerspectiveCamera(45, ...), idx:704
        self.seeds = {}                     # body.id -> material `seed` uniform, drawn at first sight (idx:492)
        self.trails: Optional[np.ndarray] = None   # float32 [n, trailLength, 6]: each trail mesh's `positions` (idx:524)
        self._depth = 0
        self._ring_ids: Optional[list] = None      # body.id per row of the device ring at the last push

    def _sync(self, bodies: List["_ph.Body"]):
        eng = self.physics.engine
        if eng.n != len(bodies):            # the list changed since the last integrate(): bring the device up to date
            self.physics._upload()
        dt = eng.dtype
        eng.set_radius(np.array([b.radius for b in bodies], dtype=dt))
        for b in bodies:
            if b.id not in self.seeds:
                self.seeds[b.id] = _ph.Math.random() * 100          # createPlanetMaterial, idx:492
        eng.set_materials(np.array([TYPE_INDEX[b.type] for b in bodies], dtype=np.int32),
                          np.array([self.seeds[b.id] for b in bodies], dtype=np.float32))
        return eng

    def update(self, bodies: List["_ph.Body"], dt: float) -> np.ndarray:
        """visuals.update(bodies, dt), idx:500-597 (history + ribbons).  Returns `self.trails`."""
        if not bodies:
            self.trails = np.zeros((0, _ph.CONFIG.trailLength, 6), np.float32)
            return self.trails
        eng = self._sync(bodies)
        if self._depth != _ph.CONFIG.trailLength:
            eng.history_enable(_ph.CONFIG.trailLength)
            self._depth = _ph.CONFIG.trailLength
            self._ring_ids = None
        ids = [b.id for b in bodies]
        if self._ring_ids is not None and ids != self._ring_ids:     # merge / fracture / spawn since the last frame
            row = {bid: k for k, bid in enumerate(self._ring_ids)}
            eng.history_remap([row.get(bid, -1) for bid in ids])
        eng.history_push()                                           # idx:540, 564-565
        self._ring_ids = ids
        self.trails = eng.trail_ribbons(self.camera_position)        # idx:567-592
        return self.trails

    def histories(self) -> np.ndarray:
        """Body.history of every body, [n, trailLength, 3], newest first."""
        return self.physics.engine.history()

    def garbageCollect(self, activeBodies: List["_ph.Body"]) -> None:   # idx:599-612
        alive = {b.id for b in activeBodies}
        self.seeds = {k: v for k, v in self.seeds.items() if k in alive}

    def render(self, width: int, height: int, ids: bool = False):
        """The planet shaders (idx:83-202) for the current bodies: float32 [h, w, 4] before tone mapping."""
        eng = self._sync(self.physics.bodies)
        return eng.render(width, height, camera=self.camera_position, target=self.camera_target, fov_deg=self.fov_deg, ids=ids)
