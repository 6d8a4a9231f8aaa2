"""B200-native all-pairs N-body engine behind the step contract of
Arecibo130117/N-body-sim (PhysicsSystem.integrate, /root/reference/index.html:247-291).

The directory name carries a hyphen; import it as `nbody_b200` (the shim at the
repo root registers this package under that name).
"""
from . import physics, scenes, snapshot  # noqa: F401
from ._lib import NBodyError, build  # noqa: F401
from .dist import create_rank_engine, padded_n, shard_range  # noqa: F401
from .engine import STATIC_MAX_BODIES, Engine, accel_f32_static, device_count, nccl_unique_id  # noqa: F401
from .physics import CONFIG, V3, Body, PhysicsSystem, Vector3, resetScenario  # noqa: F401
from .visuals import VisualSystem  # noqa: F401

__all__ = ["Engine", "PhysicsSystem", "Body", "Vector3", "V3", "CONFIG", "resetScenario", "VisualSystem", "scenes",
           "device_count", "nccl_unique_id", "accel_f32_static", "STATIC_MAX_BODIES", "create_rank_engine", "shard_range", "padded_n",
           "NBodyError", "build"]
