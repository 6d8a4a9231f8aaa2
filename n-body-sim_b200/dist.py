"""One-process-per-GPU plumbing (torchrun launch): torch.distributed carries the
128-byte NCCL id and the timing reductions, nothing else.  The position
exchange itself runs inside libnbody.so on its own communicator.

shard_range() restates the C side's partition (nbody_capi.cu: upload_t) so the
host logic can be tested on CPU with the gloo backend.
"""
from __future__ import annotations

import os
from typing import Tuple

import math

I_TILE_F32 = 1024  # the larger i-tile of the fp32 kernels: 128 threads x 8 bodies (pairs-once pass, equal masses), csrc/force_f32_sym.cuh
I_TILE_F64 = 256   # kThreads; 512 from 32,768 bodies up (two i-bodies per thread, nbody_capi.cu f64_bodies_per_thread)
J_UNIT_F32 = 512   # kJUnit: one ring stage of the fp32 kernel
J_UNIT_F64 = 64    # kJUnitF64


def i_tile_f64(n: int) -> int:
    """The fp64 kernel's i-tile for n bodies (nbody_info.i_tile reports the same)."""
    return 512 if n >= 32768 else I_TILE_F64


def padded_n(n: int, world: int, i_tile: int, j_unit: int = J_UNIT_F32) -> int:
    """nbody_capi.cu pad_unit(): whole i-tiles per shard and whole j units."""
    unit = math.lcm(i_tile * world, j_unit)
    return (n + unit - 1) // unit * unit


def shard_range(n: int, world: int, rank: int, i_tile: int, j_unit: int = J_UNIT_F32) -> Tuple[int, int]:
    """[begin, end) of the real bodies rank `rank` owns (contiguous
    i-ranges of the padded array, pads at the tail)."""
    n_local = padded_n(n, world, i_tile, j_unit) // world
    lo = min(rank * n_local, n)
    hi = min((rank + 1) * n_local, n)
    return lo, hi


def create_rank_engine(capacity: int, precision: str = "f32"):
    """Engine for this torchrun rank.  Requires torch.distributed to be initialised
    (any backend) when WORLD_SIZE > 1."""
    import torch.distributed as dist

    from .engine import Engine, nccl_unique_id

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    device = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1:
        return Engine(capacity, precision=precision, rank=0, world=1, device=device)
    box = [nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return Engine(capacity, precision=precision, rank=rank, world=world, device=device, nccl_id=box[0])


def sym32_units(n: int, bodies_per_thread: int = 8, run: int = 1, world: int = 1, warp_units: bool = False):
    """The work-unit table of the fp32 pairs-once pass (csrc/force_f32_sym.cuh; warp_units: its warp-unit form,
    csrc/force_f32_symw.cuh) and the ranks' shares of it, from the library's own host code (no device needed).
    Returns (units[k, 4] = i-tile, first position tile, tile count, diagonal flag; split[world + 1]); with warp_units
    the i-tile is a block of 32 x bodies_per_thread bodies and the tiles are groups of 64."""
    import ctypes

    import numpy as np

    from . import _lib
    L = _lib.lib()
    fn = L.nbody_selftest_symw_units if warp_units else L.nbody_selftest_sym32_units
    cnt = ctypes.c_size_t()
    split = np.zeros(world + 1, dtype=np.uint32)
    fn(n, bodies_per_thread, run, world, None, 0, ctypes.byref(cnt), None)
    units = np.zeros((cnt.value, 4), dtype=np.uint32)
    _lib.check(fn(n, bodies_per_thread, run, world, ctypes.c_void_p(units.ctypes.data), cnt.value, ctypes.byref(cnt),
                  ctypes.c_void_p(split.ctypes.data)))
    return units, split
