"""Host side of the fp32 pairs-once pass on several GPUs (csrc/nbody_capi.cu make_sym32_units).

The reference visits every pair i<j once (/root/reference/index.html:272-288).  The device pass cuts that triangle
into work units -- an i-tile against a run of 512-body position tiles -- and the GPUs share the unit table by cost,
each working for ALL bodies; the owners then add the peers' 64-bit fixed-point accumulators.  Here, without a GPU:
 * the table the library builds covers every pair of tiles exactly once, for every kernel shape and GPU count,
   and the ranks' shares are contiguous and balanced;
 * a world-size-2 gloo run of that scheme (numpy fp32 arithmetic per unit, int64 partial sums, one all_reduce)
   gives the same bits as one rank working through the whole table, and the oracle's accelerations."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
POS_TILE = 512          # csrc/common.cuh kPosTile
THREADS = 128           # csrc/force_f32.cuh kF32Threads: an i-tile is THREADS * bodies_per_thread bodies


def tiles_of(n):
    return (n + POS_TILE - 1) // POS_TILE


@pytest.mark.parametrize("n", [1, 511, 512, 513, 5000, 49151, 65536, 200001])
@pytest.mark.parametrize("R", [4, 8])
@pytest.mark.parametrize("run", [1, 2, 7])
def test_unit_table_covers_the_pair_triangle_once(nb, n, R, run):
    units, split = nb.dist.sym32_units(n, R, run, 1)
    NS, spt = tiles_of(n), THREADS * R // POS_TILE
    T = (n + THREADS * R - 1) // (THREADS * R)
    seen = np.zeros((T, NS), dtype=np.int32)
    diag = np.zeros(T, dtype=np.int32)
    for I, first, count, d in units.tolist():
        assert 1 <= count and first + count <= NS
        if d:
            # the diagonal unit: the i-tile against its own position tiles, pairs taken once inside the kernel
            assert first == I * spt and count == min((I + 1) * spt, NS) - first
            diag[I] += 1
        else:
            assert count <= run and first >= (I + 1) * spt
        seen[I, first:first + count] += 1
    assert (diag == 1).all()
    for I in range(T):
        assert (seen[I, : I * spt] == 0).all()        # nothing below the diagonal: the j side gets its share from above
        assert (seen[I, I * spt:] == 1).all()         # every tile from the diagonal up, once
    assert split.tolist() == [0, len(units)]
    # table order = i-tile order, tiles ascending inside an i-tile (the shares below rely on it)
    key = units[:, 0].astype(np.int64) * (NS + 1) + units[:, 1]
    assert (np.diff(key) > 0).all()


GROUP = 64              # csrc/force_f32_symw.cuh kSymwGroup


@pytest.mark.parametrize("n", [1, 63, 64, 65, 129, 1000, 5000, 8192, 40000])
@pytest.mark.parametrize("R", [2, 4, 8])
@pytest.mark.parametrize("run", [1, 3, 16])
def test_warp_unit_table_covers_the_pair_triangle_once(nb, n, R, run):
    """The warp-unit form (csrc/force_f32_symw.cuh): i-blocks of R/2 groups of 64 bodies against runs of groups."""
    units, split = nb.dist.sym32_units(n, R, run, 1, warp_units=True)
    NG, h = (n + GROUP - 1) // GROUP, R // 2
    NB = (NG + h - 1) // h
    seen = np.zeros((NB, NG), dtype=np.int32)
    diag = np.zeros(NB, dtype=np.int32)
    for b, first, count, d in units.tolist():
        assert 1 <= count and first + count <= NG
        if d:
            assert first == b * h and count == min((b + 1) * h, NG) - first
            diag[b] += 1
        else:
            assert count <= run and first >= (b + 1) * h
        seen[b, first:first + count] += 1
    assert (diag == 1).all()
    for b in range(NB):
        assert (seen[b, : b * h] == 0).all() and (seen[b, b * h:] == 1).all()
    assert split.tolist() == [0, len(units)]
    for world in (2, 3, 8):
        again, shares = nb.dist.sym32_units(n, R, run, world, warp_units=True)
        assert np.array_equal(again, units) and shares[0] == 0 and shares[-1] == len(units)
        assert (np.diff(shares.astype(np.int64)) >= 0).all()
        visits = units[:, 2].astype(np.int64)
        share = np.array([visits[shares[r]:shares[r + 1]].sum() for r in range(world)])
        assert np.abs(share - visits.sum() / world).max() <= 2 * int(visits.max()) + world


@pytest.mark.parametrize("n,R,run", [(5000, 8, 1), (65536, 8, 1), (70002, 4, 1), (1048576, 8, 2), (4194304, 8, 8)])
@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_ranks_share_the_table_by_cost(nb, n, R, run, world):
    one, _ = nb.dist.sym32_units(n, R, run, 1)
    units, split = nb.dist.sym32_units(n, R, run, world)
    assert np.array_equal(units, one)                 # the table is a function of n and the kernel shape only
    assert split[0] == 0 and split[-1] == len(units) and (np.diff(split.astype(np.int64)) >= 0).all()
    visits = units[:, 2].astype(np.int64)
    share = np.array([visits[split[r]:split[r + 1]].sum() for r in range(world)])
    assert share.sum() == visits.sum()
    biggest = int(visits.max())
    # each cut lands within one unit of total/world
    assert np.abs(share - visits.sum() / world).max() <= 2 * biggest + world


def unit_bodies(unit, n, R):
    I, first, count, d = (int(x) for x in unit)
    itile = THREADS * R
    i0, i1 = I * itile, min((I + 1) * itile, n)
    j0, j1 = first * POS_TILE, min((first + count) * POS_TILE, n)
    return i0, i1, j0, j1, bool(d)


def accumulate_units(pm, units, R, eps, scale):
    """One rank's pass over `units` in fp32 (numpy), both sides of every pair, into int64 fixed point: the arithmetic of
    sym_stage / sym_red (csrc/force_f32_sym.cuh) with the unit as the granule whose partial sums are rounded once."""
    n = len(pm)
    fix = np.zeros((n, 3), dtype=np.int64)
    e2 = np.float32(eps) * np.float32(eps)
    for u in units:
        i0, i1, j0, j1, d = unit_bodies(u, n, R)
        pi, pj = pm[i0:i1], pm[j0:j1]
        dx = pj[None, :, :3] - pi[:, None, :3]                                   # idx:274-276
        r2 = (dx * dx).sum(axis=2, dtype=np.float32) + e2                        # idx:277
        inv = (np.float32(1) / np.sqrt(r2)).astype(np.float32)
        inv3 = inv * inv * inv                                                   # idx:279-280
        if d:
            gi = np.arange(i0, i1)[:, None]
            gj = np.arange(j0, j1)[None, :]
            inv3 = np.where(gi < gj, inv3, np.float32(0))                        # each pair of the tile once
        f = inv3[:, :, None] * dx
        to_i = (f * pj[None, :, 3:4]).sum(axis=1, dtype=np.float32)              # idx:282-284
        to_j = (f * pi[:, None, 3:4]).sum(axis=0, dtype=np.float32)              # idx:286-288 (opposite sign)
        fix[i0:i1] += np.rint(to_i.astype(np.float64) * scale).astype(np.int64)
        fix[j0:j1] -= np.rint(to_j.astype(np.float64) * scale).astype(np.int64)
    return fix


def fixed_point_scale(pm, eps):
    """The bound the library uses for the accumulator scale: no sum can pass sum|m| * 2 / (3 sqrt(3) eps^2)."""
    bound = float(np.abs(pm[:, 3].astype(np.float64)).sum()) * 2.0 / (3.0 * np.sqrt(3.0) * eps * eps)
    return 2.0 ** (61 - int(np.ceil(np.log2(bound))))


def _worker(rank, world, port, n, R, run, q):
    sys.path.insert(0, ROOT)
    import nbody_b200 as nb
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s = nb.scenes.two_galaxies(n)
    pm = s.pos_mass.astype(np.float32)
    units, split = nb.dist.sym32_units(n, R, run, world)
    scale = fixed_point_scale(pm, s.softening)
    mine = accumulate_units(pm, units[split[rank]:split[rank + 1]], R, s.softening, scale)
    total = torch.from_numpy(mine.copy())
    dist.all_reduce(total, op=dist.ReduceOp.SUM)                                # ncclAllReduce(int64) on the device
    q.put((rank, int(split[rank + 1] - split[rank]), total.numpy().copy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n,R,run", [(3000, 4, 1), (3002, 8, 2)])
def test_two_ranks_sum_to_the_same_bits_as_one(nb, oracle, n, R, run):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, R, run, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    s = nb.scenes.two_galaxies(n)
    pm = s.pos_mass.astype(np.float32)
    units, _ = nb.dist.sym32_units(n, R, run, 1)
    scale = fixed_point_scale(pm, s.softening)
    whole = accumulate_units(pm, units, R, s.softening, scale)
    assert sum(g[1] for g in got) == len(units) and all(g[1] > 0 for g in got)
    for _, _, total in got:
        assert np.array_equal(total, whole)                                      # integer sums: any split, the same bits
    # ... and a different table (other unit size) only reorders fp32 sums
    acc = s.G * whole.astype(np.float64) / scale
    ref = oracle.compute_gravity(pm.astype(np.float64), s.G, s.softening)[:, :3]
    err = np.linalg.norm(acc - ref) / np.linalg.norm(ref)
    assert err < 2e-5, err                                                       # the fp32 gate of tests/test_gpu_sym32.py
    # Newton's third law in the integers: sum_i m_i a_i = 0 up to the rounding of each unit's partial sums
    assert np.abs((pm[:, 3:4].astype(np.float64) * acc).sum(axis=0)).max() < 1e-5 * np.abs(pm[:, 3:4] * acc).sum()
