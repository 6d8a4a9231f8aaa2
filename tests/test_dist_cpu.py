"""World-size-2 gloo test of the multi-GPU host logic: contiguous
i-range shards, every rank holding all positions, one all-gather of the freshly
drifted slices per step.  Compute here is the CPU oracle (tests may use it);
the point is that the sharded schedule reproduces the unsharded step exactly."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n, nsteps, q):
    sys.path.insert(0, ROOT)
    import nbody_b200 as nb
    from oracle import oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s = nb.scenes.two_galaxies(n)
    i_tile = 8
    lo, hi = nb.shard_range(n, world, rank, i_tile)
    n_pad = nb.padded_n(n, world, i_tile)
    n_loc = n_pad // world
    pos = np.zeros((n_pad, 4))
    pos[:n] = s.pos_mass
    vel, acc = s.vel[lo:hi].copy(), np.zeros((hi - lo, 4))
    for _ in range(nsteps):
        # kick-1 + drift on the own range, then all-gather the slices (schedule of nbody_step)
        own, vel = oracle.kick1_drift(pos[lo:hi], vel, acc, s.dt)
        mine = np.zeros((n_loc, 4))
        mine[: hi - lo] = own
        parts = [torch.zeros(n_loc, 4, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(parts, torch.from_numpy(mine))
        pos = torch.cat(parts).numpy().copy()
        pos[n:] = 0
        acc = oracle.acc_subset(pos[:n], np.arange(lo, hi), s.G, s.softening)
        vel, _ = oracle.kick2(vel, acc, None, s.dt)
    q.put((rank, lo, hi, pos[:n].copy(), vel, acc))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [250, 512])
def test_sharded_schedule_equals_unsharded(n, oracle, nb):
    world, nsteps = 2, 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, nsteps, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    s = nb.scenes.two_galaxies(n)
    p, v, a, _ = oracle.integrate(s.pos_mass, s.vel, np.zeros_like(s.vel), None, s.dt, s.G, s.softening, nsteps=nsteps)
    covered = 0
    for rank, lo, hi, pos, vel, acc in got:
        assert np.array_equal(pos, p)
        assert np.array_equal(vel, v[lo:hi]) and np.array_equal(acc, a[lo:hi])
        covered += hi - lo
    assert covered == n


def test_shard_ranges_partition(nb):
    for n in (1, 151, 512, 1000, 1048576):
        for world in (1, 2, 3, 4, 8):
            for tile in (256, 512):
                r = [nb.shard_range(n, world, k, tile) for k in range(world)]
                assert r[0][0] == 0 and r[-1][1] == n
                assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
                assert nb.padded_n(n, world, tile) % (world * tile) == 0
