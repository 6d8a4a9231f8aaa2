"""One rank of tests/test_gpu_rank_mode.py (run under torchrun, one process per GPU): the nbody_create_rank model with CUDA-IPC
peer mappings.  Rank-mode results must equal a one-GPU context's bit for bit -- with one rank stalled between upload and step,
so that a missing cross-rank handshake around nbody_upload shows -- and the union of the ranks' contact lists
must be the oracle's complete list."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    failures = []

    def gather(obj):
        out = [None] * world
        dist.all_gather_object(out, obj)
        return out

    for precision, n in (("f32", 70002), ("f64", 6000)):
        dt = np.float32 if precision == "f32" else np.float64
        s = nb.scenes.two_galaxies(n - n % 2)
        pm, vel, temp = s.astype(dt)
        pm[:, 3] *= np.random.default_rng(3).uniform(0.5, 2.0, s.n).astype(dt)          # general masses
        pm2 = pm.copy()
        pm2[:, :3] += dt(0.125)
        eng = nb.create_rank_engine(s.n, precision)
        info0 = None
        # --- a step with rank (world-1) late after its upload, then a second upload with rank 0 late
        eng.upload(pm, vel, None, temp)
        if rank == world - 1:
            time.sleep(0.7)
        eng.step(s.dt, s.G, s.softening, 3)
        if rank == 0:
            time.sleep(0.7)
        eng.upload(pm2, vel, None, temp)          # peers may still be in step 3 / already past their upload
        if rank == world - 1:
            time.sleep(0.4)
        eng.step(s.dt, s.G, s.softening, 2)
        eng.force(s.G, s.softening)
        eng.force(s.G, s.softening)
        eng.step(s.dt, s.G, s.softening, 1)
        info0 = eng.info()
        out = eng.download(temp=True)
        own = {k: np.zeros_like(v) for k, v in out.items()}
        eng.download(temp=True, out=own, own_only=True)
        lo, hi = info0["i_begin"], info0["i_end"]
        # --- the same on one GPU, in this process
        with nb.Engine(s.n, precision=precision, rank=0, world=1, device=local) as one:
            one.upload(pm, vel, None, temp)
            one.step(s.dt, s.G, s.softening, 3)
            one.upload(pm2, vel, None, temp)
            one.step(s.dt, s.G, s.softening, 2)
            one.force(s.G, s.softening)
            one.force(s.G, s.softening)
            one.step(s.dt, s.G, s.softening, 1)
            ref = one.download(temp=True)
            if precision == "f32" and one.info()["f32_sym_last_pass"] != info0["f32_sym_last_pass"]:
                failures.append(f"{precision}: kernel choice differs between 1 GPU and rank mode")
        if not np.array_equal(out["pos_mass"], ref["pos_mass"]):
            failures.append(f"{precision} rank {rank}: pos_mass differs from the one-GPU run")
        for k in ("vel", "acc", "temp"):
            if not np.array_equal(out[k][lo:hi], ref[k][lo:hi]):
                failures.append(f"{precision} rank {rank}: {k} differs from the one-GPU run")
        for k in ("pos_mass", "vel", "acc", "temp"):
            if not np.array_equal(own[k][lo:hi], ref[k][lo:hi]) or np.any(own[k][:lo]) or np.any(own[k][hi:]):
                failures.append(f"{precision} rank {rank}: download_own {k}")
        if info0["exchange"] != 2:
            failures.append(f"{precision}: exchange mode {info0['exchange']}, expected peer stores (2)")
        if info0["upload_h2d_bytes"] >= s.n * 16 * (3 if precision == "f32" else 6) // 2 and world > 1:
            failures.append(f"{precision} rank {rank}: upload copied {info0['upload_h2d_bytes']} bytes, not just its slice")
        # --- contacts in rank mode: the union over ranks is the complete list
        if precision == "f32":
            from oracle import oracle
            radius = (0.6 * np.cbrt(np.abs(pm2[:, 3]) + 1e-3)).astype(dt)
            eng.upload(pm2, vel, None, temp)
            eng.set_radius(radius)
            eng.force(s.G, s.softening)             # tracks nearest neighbours: the flagged-candidate path
            pairs = eng.contacts(0.0)
            allp = gather(pairs)
            if rank == 0:
                got = np.unique(np.concatenate([p.reshape(-1, 2) for p in allp]), axis=0)
                want = oracle.overlap_pairs(pm2.astype(np.float64), radius.astype(np.float64), 0.0, mt=True)
                if len(want) == 0 or not np.array_equal(got, want):
                    failures.append(f"rank-mode contacts: union has {len(got)} pairs, the oracle {len(want)}")
        eng.close()
    allf = [f for fs in gather(failures) for f in fs]
    if rank == 0:
        print("RANK-MODE " + ("OK" if not allf else "FAIL:\n  " + "\n  ".join(allf)), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if allf else 0)


if __name__ == "__main__":
    main()
