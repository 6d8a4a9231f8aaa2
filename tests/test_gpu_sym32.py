"""The fp32 pairs-once pass (csrc/force_f32_sym.cuh): every pair evaluated once and applied to
both bodies, as the reference's i<j loop does (/root/reference/index.html:272-288), through exact
fixed-point accumulators.  Parity against the oracle at the fp32 tolerance, agreement with the
one-sided sweep, and the properties the fixed-point accumulation buys: the same bits run after
run, fused or split, for every unit size's schedule and on every GPU count."""
import os

import numpy as np
import pytest

from .conftest import norm_rel_err

pytestmark = pytest.mark.gpu

F32_ACC_TOL = 2e-5   # same gate as the one-sided sweep (tests/test_gpu_parity.py)


class env:
    def __init__(self, **kv):
        self.kv = {k: str(v) for k, v in kv.items()}

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.kv}
        os.environ.update(self.kv)

    def __exit__(self, *exc):
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def scene_arrays(nb, name, n, general):
    s = nb.scenes.SCENES[name](n)
    pm = s.pos_mass.astype(np.float32)
    if general:
        pm[:, 3] = (pm[:, 3].mean() * np.random.default_rng(n).uniform(0.2, 2.5, s.n)).astype(np.float32)
    return s, pm, s.vel.astype(np.float32)


def force(nb, pm, vel, G, eps, **kv):
    with env(**kv):
        with nb.Engine(len(pm), precision="f32") as e:
            e.upload(pm, vel)
            e.force(G, eps)
            return e.download()["acc"], e.info()


@pytest.mark.parametrize("R", [4, 8])
@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("scene,n", [("plummer", 1), ("plummer", 2), ("plummer", 513), ("two_galaxies", 1026), ("plummer", 3001),
                                     ("uniform_cube", 5000), ("cold_collapse", 40000)])
def test_sym32_force_vs_oracle(nb, oracle, scene, n, general, R):
    s, pm, vel = scene_arrays(nb, scene, n, general)
    ref = oracle.compute_gravity(pm.astype(np.float64), s.G, s.softening, mt=True)
    acc, info = force(nb, pm, vel, s.G, s.softening, NBODY_F32_SYM_MIN=1, NBODY_F32_SYM_R=R, NBODY_F32_SYMW=0)
    assert info["f32_symmetric"] > 0 and info["f32_sym_bodies_per_thread"] == R and info["f32_sym_last_pass"] == 1
    assert info["f32_sym_warp_units"] == 0
    err = norm_rel_err(acc, ref) if n > 1 else float(np.abs(acc[:, :3]).max())
    print(f"pairs-once R={R} {scene} n={s.n} {'general' if general else 'scene'} masses: acc norm-rel err {err:.3e}")
    assert err < F32_ACC_TOL
    one, info1 = force(nb, pm, vel, s.G, s.softening, NBODY_F32_SYM=0, NBODY_SMALL=0, NBODY_MEDIUM=0)
    assert info1["f32_symmetric"] == 0 and info1["f32_sym_last_pass"] == 0
    if n > 1:
        assert norm_rel_err(acc, one) < F32_ACC_TOL


@pytest.mark.parametrize("R", [4, 8])
def test_sym32_newtons_third_law(nb, R):
    """sum m_i a_i = 0 (idx:282-288): each pair's two contributions are the same fp32 products with opposite signs."""
    s, pm, vel = scene_arrays(nb, "two_galaxies", 50000, True)
    acc, _ = force(nb, pm, vel, s.G, s.softening, NBODY_F32_SYM_R=R, NBODY_F32_SYMW=0)
    ma = pm[:, 3:4].astype(np.float64) * acc[:, :3].astype(np.float64)
    assert np.abs(ma.sum(axis=0)).max() < 1e-6 * np.abs(ma).sum()


@pytest.mark.parametrize("general", [False, True])
def test_sym32_bits_do_not_depend_on_the_schedule(nb, general):
    """Dynamic unit scheduling + integer accumulation: two runs agree bit for bit; so do nsteps fused into one
    call, single steps, and the split entry points."""
    s, pm, vel = scene_arrays(nb, "plummer", 40000, general)
    s.softening = 0.02
    outs = []
    for how in ("fused", "fused", "single", "split"):
        with nb.Engine(s.n, precision="f32") as e:
            e.upload(pm, vel)
            assert e.info()["f32_symmetric"] > 0
            if how == "fused":
                e.step(s.dt, s.G, s.softening, 4)
            elif how == "single":
                for _ in range(4):
                    e.step(s.dt, s.G, s.softening, 1)
            else:
                for _ in range(4):
                    e.kick1_drift(s.dt)
                    e.force(s.G, s.softening)
                    e.kick2(s.dt, 0.998)
            outs.append(e.download())
    for o in outs[1:]:
        for k in ("pos_mass", "vel", "acc"):
            assert np.array_equal(o[k], outs[0][k]), k


def test_sym32_unit_size_only_reorders_fp32_sums(nb):
    s, pm, vel = scene_arrays(nb, "plummer", 65536, False)
    base, info = force(nb, pm, vel, s.G, s.softening, NBODY_F32_SYMW=0)
    assert info["f32_symmetric"] == 1 and info["f32_sym_warp_units"] == 0          # 128 position tiles: one tile per unit
    for run in (2, 7, 64):
        acc, info = force(nb, pm, vel, s.G, s.softening, NBODY_F32_SYM_RUN=run, NBODY_F32_SYMW=0)
        assert info["f32_symmetric"] == run
        assert norm_rel_err(acc, base) < F32_ACC_TOL


def test_sym32_uniform_mass_with_outliers(nb, oracle):
    """Config 3's shape: two mass-500 centres among equal disc bodies (the equal-mass sweep + per-body corrections)."""
    s = nb.scenes.two_galaxies(40000)
    pm, vel, _ = s.astype(np.float32)
    ref = oracle.compute_gravity(pm.astype(np.float64), s.G, s.softening, mt=True)
    acc, info = force(nb, pm, vel, s.G, s.softening)
    assert info["f32_symmetric"] > 0 and info["uniform_mass"] == 3
    assert norm_rel_err(acc, ref) < F32_ACC_TOL
    gen, info = force(nb, pm, vel, s.G, s.softening, NBODY_UNIFORM=0)
    assert info["uniform_mass"] == 0
    assert norm_rel_err(gen, ref) < F32_ACC_TOL


def test_sym32_non_finite_input_poisons_the_step(nb):
    """The reference turns every acceleration into NaN when one body's position is NaN (every pair with it is NaN,
    idx:277-288); a fixed-point accumulator cannot hold a NaN, so the pass carries a poison mark instead."""
    s, pm, vel = scene_arrays(nb, "plummer", 33000, False)
    pm[12345, 1] = np.nan
    acc, info = force(nb, pm, vel, s.G, s.softening)
    assert info["f32_symmetric"] > 0
    assert np.isnan(acc[:, :3]).all()
    pm[12345, 1] = 0.25
    acc, _ = force(nb, pm, vel, s.G, s.softening)      # ... and the next pass is clean again
    assert np.isfinite(acc).all()


def test_sym32_falls_back_when_it_cannot_apply(nb, oracle):
    s, pm, vel = scene_arrays(nb, "plummer", 33000, False)
    acc0, info = force(nb, pm, vel, s.G, 0.0)                        # no bound without softening: the one-sided sweep runs
    assert info["f32_sym_last_pass"] == 0
    ref = oracle.acc_subset(pm.astype(np.float64), np.arange(0, s.n, 1111), s.G, 0.0)
    assert norm_rel_err(acc0[::1111], ref) < 5e-5
    with nb.Engine(s.n, precision="f32") as e:                        # radii set: the sweep that tracks nearest neighbours
        e.upload(pm, vel)
        e.set_radius(np.full(s.n, 1e-4, np.float32))
        e.force(s.G, s.softening)
        a = e.download()["acc"]
        assert e.info()["f32_sym_last_pass"] == 0
    acc1, _ = force(nb, pm, vel, s.G, s.softening)
    assert norm_rel_err(a, acc1) < F32_ACC_TOL


# ---- the warp-unit form (csrc/force_f32_symw.cuh): mid-size systems, and general masses at every size -------------------------

WIDE = dict(NBODY_F32_SYMW_MIN=1, NBODY_F32_SYMW_MAX=1 << 30)


@pytest.mark.parametrize("R", [2, 4, 8])
@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("scene,n", [("plummer", 1), ("plummer", 2), ("plummer", 65), ("plummer", 513), ("two_galaxies", 1026), ("plummer", 3001),
                                     ("uniform_cube", 8192), ("cold_collapse", 20000)])
def test_warp_units_force_vs_oracle(nb, oracle, scene, n, general, R):
    s, pm, vel = scene_arrays(nb, scene, n, general)
    ref = oracle.compute_gravity(pm.astype(np.float64), s.G, s.softening, mt=True)
    acc, info = force(nb, pm, vel, s.G, s.softening, NBODY_F32_SYMW_R=R, **WIDE)
    assert info["f32_sym_warp_units"] == 1 and info["f32_sym_bodies_per_thread"] == R and info["f32_sym_last_pass"] == 1
    err = norm_rel_err(acc, ref) if n > 1 else float(np.abs(acc[:, :3]).max())
    print(f"warp units R={R} {scene} n={s.n} {'general' if general else 'scene'} masses: acc norm-rel err {err:.3e}")
    assert err < F32_ACC_TOL
    for run in (1, 5):     # the unit size only regroups fp32 partial sums
        other, info = force(nb, pm, vel, s.G, s.softening, NBODY_F32_SYMW_R=R, NBODY_F32_SYMW_RUN=run, **WIDE)
        assert info["f32_symmetric"] == run
        if n > 1:
            assert norm_rel_err(other, acc) < F32_ACC_TOL


def test_warp_units_where_they_are_the_default(nb, oracle):
    """[5,121, 393,216) bodies whatever the masses; general masses from 5,121 bodies up; equal masses above keep the CTA-wide kernel."""
    for n, general, fine, R in ((4096, True, None, 0), (6500, False, True, 2), (8192, False, True, 4), (20000, True, True, 4), (40000, False, True, 8),
                                (65536, False, True, 8), (65536, True, True, 8), (400000, True, True, 8), (400000, False, False, 8)):
        s, pm, vel = scene_arrays(nb, "plummer", n, general)
        acc, info = force(nb, pm, vel, s.G, s.softening)
        if fine is not None:
            assert info["f32_sym_last_pass"] == 1 and info["f32_sym_warp_units"] == int(fine), (n, info)
            assert info["f32_sym_bodies_per_thread"] == R, (n, info)
        pick = np.arange(0, s.n, max(1, s.n // 300))
        ref = oracle.acc_subset(pm.astype(np.float64), pick, s.G, s.softening)
        assert norm_rel_err(acc[pick], ref) < F32_ACC_TOL


@pytest.mark.parametrize("general", [False, True])
def test_warp_units_bits_do_not_depend_on_the_schedule(nb, general):
    s, pm, vel = scene_arrays(nb, "plummer", 12000, general)
    s.softening = 0.02
    outs = []
    for how in ("fused", "fused", "single", "split"):
        with nb.Engine(s.n, precision="f32") as e:
            e.upload(pm, vel)
            assert e.info()["f32_sym_warp_units"] == 1
            if how == "fused":
                e.step(s.dt, s.G, s.softening, 4)
            elif how == "single":
                for _ in range(4):
                    e.step(s.dt, s.G, s.softening, 1)
            else:
                for _ in range(4):
                    e.kick1_drift(s.dt)
                    e.force(s.G, s.softening)
                    e.kick2(s.dt, 0.998)
            outs.append(e.download())
    for o in outs[1:]:
        for k in ("pos_mass", "vel", "acc"):
            assert np.array_equal(o[k], outs[0][k]), k


def test_warp_units_non_finite_input_and_third_law(nb):
    s, pm, vel = scene_arrays(nb, "two_galaxies", 16002, True)
    acc, info = force(nb, pm, vel, s.G, s.softening)
    assert info["f32_sym_warp_units"] == 1
    ma = pm[:, 3:4].astype(np.float64) * acc[:, :3].astype(np.float64)
    assert np.abs(ma.sum(axis=0)).max() < 1e-6 * np.abs(ma).sum()
    with nb.Engine(s.n, precision="f32") as e:        # poisoned pass, clean pass, poisoned pass on ONE context: the two poison
        e.upload(pm, vel)                              # words of the memset-free single-GPU path take turns
        for k, bad in enumerate((True, False, False, True, False)):
            pm2 = pm.copy()
            if bad:
                pm2[777, 2] = np.inf
            e.upload(pm2, vel)
            for _ in range(k % 2 + 1):
                e.force(s.G, s.softening)
            a = e.download()["acc"]
            assert np.isnan(a[:, :3]).all() if bad else np.isfinite(a).all(), (k, bad)


# ---- all steps of a call in one cooperative launch (csrc/sym_coop.cuh) -----------------------------------------------------

def run_steps(nb, pm, vel, temp, s, calls, **kv):
    with env(**kv):
        with nb.Engine(len(pm), precision="f32") as e:
            e.upload(pm, vel, None, temp)
            for k in calls:
                e.step(s.dt, s.G, s.softening, k)
            out = e.download(temp=True)
            return out, e.info()


@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("scene,n", [("plummer", 6144), ("two_galaxies", 7002), ("plummer", 12000), ("uniform_cube", 16384), ("cold_collapse", 30001),
                                     ("plummer", 32768)])
def test_cooperative_steps_equal_the_launch_per_step_path(nb, oracle, scene, n, general):
    """The same unit table, integer accumulation and rounded integrator operations: every bit equal, call after call."""
    s, pm, vel = scene_arrays(nb, scene, n, general)
    temp = np.random.default_rng(n).uniform(0.0, 50.0, s.n).astype(np.float32)
    calls = (1, 3, 2)
    coop, info = run_steps(nb, pm, vel, temp, s, calls, NBODY_SYM_COOP_MAX=32768)   # (the default stops at 14,336 bodies)
    per_step, info0 = run_steps(nb, pm, vel, temp, s, calls, NBODY_SYM_COOP=0)
    assert info["f32_sym_warp_units"] == 1 and info["f32_sym_last_pass"] == 1
    if scene == "two_galaxies" and not general:      # equal masses with two heavy centres: corrected per body by the finish kernel,
        assert info["kernel_launches"] == info0["kernel_launches"]   # so the launch-per-step path runs on both sides
    else:                                              # 6 steps in 3 launches (+ upload kernels) against two or three launches per step
        assert info["kernel_launches"] < info0["kernel_launches"] - 8, (info["kernel_launches"], info0["kernel_launches"])
    for k in ("pos_mass", "vel", "acc", "temp"):
        assert np.array_equal(coop[k], per_step[k]), k
    if s.n > 12000:
        return          # (the CPU oracle's six steps take a minute there; the launch-per-step path has its own oracle gates)
    p, v, a, t = oracle.integrate(pm.astype(np.float64), vel.astype(np.float64), np.zeros((s.n, 4)), temp.astype(np.float64), s.dt, s.G, s.softening,
                                  nsteps=sum(calls))
    assert norm_rel_err(coop["acc"], a) < 2e-4 and norm_rel_err(coop["pos_mass"], p) < 1e-5 and norm_rel_err(coop["vel"], v) < 5e-5   # six fp32 steps


def test_cooperative_steps_default_range(nb):
    for n, coop in ((5000, False), (6144, True), (14000, True), (16384, False)):
        s, pm, vel = scene_arrays(nb, "plummer", n, True)
        _, info = run_steps(nb, pm, vel, None, s, (5,))
        assert (info["kernel_launches"] <= 4) == (coop or n <= 5120), (n, info["kernel_launches"])   # (up to 5,120 bodies: medium.cuh's one launch)


def test_cooperative_steps_poison_and_recovery(nb):
    s, pm, vel = scene_arrays(nb, "plummer", 9000, True)
    with nb.Engine(s.n, precision="f32") as e:
        bad = pm.copy()
        bad[4321, 0] = np.nan
        e.upload(bad, vel)
        e.step(s.dt, s.G, s.softening, 3)
        o = e.download()
        assert np.isnan(o["acc"][:, :3]).all() and np.isnan(o["pos_mass"][:, :3]).all()
        e.upload(pm, vel)                      # the poison word does not outlive the launch
        e.step(s.dt, s.G, s.softening, 2)
        e.force(s.G, s.softening)              # ... nor do the counter and the accumulators: the launch-per-step kernels follow cleanly
        o = e.download()
        assert np.isfinite(o["acc"]).all() and np.isfinite(o["pos_mass"]).all()
    ref, _ = run_steps(nb, pm, vel, None, s, (2,), NBODY_SYM_COOP=0)
    assert np.array_equal(o["pos_mass"], ref["pos_mass"]) and np.array_equal(o["vel"], ref["vel"])


def test_warp_units_multi_gpu_same_bits(nb):
    ng = nb.device_count()
    if ng < 2:
        pytest.skip("needs >= 2 GPUs")
    for n in (20002, 70002):
        s, pm, vel = scene_arrays(nb, "two_galaxies", n, True)
        outs = []
        for exchange in ("peer", "nccl"):
            for g in sorted({1, 2, min(ng, 4), min(ng, 8)}):
                with env(NBODY_EXCHANGE=exchange):
                    with nb.Engine(s.n, precision="f32", n_gpus=g) as e:
                        e.upload(pm, vel)
                        assert e.info()["f32_sym_warp_units"] == 1
                        e.step(s.dt, s.G, s.softening, 3)
                        e.force(s.G, s.softening)
                        e.step(s.dt, s.G, s.softening, 1)
                        outs.append((exchange, g, e.download()))
        for exchange, g, o in outs[1:]:
            for k in ("pos_mass", "vel", "acc"):
                assert np.array_equal(o[k], outs[0][2][k]), (n, exchange, g, k)


def test_sym32_multi_gpu_same_bits(nb):
    """Each GPU works through its share of the unit table for ALL bodies and the owners add the peers' accumulators:
    integer sums, so 1, 2, 4 and 8 GPUs give the same bits -- with peer mappings and with the NCCL all-reduce."""
    ng = nb.device_count()
    if ng < 2:
        pytest.skip("needs >= 2 GPUs")
    s, pm, vel = scene_arrays(nb, "two_galaxies", 70002, True)
    outs = []
    for exchange in ("peer", "nccl"):
        for g in sorted({1, 2, min(ng, 4), min(ng, 8)}):
            with env(NBODY_EXCHANGE=exchange, NBODY_F32_SYMW=0):     # (the CTA-wide form; the warp units have their own test above)
                with nb.Engine(s.n, precision="f32", n_gpus=g) as e:
                    e.upload(pm, vel)
                    assert e.info()["f32_symmetric"] > 0 and e.info()["f32_sym_warp_units"] == 0
                    e.step(s.dt, s.G, s.softening, 3)
                    e.force(s.G, s.softening)
                    e.force(s.G, s.softening)
                    e.step(s.dt, s.G, s.softening, 1)
                    outs.append((exchange, g, e.download()))
    for exchange, g, o in outs[1:]:
        for k in ("pos_mass", "vel", "acc"):
            assert np.array_equal(o[k], outs[0][2][k]), (exchange, g, k)
