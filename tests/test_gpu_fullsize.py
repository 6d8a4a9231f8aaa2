"""Full-size and multi-GPU checks through size-independent properties: sampled
bodies against the oracle in O(k*N), Newton's third law (sum m*a = 0), energy /
momentum conservation over many steps, invariance to the j-chunk decomposition
and to the number of GPUs."""
import os

import numpy as np
import pytest

from .conftest import norm_rel_err

pytestmark = pytest.mark.gpu


def test_f32_1m_uniform_cube_sampled_vs_oracle(nb, oracle):
    """Config 4 at full size: N=1,048,576, G=1, eps=1."""
    s = nb.scenes.uniform_cube(1048576)
    pm32 = s.pos_mass.astype(np.float32)
    with nb.Engine(s.n, precision="f32") as e:
        e.upload(pm32, np.zeros((s.n, 4), np.float32))
        e.force(s.G, s.softening)
        acc = e.download()["acc"].astype(np.float64)
        info = e.info()
    idx = np.unique(np.concatenate([[0, 1, 511, 512, s.n - 1], np.random.default_rng(4).integers(0, s.n, 43)]))
    ref = oracle.acc_subset(pm32.astype(np.float64), idx, s.G, s.softening)
    err = norm_rel_err(acc[idx], ref)
    per_body = np.linalg.norm(acc[idx, :3] - ref[:, :3], axis=1) / np.linalg.norm(ref[:, :3], axis=1)
    print(f"1M cube f32: sampled norm-rel err {err:.3e}, worst per-body {per_body.max():.3e}, j_chunks={info['j_chunks']}")
    assert err < 2e-5 and per_body.max() < 1e-3     # a uniform cube nearly cancels: per-body is the loose one
    net = (pm32[:, 3:4].astype(np.float64) * acc[:, :3]).sum(axis=0)
    scale = np.abs(pm32[:, 3:4] * acc[:, :3]).sum()
    assert np.abs(net).max() < 1e-5 * scale          # sum m_i a_i = 0 (idx:282-288)


def test_f32_4m_cold_collapse_sampled_vs_oracle(nb, oracle):
    """Config 5's scene at full size on ONE GPU: N=4,194,304 cold collapse, G=1, eps=0.005 -- one force pass
    (1.76e13 interactions), sampled rows against the oracle's fp64 sums over all N bodies, and sum m_i a_i = 0."""
    s = nb.scenes.cold_collapse(4194304)
    pm32 = s.pos_mass.astype(np.float32)
    with nb.Engine(s.n, precision="f32") as e:
        e.upload(pm32, np.zeros((s.n, 4), np.float32))
        e.force(s.G, s.softening)
        acc = e.download()["acc"].astype(np.float64)
        info = e.info()
    assert info["f32_sym_last_pass"] == 1 and info["uniform_mass"] == 1
    idx = np.unique(np.concatenate([[0, 1, 511, 512, 1023, 1024, s.n - 1], np.random.default_rng(5).integers(0, s.n, 25)]))
    ref = oracle.acc_subset(pm32.astype(np.float64), idx, s.G, s.softening)
    err = norm_rel_err(acc[idx], ref)
    per_body = np.linalg.norm(acc[idx, :3] - ref[:, :3], axis=1) / np.linalg.norm(ref[:, :3], axis=1)
    print(f"4M cold collapse f32: sampled norm-rel err {err:.3e}, worst per-body {per_body.max():.3e}, "
          f"pairs-once units of {info['f32_symmetric']} tiles, {info['f32_sym_bodies_per_thread']} bodies/thread")
    assert err < 2e-5 and per_body.max() < 1e-4
    assert np.isfinite(acc).all()
    net = (pm32[:, 3:4].astype(np.float64) * acc[:, :3]).sum(axis=0)
    scale = np.abs(pm32[:, 3:4] * acc[:, :3]).sum()
    assert np.abs(net).max() < 1e-6 * scale          # sum m_i a_i = 0 (idx:282-288)


def test_f32_two_galaxies_131072_conservation(nb):
    """Config 3 at full size, shortened to 200 steps here (bench/energy script runs 10k):
    momentum stays put and energy drift stays small."""
    s = nb.scenes.two_galaxies(131072)
    pm, vel, _ = s.astype(np.float32)
    with nb.Engine(s.n, precision="f32") as e:
        e.upload(pm, vel)
        ke0, pe0, m0 = e.energy(s.G, s.softening)
        e.step(s.dt, s.G, s.softening, 200)
        ke1, pe1, m1 = e.energy(s.G, s.softening)
    drift = abs((ke1 + pe1) - (ke0 + pe0)) / abs(ke0 + pe0)
    print(f"two_galaxies 131072 f32: E0={ke0 + pe0:.6e} drift after 200 steps {drift:.3e}; dP={np.abs(m1 - m0).max():.3e}")
    assert drift < 1e-4
    assert np.abs(m1 - m0).max() < 1e-3 * 1100.0 * 1.0   # total mass 1100, bulk speed 1


def test_chunking_invariance(nb):
    """The j-range split only changes fp32 summation order."""
    s = nb.scenes.plummer(65536)
    pm, vel, _ = s.astype(np.float32)
    res = []
    os.environ["NBODY_F32_SYM"] = "0"     # (65,536 bodies would take the pairs-once pass, which has no j-chunks: tests/test_gpu_sym32.py)
    for chunk in ("0", "65536", "4096", "1024"):
        os.environ["NBODY_CHUNK_LEN"] = chunk
        try:
            with nb.Engine(s.n, precision="f32") as e:
                e.upload(pm, vel)
                e.force(s.G, s.softening)
                res.append((e.info()["j_chunks"], e.download()["acc"]))
        finally:
            os.environ.pop("NBODY_CHUNK_LEN", None)
    os.environ.pop("NBODY_F32_SYM", None)
    assert res[1][0] == 1 and res[3][0] == 64
    for jc, acc in res[1:]:
        assert norm_rel_err(acc, res[0][1]) < 2e-5, jc   # measured 5.6e-6 on B200 (fp32 reordering at N=65536)


def _multi_gpu(nb):
    n = nb.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    return n


@pytest.mark.parametrize("exchange", ["peer", "nccl"])
@pytest.mark.parametrize("precision", ["f32", "f64"])
def test_multi_gpu_bit_identical(nb, precision, exchange):
    """i-range sharding + position exchange (fused peer stores over NVLink, or
    ncclAllGather): same bits on 1 and P GPUs (SURVEY.md 8e 'Determinism')."""
    ng = _multi_gpu(nb)
    s = nb.scenes.two_galaxies(32768)
    dt = np.float32 if precision == "f32" else np.float64
    pm, vel, t = s.astype(dt)
    outs = []
    os.environ["NBODY_EXCHANGE"] = exchange
    # (fp32 and fp64 both take their pairs-once passes here -- n = 32,768, softening > 0 -- whose integer accumulators make the
    # result independent of the GPU count; the one-sided sweeps' fixed j-chunk order is covered by the softening = 0 run below)
    try:
        for g in sorted({1, 2, min(ng, 4), min(ng, 8)}):
            with nb.Engine(s.n, precision=precision, n_gpus=g) as e:
                e.upload(pm, vel, None, t)
                e.step(s.dt, s.G, s.softening, 5)
                e.step(s.dt, s.G, s.softening, 1)       # a second call: kick1_drift stores to the peers again
                assert e.info()["exchange"] == (0 if g == 1 else (2 if exchange == "peer" else 1))
                outs.append((g, e.download(temp=True), e.energy(s.G, s.softening)))
    finally:
        os.environ.pop("NBODY_EXCHANGE", None)
    for g, o, en in outs[1:]:
        for k in ("pos_mass", "vel", "acc", "temp"):
            assert np.array_equal(o[k], outs[0][1][k]), (g, k)
        assert abs(en[0] - outs[0][2][0]) < 1e-12 * abs(en[0])


@pytest.mark.parametrize("precision", ["f32", "f64"])
def test_multi_gpu_bit_identical_one_sided_sweeps(nb, precision):
    """softening = 0 selects the one-sided sweeps (no bound for the fixed-point scale): i-range shards with a j-chunk
    decomposition that is a function of N only -- the same bits on 1 and P GPUs as well."""
    ng = _multi_gpu(nb)
    s = nb.scenes.plummer(20000)
    dt = np.float32 if precision == "f32" else np.float64
    pm, vel, t = s.astype(dt)
    outs = []
    os.environ["NBODY_F64_SYM"] = "0"   # (fp64 on ONE GPU would take the array form of the pairs-once pass at softening 0: another order)
    try:
        for g in sorted({1, 2, min(ng, 4)}):
            with nb.Engine(s.n, precision=precision, n_gpus=g) as e:
                e.upload(pm, vel, None, t)
                e.step(s.dt, s.G, 0.0, 3)
                assert e.info()["f32_sym_last_pass"] == 0
                outs.append(e.download(temp=True))
    finally:
        os.environ.pop("NBODY_F64_SYM", None)
    for o in outs[1:]:
        for k in ("pos_mass", "vel", "acc", "temp"):
            assert np.array_equal(o[k], outs[0][k]), k


def test_f64_pairs_once_on_several_gpus(nb, oracle):
    """The fp64 pairs-once pass with exact hi/lo fixed-point accumulators (force_f64_sym.cuh, FIXED) on 1, 2 and 4 GPUs:
    the same bits, within the 1e-12 gate of the oracle, and the rows of the pair triangle shared by cost."""
    ng = _multi_gpu(nb)
    s = nb.scenes.plummer(24000)
    pm = s.pos_mass.copy()
    pm[:, 3] *= np.random.default_rng(5).uniform(0.5, 2.0, s.n)
    idx = np.arange(0, s.n, 997)
    ref = oracle.acc_subset(pm, idx, s.G, s.softening)
    outs = []
    for g in sorted({1, 2, min(ng, 4)}):
        with nb.Engine(s.n, precision="f64", n_gpus=g) as e:
            e.upload(pm, s.vel)
            assert e.info()["f64_symmetric"] > 0
            e.force(s.G, s.softening)
            acc = e.download()["acc"]
            e.step(s.dt, s.G, s.softening, 3)
            outs.append((acc, e.download()))
        assert norm_rel_err(acc[idx], ref) < 1e-12
    for acc, o in outs[1:]:
        assert np.array_equal(acc, outs[0][0])
        for k in ("pos_mass", "vel", "acc"):
            assert np.array_equal(o[k], outs[0][1][k]), k


def test_multi_gpu_ragged_and_split_api(nb, oracle):
    """Ragged N on 2 GPUs through the split entry points (kick1_drift / force / kick2)."""
    _multi_gpu(nb)
    s = nb.scenes.plummer(3001)
    p, v, a, _ = oracle.integrate(s.pos_mass, s.vel, np.zeros_like(s.vel), None, s.dt, s.G, s.softening, nsteps=2)
    with nb.Engine(s.n, precision="f64", n_gpus=2) as e:
        e.upload(s.pos_mass, s.vel)
        for _ in range(2):
            e.kick1_drift(s.dt)
            e.force(s.G, s.softening)
            e.kick2(s.dt, 0.0)
        o = e.download()
    assert norm_rel_err(o["pos_mass"], p) < 1e-12 and norm_rel_err(o["vel"], v) < 1e-12 and norm_rel_err(o["acc"], a) < 1e-12


def test_f32_1m_contacts_flagged_equals_full_scan_and_sampled_rows(nb):
    """Collision broad phase at config 4's full size (N=1,048,576, radii inflated until the
    cube has a few thousand contacts).  Size-independent properties: the list from the force
    sweep's nearest-neighbour flags equals the list from scanning every body; it is sorted
    (i ascending, then j) with i < j; and the rows of sampled bodies equal an O(k*N) restatement
    of the overlap test of resolveCollisions (idx:301-313) in float64 on the stored fp32 state."""
    s = nb.scenes.uniform_cube(1048576)
    pm, vel, _ = s.astype(np.float32)
    n = s.n
    rad = np.full(n, 0.1, np.float32)                             # ~18,000 contacts at one body per unit volume
    rad[:: 4099] = 0.25                                           # ragged radii, none dominant
    with nb.Engine(n, precision="f32") as e:
        e.upload(pm, vel)
        e.set_radius(rad)
        e.force(s.G, s.softening)
        flagged = e.contacts(0.0)
        st = e.contact_stats()
        assert st["used_flags"] and st["flagged"] < n // 4         # reach = r_i + largest ordinary radius
        assert st["candidates_only"]                                # 178k flagged: F^2 tests instead of F*N
        e.set_radius(rad)                                         # drops the flags
        full = e.contacts(0.0)
        assert not e.contact_stats()["used_flags"]
    assert len(full) > 1000
    assert np.array_equal(flagged, full)
    key = full[:, 0].astype(np.uint64) * np.uint64(n) + full[:, 1].astype(np.uint64)
    assert (full[:, 0] < full[:, 1]).all() and (np.diff(key.astype(np.int64)) > 0).all()
    # sampled rows: every body named in the first contacts plus random ones
    rng = np.random.default_rng(8)
    rows = np.unique(np.concatenate([full[:40, 0], full[-40:, 0], rng.integers(0, n, 120)]))
    p64, r64 = pm[:, :3].astype(np.float64), rad.astype(np.float64)
    for i in rows:
        d = p64[i + 1:] - p64[i]
        d2 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
        rs = r64[i] + r64[i + 1:]
        ref_j = np.nonzero(d2 < rs * rs)[0] + i + 1               # distSq < minDist * minDist, idx:309-313
        got_j = full[full[:, 0] == i, 1]
        assert np.array_equal(got_j, ref_j), i


@pytest.mark.parametrize("scene", ["plummer", "two_galaxies"])
def test_f64_32768_two_bodies_per_thread_sampled(nb, oracle, scene):
    """From 32,768 bodies up the fp64 kernel gives each thread two i-bodies; the Plummer sphere
    (masses 1/N) also takes its equal-mass variant, the two-galaxy scene (two heavy centres) the
    general one.  Sampled bodies against the oracle at the fp64 gate, and the one-body-per-thread
    shape must agree to rounding."""
    s = nb.scenes.SCENES[scene](32768)
    idx = np.unique(np.concatenate([[0, 1, 255, 256, 511, 512, 16384, s.n - 1], np.random.default_rng(9).integers(0, s.n, 56)]))
    ref = oracle.acc_subset(s.pos_mass, idx, s.G, s.softening)
    accs = {}
    for forced in ("0", "1"):
        os.environ["NBODY_F64_R"] = forced
        try:
            with nb.Engine(s.n, precision="f64") as e:
                e.upload(s.pos_mass, s.vel)
                info = e.info()
                assert info["i_tile"] == (512 if forced == "0" else 256)
                assert info["uniform_mass"] == (1 if scene == "plummer" else 0)
                e.force(s.G, s.softening)
                accs[forced] = e.download()["acc"]
        finally:
            del os.environ["NBODY_F64_R"]
    assert norm_rel_err(accs["0"][idx], ref) < 1e-12
    assert norm_rel_err(accs["0"], accs["1"]) < 1e-14
