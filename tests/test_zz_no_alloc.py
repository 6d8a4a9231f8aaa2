"""nbody_accel_f32_static (include/nbody.h, csrc/static_scratch.cu): the entry point that allocates nothing -- its scratch is a
static device array -- against the existing entry points (nbody_force + nbody_download) on the same bodies, at the fp32
tolerance of tests/test_gpu_sym32.py, on calls in a row that share the scratch.  No counterpart in the reference.
Collected last on purpose."""
import ctypes
import os
import re

import numpy as np
import pytest

from .conftest import norm_rel_err

F32_ACC_TOL = 2e-5   # the gate of tests/test_gpu_sym32.py and tests/test_gpu_parity.py
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_the_static_entry_point_never_allocates():
    """Source check (runs without a GPU): no allocating call in the file, one static device array, its address looked up."""
    src = open(os.path.join(ROOT, "n-body-sim_b200", "csrc", "static_scratch.cu")).read()
    code = re.sub(r"//[^\n]*", "", src)
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocAsync", "new ", "malloc(", "std::vector"):
        assert banned not in code, banned
    assert re.search(r"static\s+__device__\s+__align__\(256\)\s+unsigned char g_scratch\[", code)
    assert "cudaGetSymbolAddress(&ptr, g_scratch)" in code
    # host code never hands the array's host-side name to a kernel: every launch that takes the scratch takes d.scratch
    for launch in re.findall(r"<<<[^;]*;", code):
        assert "g_scratch" not in launch, launch


def reference_acc(nb, pm, G, eps):
    with nb.Engine(len(pm), precision="f32") as e:
        e.upload(pm, np.zeros_like(pm))
        e.force(G, eps)
        return e.download()["acc"]


@pytest.mark.gpu
def test_static_scratch_matches_the_existing_entry_point_on_calls_in_a_row(nb, oracle):
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    cap = nb.STATIC_MAX_BODIES
    pos = torch.empty((cap, 4), dtype=torch.float32, device=dev)       # the caller's buffers: inputs and outputs only
    acc = torch.empty((cap, 4), dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    rng = np.random.default_rng(7)
    # two calls in a row (and more): sizes going down as well as up, so that stale scratch would show; equal and general masses
    cases = [("plummer", 12000, True), ("two_galaxies", 3002, False), ("uniform_cube", cap, True), ("plummer", 65, True),
             ("plummer", 12000, True), ("plummer", 1, False)]
    first = None
    for k, (scene, n, general) in enumerate(cases):
        s = nb.scenes.SCENES[scene](n)
        pm = s.pos_mass.astype(np.float32)
        if general:
            pm[:, 3] = (pm[:, 3].mean() * np.random.default_rng(n).uniform(0.2, 2.5, s.n)).astype(np.float32)
        pos[:s.n].copy_(torch.from_numpy(pm))
        acc.fill_(float("nan"))
        nb.accel_f32_static(pos.data_ptr(), acc.data_ptr(), s.n, s.G, s.softening, stream)
        got = acc[:s.n].cpu().numpy()
        want = reference_acc(nb, pm, s.G, s.softening)
        assert np.isfinite(got).all() and (got[:, 3] == 0).all()
        assert np.isnan(acc[s.n:].cpu().numpy()).all()                  # nothing written beyond n
        if s.n > 1:
            err = norm_rel_err(got, want)
            ref = oracle.acc_subset(pm.astype(np.float64), np.arange(0, s.n, max(1, s.n // 200)), s.G, s.softening)
            err_o = norm_rel_err(got[::max(1, s.n // 200)], ref)
            print(f"call {k}: {scene} n={s.n}: vs nbody_force {err:.2e}, vs the oracle {err_o:.2e}")
            assert err < F32_ACC_TOL and err_o < F32_ACC_TOL
        else:
            assert (got == 0).all()
        if k == 0:
            first = got.copy()
        if k == 4:
            assert np.array_equal(got, first)                           # integer accumulation: the same bits, call after call
    # back to back on one stream without a host synchronise in between: the second call's result stands
    s = nb.scenes.plummer(5000)
    pm = s.pos_mass.astype(np.float32)
    pos[:5000].copy_(torch.from_numpy(pm))
    acc2 = torch.empty((5000, 4), dtype=torch.float32, device=dev)
    nb.accel_f32_static(pos.data_ptr(), acc.data_ptr(), 5000, s.G, s.softening, stream)
    nb.accel_f32_static(pos.data_ptr(), acc2.data_ptr(), 4000, s.G, s.softening, stream)
    torch.cuda.synchronize()
    assert norm_rel_err(acc[:5000].cpu().numpy(), reference_acc(nb, pm, s.G, s.softening)) < F32_ACC_TOL
    assert norm_rel_err(acc2[:4000].cpu().numpy(), reference_acc(nb, pm[:4000], s.G, s.softening)) < F32_ACC_TOL
    # a non-finite coordinate poisons the pass, and the next call is clean again
    bad = pm.copy()
    bad[123, 0] = np.nan
    pos[:5000].copy_(torch.from_numpy(bad))
    nb.accel_f32_static(pos.data_ptr(), acc.data_ptr(), 5000, s.G, s.softening, stream)
    assert np.isnan(acc[:5000, :3].cpu().numpy()).all()
    pos[:5000].copy_(torch.from_numpy(pm))
    nb.accel_f32_static(pos.data_ptr(), acc.data_ptr(), 5000, s.G, s.softening, stream)
    assert np.isfinite(acc[:5000].cpu().numpy()).all()


@pytest.mark.gpu
def test_static_scratch_rejects_what_it_cannot_hold(nb):
    import torch
    torch.cuda.set_device(0)
    buf = torch.zeros((8, 4), dtype=torch.float32, device="cuda:0")
    L = nb._lib.lib()
    err = ctypes.create_string_buffer(256)
    args = (ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(buf.data_ptr()))
    rc = L.nbody_accel_f32_static(*args, nb.STATIC_MAX_BODIES + 1, 1.0, 0.1, None, err, len(err))
    assert rc == nb._lib.NBODY_EINVAL and b"NBODY_STATIC_MAX_BODIES=16384" in err.value
    rc = L.nbody_accel_f32_static(*args, 8, 1.0, 0.0, None, err, len(err))
    assert rc == nb._lib.NBODY_EINVAL and b"softening" in err.value
    rc = L.nbody_accel_f32_static(None, args[1], 8, 1.0, 0.1, None, None, 0)
    assert rc == nb._lib.NBODY_EINVAL
    assert L.nbody_accel_f32_static(None, None, 0, 1.0, 0.1, None, err, len(err)) == nb._lib.NBODY_OK   # nothing to do
    with pytest.raises(nb.NBodyError, match="exceeds"):
        nb.accel_f32_static(buf.data_ptr(), buf.data_ptr(), 10 ** 6, 1.0, 0.1)
