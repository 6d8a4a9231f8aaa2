"""ctypes binding of libnbody.so (include/nbody.h).

There is no CPU fallback: if the shared library is missing, or a call fails,
this module raises.  The library is built in-tree by `build()` (nvcc, sm_100a).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# NBODY_LIB: load another build of the same library (kernel-variant experiments, scripts/)
LIB_PATH = os.environ.get("NBODY_LIB") or os.path.join(_HERE, "libnbody.so")
CSRC = os.path.join(_HERE, "csrc")

NBODY_F32, NBODY_F64 = 0, 1
NBODY_OK, NBODY_EINVAL, NBODY_ECUDA, NBODY_ENCCL, NBODY_ENOMEM, NBODY_EIO = 0, -1, -2, -3, -4, -5
NCCL_ID_BYTES = 128

# every symbol include/nbody.h declares (tests check the library exports them all)
SYMBOLS = (
    "nbody_create", "nbody_create_rank", "nbody_nccl_unique_id", "nbody_destroy",
    "nbody_upload", "nbody_download", "nbody_download_own", "nbody_step", "nbody_step_host", "nbody_kick1_drift", "nbody_force",
    "nbody_kick2", "nbody_set_heat_decay", "nbody_energy", "nbody_sync", "nbody_last_error",
    "nbody_strerror", "nbody_version", "nbody_device_count", "nbody_set_stream",
    "nbody_enable_timing", "nbody_read_timing", "nbody_get_info",
    "nbody_set_radius", "nbody_contacts", "nbody_contact_stats",
    "nbody_history_enable", "nbody_history_push", "nbody_history_remap", "nbody_history_download", "nbody_trail_ribbons",
    "nbody_snapshot_save", "nbody_snapshot_load", "nbody_set_materials", "nbody_render",
    "nbody_world_enable", "nbody_world_config", "nbody_world_set_bodies", "nbody_world_step", "nbody_world_sync", "nbody_world_frame", "nbody_world_contacts",
    "nbody_selftest_pow_third", "nbody_selftest_sym32_units", "nbody_selftest_symw_units", "nbody_accel_f32_static",
)


class NBodyError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"libnbody status {status}: {message}")
        self.status = status


class Info(ctypes.Structure):
    _fields_ = [
        ("n", ctypes.c_size_t), ("n_padded", ctypes.c_size_t), ("i_begin", ctypes.c_size_t),
        ("i_end", ctypes.c_size_t), ("n_gpus", ctypes.c_int), ("world", ctypes.c_int),
        ("rank", ctypes.c_int), ("sm_count", ctypes.c_int), ("j_chunks", ctypes.c_int),
        ("i_tile", ctypes.c_int), ("kernel_launches", ctypes.c_longlong), ("exchange", ctypes.c_int),
        ("uniform_mass", ctypes.c_int), ("f64_symmetric", ctypes.c_int),
        ("f32_symmetric", ctypes.c_int), ("f32_sym_bodies_per_thread", ctypes.c_int),
        ("f32_sym_last_pass", ctypes.c_int), ("upload_h2d_bytes", ctypes.c_longlong),
        ("f32_sym_warp_units", ctypes.c_int),
    ]


def build(force: bool = False) -> str:
    """Compile libnbody.so for sm_100a with nvcc."""
    args = ["make", "-C", CSRC, "-s"]
    if force:
        args.append("-B")
    subprocess.run(args, check=True)
    return LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    """Load libnbody.so.  Raises if it has not been built: no fallback exists."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA engine has not been built "
            "(run `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
    L = ctypes.CDLL(LIB_PATH)
    vp, sz, d, i = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_double, ctypes.c_int
    pd = ctypes.POINTER(ctypes.c_double)
    L.nbody_create.argtypes = [ctypes.POINTER(vp), i, i, ctypes.POINTER(i), sz]
    L.nbody_create_rank.argtypes = [ctypes.POINTER(vp), i, i, i, i, vp, sz]
    L.nbody_nccl_unique_id.argtypes = [vp]
    L.nbody_destroy.argtypes = [vp]
    L.nbody_destroy.restype = None
    L.nbody_upload.argtypes = [vp, sz, vp, vp, vp, vp]
    L.nbody_download.argtypes = [vp, vp, vp, vp, vp]
    L.nbody_download_own.argtypes = [vp, vp, vp, vp, vp]
    L.nbody_step.argtypes = [vp, d, d, d, i]
    L.nbody_step_host.argtypes = [vp, sz, vp, vp, vp, vp, d, d, d, i]
    L.nbody_kick1_drift.argtypes = [vp, d]
    L.nbody_force.argtypes = [vp, d, d]
    L.nbody_kick2.argtypes = [vp, d, d]
    L.nbody_set_heat_decay.argtypes = [vp, d]
    L.nbody_energy.argtypes = [vp, d, d, pd, pd, pd]
    L.nbody_sync.argtypes = [vp]
    L.nbody_last_error.argtypes = [vp]
    L.nbody_last_error.restype = ctypes.c_char_p
    L.nbody_strerror.argtypes = [i]
    L.nbody_strerror.restype = ctypes.c_char_p
    L.nbody_version.restype = ctypes.c_char_p
    L.nbody_device_count.restype = i
    L.nbody_set_stream.argtypes = [vp, vp]
    L.nbody_enable_timing.argtypes = [vp, i]
    L.nbody_read_timing.argtypes = [vp, ctypes.POINTER(ctypes.c_longlong), pd]
    L.nbody_get_info.argtypes = [vp, ctypes.POINTER(Info)]
    L.nbody_set_radius.argtypes = [vp, vp]
    L.nbody_contacts.argtypes = [vp, d, vp, sz, ctypes.POINTER(sz)]
    L.nbody_contact_stats.argtypes = [vp, ctypes.POINTER(ctypes.c_longlong)]
    L.nbody_history_enable.argtypes = [vp, i]
    L.nbody_history_push.argtypes = [vp]
    L.nbody_history_remap.argtypes = [vp, vp]
    L.nbody_history_download.argtypes = [vp, vp]
    L.nbody_trail_ribbons.argtypes = [vp, pd, vp]
    L.nbody_set_materials.argtypes = [vp, vp, vp]
    L.nbody_render.argtypes = [vp, i, i, pd, pd, d, vp, vp]
    ull = ctypes.c_ulonglong
    L.nbody_snapshot_save.argtypes = [vp, ctypes.c_char_p, ull, d, d, d]
    L.nbody_snapshot_load.argtypes = [vp, ctypes.c_char_p, ctypes.POINTER(ull), pd, pd, pd]
    pi, pll = ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_longlong)
    L.nbody_world_enable.argtypes = [vp, ull, sz]
    L.nbody_world_config.argtypes = [vp, d, d, d]
    L.nbody_world_set_bodies.argtypes = [vp, vp, vp, vp]
    L.nbody_world_step.argtypes = [vp, d, d, d, i, i]
    L.nbody_world_sync.argtypes = [vp, ctypes.POINTER(sz), vp, vp, vp, pll, ctypes.POINTER(sz), ctypes.POINTER(ull)]
    L.nbody_world_frame.argtypes = [vp, d, d, d, i, i, sz, ctypes.POINTER(sz), vp, vp, vp, vp, vp, vp, vp, pll, ctypes.POINTER(sz),
                                    ctypes.POINTER(ull)]
    L.nbody_world_contacts.argtypes = [vp, vp, sz, ctypes.POINTER(sz)]
    L.nbody_selftest_pow_third.argtypes = [vp, vp, vp, sz]
    L.nbody_selftest_sym32_units.argtypes = [sz, i, i, i, vp, sz, ctypes.POINTER(sz), vp]
    L.nbody_selftest_symw_units.argtypes = [sz, i, i, i, vp, sz, ctypes.POINTER(sz), vp]
    L.nbody_accel_f32_static.argtypes = [vp, vp, sz, d, d, vp, ctypes.c_char_p, sz]
    _lib = L
    return L


def check(status: int, ctx=None) -> None:
    if status == NBODY_OK:
        return
    L = lib()
    msg = L.nbody_strerror(status).decode()
    if ctx:
        detail = L.nbody_last_error(ctx).decode()
        if detail:
            msg = f"{msg}: {detail}"
    raise NBodyError(status, msg)
