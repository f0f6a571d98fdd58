"""ctypes binding of libslamcuda.so (the C ABI of include/slamcuda.h).

The library is the product: if it is missing this module raises at import (no fallback).
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent.parent
LIB_PATH = Path(os.environ.get("SLAMCUDA_LIB", _PKG / "lib" / "libslamcuda.so"))

if not LIB_PATH.exists():
    raise ImportError(
        f"{LIB_PATH} not found: build it with `python slam.jl_b200/build.py` "
        "(libslamcuda has no CPU fallback)")

lib = C.CDLL(str(LIB_PATH), mode=C.RTLD_GLOBAL)

OK = 0
ERR_NOT_PD, ERR_BAD_INDEX, ERR_CAPACITY, ERR_CUDA, ERR_NCCL, ERR_ARG = -1, -2, -3, -4, -5, -6
MAX_OBS = 1024

c_double_p = C.POINTER(C.c_double)
c_int64_p = C.POINTER(C.c_int64)
c_int_p = C.POINTER(C.c_int)
vp = C.c_void_p


class SlamError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libslamcuda error {code}: {msg}")
        self.code = code


class NotPositiveDefinite(SlamError):
    pass


class BadIndex(SlamError, IndexError):
    pass


class CapacityError(SlamError):
    pass


def check(rc: int) -> int:
    if rc >= 0:
        return rc
    msg = lib.slam_last_error().decode()
    cls = {ERR_NOT_PD: NotPositiveDefinite, ERR_BAD_INDEX: BadIndex,
           ERR_CAPACITY: CapacityError}.get(rc, SlamError)
    raise cls(rc, msg)


def _sig(name, restype, *argtypes):
    f = getattr(lib, name)
    f.restype = restype
    f.argtypes = list(argtypes)
    return f


I, D, U64, U32, I64 = C.c_int, C.c_double, C.c_uint64, C.c_uint32, C.c_int64

_sig("slam_last_error", C.c_char_p)
_sig("slam_version", C.c_char_p)
_sig("slam_device_count", I)
_sig("slam_fp64_peak", I, I, c_double_p, c_double_p)
_sig("slam_create", I, I, I, C.POINTER(vp))
_sig("slam_destroy", I, vp)
_sig("slam_reserve", I, vp, I)
_sig("slam_capacity", I, vp)
_sig("slam_set_state", I, vp, c_double_p, c_double_p, I, I)
_sig("slam_get_state", I, vp, c_double_p, c_double_p, I)
_sig("slam_get_pose", I, vp, c_double_p)
_sig("slam_n", I, vp)
_sig("slam_get_cov_block", I, vp, I, I, I, I, c_double_p, I)
_sig("slam_get_cov_block_static", I, vp, I, I, I, I, c_double_p, I)
COV_BLOCK_STATIC_MAX = 65536    # SLAM_COV_BLOCK_STATIC_MAX of include/slamcuda.h
_sig("slam_downdate_kernel", I, vp)
_sig("slam_shard_downdate_kernel", I, vp)
_sig("slam_shard_debug_guard", I, vp, I, C.POINTER(C.c_longlong))
_sig("slam_debug_guard", I, vp, I, C.POINTER(C.c_longlong))
_sig("slam_compute_landmark_ellipses", I, vp, D, I, c_double_p, I)
_sig("slam_vehicle_ellipse", I, vp, D, I, c_double_p)
_sig("slam_feature_ellipses", I, vp, c_double_p)
_sig("slam_vehicle_ellipse_params", I, vp, c_double_p)
_sig("slam_laser_lines", I, vp, c_double_p, I, c_double_p)
_sig("slam_predict", I, vp, D, D, D, c_double_p, D)
_sig("slam_associate", I, vp, c_double_p, I, c_double_p, D, D, c_double_p, c_int64_p, c_int_p,
     c_double_p, c_int_p)
_sig("slam_set_pregate", I, vp, I)
_sig("slam_gate_stats", I, vp, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong), c_double_p)
_sig("slam_compute_association", I, vp, c_double_p, c_double_p, I64, c_double_p, c_double_p)
_sig("slam_predict_observation", I, vp, I64, c_double_p, c_double_p)
_sig("slam_update", I, vp, c_double_p, I, c_double_p, c_int64_p)
_sig("slam_add_features", I, vp, c_double_p, I, c_double_p)
_sig("slam_update_resident", I, vp, vp, I, c_double_p, vp)
_sig("slam_sync", I, vp)
_sig("slam_stream", vp, vp)
_sig("slam_set_downdate_variant", I, vp, I)
_sig("slam_launch_count", C.c_longlong, vp)
_sig("slam_profile", I, vp, I)
_sig("slam_profile_read", I, vp, c_double_p, c_int_p)
_sig("slam_synth_state", I, vp, I, D, U64)
_sig("slam_synth_observations", I, vp, I, D, U64, I, I, c_double_p, c_int64_p)

_sig("slam_ens_create", I, I, c_double_p, I, c_double_p, I, I, U64, U32, I, C.POINTER(vp))
_sig("slam_ens_destroy", I, vp)
_sig("slam_ens_steps", I, vp, c_int_p, c_int_p)
_sig("slam_ens_set_extra_landmarks", I, vp, I)
_sig("slam_ens_capacity", I, vp)
_sig("slam_ens_set_growth", I, vp, I)
_sig("slam_ens_set_max_shared_extra", I, vp, I)
_sig("slam_ens_tier_info", I, vp, c_int_p, c_int_p, I)
_sig("slam_ens_count_flops", I, vp, I, c_double_p)
_sig("slam_ens_set_threads", I, vp, I)
_sig("slam_ens_set_first_pass_extra", I, vp, I)
_sig("slam_ens_pass_info", I, vp, c_int_p, c_int_p)
_sig("slam_ens_steps_done", I, vp, C.POINTER(C.c_longlong), c_int_p)
_sig("slam_ens_run", I, vp)
_sig("slam_ens_sync", I, vp)
_sig("slam_ens_get_results", I, vp, c_double_p, c_int_p, c_int_p)
_sig("slam_ens_keep_tracks", I, vp, I)
_sig("slam_ens_get_track", I, vp, I, c_double_p)
_sig("slam_ens_get_state", I, vp, I, c_double_p, c_double_p, I)
_sig("slam_ens_set_noise", I, vp, c_double_p, c_double_p, I)
_sig("slam_ens_stream", vp, vp)
_sig("slam_ens_launch_count", C.c_longlong, vp)

_sig("slam_shard_unique_id", I, C.c_char_p)
_sig("slam_shard_create", I, I, I, I, C.c_char_p, I, C.POINTER(vp))
_sig("slam_shard_destroy", I, vp)
_sig("slam_shard_plan", I, I, I, I, c_int_p, c_int_p, I)
_sig("slam_shard_synth_state", I, vp, I, D, U64)
_sig("slam_shard_n", I, vp)
_sig("slam_shard_get_pose", I, vp, c_double_p)
_sig("slam_shard_get_x", I, vp, c_double_p)
_sig("slam_shard_get_cov_block", I, vp, I, I, I, I, c_double_p, I)
_sig("slam_shard_predict", I, vp, D, D, D, c_double_p, D)
_sig("slam_shard_associate", I, vp, c_double_p, I, c_double_p, D, D, c_double_p, c_int64_p, c_int_p,
     c_double_p, c_int_p)
_sig("slam_shard_update", I, vp, c_double_p, I, c_double_p, c_int64_p)
_sig("slam_shard_p2p", I, vp)
_sig("slam_shard_profile_phases", I, vp, c_double_p, c_int_p)
_sig("slam_shard_add_features", I, vp, c_double_p, I, c_double_p)
_sig("slam_shard_update_resident", I, vp, vp, I, c_double_p, vp)
_sig("slam_shard_sync", I, vp)
_sig("slam_shard_stream", vp, vp)
_sig("slam_shard_launch_count", C.c_longlong, vp)
_sig("slam_shard_profile", I, vp, I)
_sig("slam_shard_profile_read", I, vp, c_double_p, c_int_p)

# every symbol include/slamcuda.h declares (tests/test_abi.py checks the header against this)
SYMBOLS = sorted(n for n in dir(lib) if n.startswith("slam_"))


def dptr(a):
    """numpy float64 array -> double*"""
    return a.ctypes.data_as(c_double_p)


def iptr(a):
    return a.ctypes.data_as(c_int64_p)
