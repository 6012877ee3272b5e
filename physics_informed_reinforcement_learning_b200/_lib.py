"""ctypes binding of include/pirl_rcwa.h.  Fails loudly when the CUDA library is missing: no CPU fallback."""
from __future__ import annotations

import ctypes as C
import os

from .build import LIB_PATH

SYMBOLS = [
    "pirl_abi_version", "pirl_last_error", "pirl_rcwa_create", "pirl_rcwa_destroy", "pirl_rcwa_eval",
    "pirl_rcwa_eval_host", "pirl_env_flip", "pirl_env_reward", "pirl_rcwa_launch_count", "pirl_rcwa_debug_fetch",
    "pirl_rcwa_num_stages", "pirl_rcwa_stage_name", "pirl_rcwa_set_stage_timing", "pirl_rcwa_stage_times",
    "pirl_rcwa_eval_grad", "pirl_rcwa_eval_grad_host", "pirl_rcwa_set_subchunk", "pirl_rcwa_eval_stack",
    "pirl_rcwa_eval_stack_host", "pirl_rcwa_eval_static",
]


class PirlCfg(C.Structure):
    _fields_ = [
        ("n_cells", C.c_int32), ("order", C.c_int32),
        ("n_inc", C.c_double), ("n_exit", C.c_double), ("n_groove", C.c_double), ("theta_deg", C.c_double),
        ("pol", C.c_int32), ("max_batch", C.c_int32), ("device", C.c_int32), ("algo", C.c_int32),
        ("precision", C.c_int32), ("reserved", C.c_int32 * 7),
    ]


_lib = None


def load():
    """dlopen libpirl_rcwa.so (built by build.py / __graft_entry__.build()) and declare every prototype."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing. Build it with `python -m physics_informed_reinforcement_learning_b200.build` "
            "(nvcc, sm_100a). This package has no CPU or PyTorch fallback for the RCWA path.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, dp = C.c_void_p, C.c_int32, C.c_int64, C.c_void_p
    lib.pirl_abi_version.restype = C.c_int
    lib.pirl_last_error.restype = C.c_char_p
    lib.pirl_rcwa_create.argtypes = [C.POINTER(PirlCfg), C.POINTER(vp)]
    lib.pirl_rcwa_destroy.argtypes = [vp]
    lib.pirl_rcwa_eval.argtypes = [vp, dp, dp, i32, i64, dp, dp, dp, dp, vp]
    lib.pirl_rcwa_eval_host.argtypes = [vp, dp, dp, i32, i64, dp, dp, dp, dp]
    lib.pirl_rcwa_eval_static.argtypes = [C.POINTER(PirlCfg), dp, dp, i32, i64, dp, dp, dp, dp, vp]
    lib.pirl_rcwa_eval_grad.argtypes = [vp, dp, dp, i32, i64, dp, dp, dp, vp]
    lib.pirl_rcwa_eval_grad_host.argtypes = [vp, dp, dp, i32, i64, dp, dp, dp]
    lib.pirl_rcwa_eval_stack.argtypes = [vp, dp, dp, i32, dp, i32, i32, i64, dp, dp, dp, dp, vp]
    lib.pirl_rcwa_eval_stack_host.argtypes = [vp, dp, dp, i32, dp, i32, i32, i64, dp, dp, dp, dp]
    lib.pirl_env_flip.argtypes = [dp, dp, i64, i32, vp]
    lib.pirl_env_reward.argtypes = [dp, dp, dp, dp, dp, dp, i64, i32, vp]
    lib.pirl_rcwa_launch_count.argtypes = [vp]
    lib.pirl_rcwa_launch_count.restype = C.c_int64
    lib.pirl_rcwa_debug_fetch.argtypes = [vp, i32, dp, i64]
    lib.pirl_rcwa_num_stages.restype = C.c_int
    lib.pirl_rcwa_stage_name.argtypes = [i32]
    lib.pirl_rcwa_stage_name.restype = C.c_char_p
    lib.pirl_rcwa_set_stage_timing.argtypes = [vp, i32]
    lib.pirl_rcwa_set_subchunk.argtypes = [vp, i32]
    lib.pirl_rcwa_set_subchunk.restype = C.c_int
    lib.pirl_rcwa_stage_times.argtypes = [vp, dp, dp]
    for name in ("pirl_rcwa_create", "pirl_rcwa_destroy", "pirl_rcwa_eval", "pirl_rcwa_eval_host", "pirl_env_flip",
                 "pirl_env_reward", "pirl_rcwa_debug_fetch", "pirl_rcwa_set_stage_timing", "pirl_rcwa_stage_times", "pirl_rcwa_eval_grad",
                 "pirl_rcwa_eval_grad_host", "pirl_rcwa_eval_stack", "pirl_rcwa_eval_stack_host", "pirl_rcwa_eval_static"):
        getattr(lib, name).restype = C.c_int
    _lib = lib
    return lib


def check(rc: int):
    if rc != 0:
        msg = load().pirl_last_error().decode("utf-8", "replace")
        if rc < 0:
            raise ValueError(f"pirl_rcwa argument error {rc}: {msg}")
        raise RuntimeError(f"pirl_rcwa CUDA error {rc}: {msg}")
