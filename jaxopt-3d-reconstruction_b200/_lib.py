"""ctypes binding of libb200ba.so (the C-ABI declared in include/b200ba.h).

There is no CPU fallback: if the shared library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200BA_LIB") or os.path.join(_HERE, "libb200ba.so")  # B200BA_LIB: kernel-variant experiments

SYMBOLS = [
    "b200ba_last_error", "b200ba_version", "b200ba_plan_create", "b200ba_plan_destroy",
    "b200ba_plan_set_observations", "b200ba_nccl_unique_id", "b200ba_plan_set_comm",
    "b200ba_residual_jacobian", "b200ba_lm_default_opts", "b200ba_lm_solve", "b200ba_linearize",
    "b200ba_schur_prepare", "b200ba_schur_matvec", "b200ba_schur_solve", "b200ba_cost",
    "b200ba_pose_lm_batched", "b200ba_launch_count", "b200ba_profile_enable", "b200ba_profile_get",
    "b200ba_reprojection_errors", "b200ba_plan_set_residual_model", "b200ba_plan_info", "b200ba_with_noise",
    "b200ba_cost_noalloc", "b200ba_plan_set_robust_loss",
]


class LMOpts(C.Structure):
    _fields_ = [("maxiter", C.c_int), ("tol", C.c_double), ("damping_parameter", C.c_double),
                ("pcg_rtol", C.c_double), ("pcg_maxiter", C.c_int), ("fixed_pcg_iters", C.c_int),
                ("verbose", C.c_int)]


class LMResult(C.Structure):
    _fields_ = [("iter_num", C.c_int), ("status", C.c_int), ("loss", C.c_double),
                ("grad_l2", C.c_double), ("damping_factor", C.c_double),
                ("increase_factor", C.c_double), ("n_jac_evals", C.c_int64),
                ("n_res_evals", C.c_int64), ("n_pcg_iters_total", C.c_int64),
                ("loss_trajectory_host", C.POINTER(C.c_double)),
                ("grad_trajectory_host", C.POINTER(C.c_double)),
                ("mu_trajectory_host", C.POINTER(C.c_double)),
                ("pcg_iters_host", C.POINTER(C.c_int)), ("accepted_host", C.POINTER(C.c_int))]


_lib = None


class B200BAError(RuntimeError):
    pass


def load():
    """Load the library once; raise loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200BAError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C jaxopt-3d-reconstruction_b200/csrc`. There is no CPU fallback.")
    # torch ships the NCCL this library resolves against; importing it first makes the loader
    # reuse torch's libnccl.so.2 instead of looking for a system copy.
    import torch  # noqa: F401
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, i32, i64, dbl = C.c_void_p, C.c_int, C.c_int64, C.c_double
    lib.b200ba_last_error.restype = C.c_char_p
    lib.b200ba_version.restype = i32
    lib.b200ba_launch_count.restype = i64
    lib.b200ba_plan_create.argtypes = [i32, i32, i64, vp, vp, i32, vp, C.POINTER(vp)]
    lib.b200ba_plan_destroy.argtypes = [vp]
    lib.b200ba_plan_set_observations.argtypes = [vp, vp, vp]
    lib.b200ba_nccl_unique_id.argtypes = [C.c_char_p]
    lib.b200ba_plan_set_comm.argtypes = [vp, i32, i32, C.c_char_p]
    lib.b200ba_residual_jacobian.argtypes = [vp, vp, vp, vp, dbl, vp, vp, vp, C.POINTER(dbl), vp]
    lib.b200ba_lm_default_opts.argtypes = [C.POINTER(LMOpts)]
    lib.b200ba_lm_default_opts.restype = None
    lib.b200ba_lm_solve.argtypes = [vp, vp, vp, vp, dbl, C.POINTER(LMOpts), C.POINTER(LMResult), vp]
    lib.b200ba_linearize.argtypes = [vp, vp, vp, vp, dbl, vp, vp, vp, vp, C.POINTER(dbl), vp]
    lib.b200ba_schur_prepare.argtypes = [vp, dbl, vp, vp]
    lib.b200ba_schur_matvec.argtypes = [vp, vp, vp, vp]
    lib.b200ba_schur_solve.argtypes = [vp, dbl, i32, vp, vp, C.POINTER(i32), vp]
    lib.b200ba_cost.argtypes = [vp, vp, vp, vp, dbl, C.POINTER(dbl), vp]
    lib.b200ba_cost_noalloc.argtypes = [i32, i32, i64, vp, vp, vp, vp, vp, vp, dbl, i32, vp, vp]
    lib.b200ba_plan_set_residual_model.argtypes = [vp, i32]
    lib.b200ba_plan_set_robust_loss.argtypes = [vp, i32, dbl]
    lib.b200ba_reprojection_errors.argtypes = [vp, vp, vp, vp, i32, dbl, vp, vp, vp]
    lib.b200ba_pose_lm_batched.argtypes = [i32, i32, vp, vp, vp, vp, vp, dbl, i32, i32, dbl, dbl,
                                           vp, vp, vp, vp, vp, vp]
    lib.b200ba_plan_info.argtypes = [vp, C.POINTER(i32)]
    lib.b200ba_with_noise.argtypes = [i32, i32, i64, vp, vp, vp, vp, C.c_uint64, dbl, dbl, dbl, dbl, dbl, i32, vp, vp, vp, vp, vp, vp]
    lib.b200ba_profile_enable.argtypes = [vp, i32]
    lib.b200ba_profile_get.argtypes = [vp, C.POINTER(dbl), C.POINTER(i64)]
    for name in SYMBOLS:  # a stale library (older build without a newer entry point) must fail here, not at first use
        if not hasattr(lib, name):
            raise B200BAError(f"{LIB_PATH} lacks {name}: rebuild it (make -C jaxopt-3d-reconstruction_b200/csrc)")
    _lib = lib
    return lib


def default_pcg_rtol():
    """The library's default PCG tolerance (b200ba_lm_default_opts)."""
    o = LMOpts()
    load().b200ba_lm_default_opts(C.byref(o))
    return float(o.pcg_rtol)


def check(rc, what=""):
    if rc != 0:
        msg = load().b200ba_last_error().decode(errors="replace")
        raise B200BAError(f"{what} failed (code {rc}): {msg}")
