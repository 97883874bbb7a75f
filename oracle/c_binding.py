"""ORACLE (test infrastructure, NOT product code) — ctypes binding of oracle/ba_oracle_c.c, the
OpenMP C restatement of the reference's BA path in its own formulation (full-system CG).  Used by
bench.py's cpu_baseline / `--impl reference` legs and tests/ only."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "libba_oracle_c.so")
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            import subprocess
            subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
        # spinning OpenMP workers oversubscribe shared host cores (measured 10x slower here)
        os.environ.setdefault("OMP_WAIT_POLICY", "PASSIVE")
        lib = C.CDLL(_PATH)
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int32)
        lib.ba_c_num_threads.restype = C.c_int
        lib.ba_c_set_threads.argtypes = [C.c_int]
        lib.ba_c_set_threads.restype = None
        lib.ba_c_eval.restype = C.c_double
        lib.ba_c_eval.argtypes = [C.c_int, C.c_int, C.c_int64, dp, ip, ip, dp, dp, C.c_double, dp, dp, dp, C.c_int]
        lib.ba_c_lm.restype = C.c_int
        lib.ba_c_lm.argtypes = [C.c_int, C.c_int, C.c_int64, dp, ip, ip, dp, dp, C.c_double, C.c_double, C.c_int,
                                C.c_double, C.c_double, C.c_int64, dp, dp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        lib.ba_c_schur_dense.restype = C.c_int
        lib.ba_c_schur_dense.argtypes = [C.c_int, C.c_int, C.c_int64, ip, ip, dp, dp, dp, C.c_double, dp, dp, dp, dp]
        _lib = lib
    return _lib


def _d(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _i(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def num_threads():
    return int(load().ba_c_num_threads())


def set_num_threads(n):
    """OpenMP team size of the C restatement (torch.distributed.run exports OMP_NUM_THREADS=1 to its workers)."""
    load().ba_c_set_threads(int(n))


def ba_eval(x, Cn, P, cam_idx, pt_idx, obs, k3, w2, want_jac=True):
    lib = load()
    x = np.ascontiguousarray(x, dtype=np.float64)
    cam = np.ascontiguousarray(cam_idx, dtype=np.int32)
    pt = np.ascontiguousarray(pt_idx, dtype=np.int32)
    obs = np.ascontiguousarray(obs, dtype=np.float64)
    k3 = np.ascontiguousarray(k3, dtype=np.float64)
    N = len(cam)
    r = np.empty(N)
    a = np.empty((N, 8))
    b = np.empty((N, 3))
    cost = lib.ba_c_eval(Cn, P, N, _d(x), _i(cam), _i(pt), _d(obs), _d(k3), float(w2), _d(r), _d(a), _d(b), int(want_jac))
    return r, a, b, cost


def lm_solve(x0, Cn, P, cam_idx, pt_idx, obs, k3, avg_cam_width, tol=1e-6, maxiter=100, damping=1e-6, cg_tol=1e-5,
             cg_cap=0):
    """jaxopt LM with its default CG inner solver (cg_cap > 0 truncates the inner CG)."""
    lib = load()
    x = np.array(x0, dtype=np.float64)
    cam = np.ascontiguousarray(cam_idx, dtype=np.int32)
    pt = np.ascontiguousarray(pt_idx, dtype=np.int32)
    obs = np.ascontiguousarray(obs, dtype=np.float64)
    k3 = np.ascontiguousarray(k3, dtype=np.float64)
    m = max(int(maxiter), 1)
    loss = np.zeros(m + 1)
    grad = np.zeros(m + 1)
    cg = np.zeros(m, dtype=np.int32)
    acc = np.zeros(m, dtype=np.int32)
    it = lib.ba_c_lm(Cn, P, len(cam), _d(x), _i(cam), _i(pt), _d(obs), _d(k3), float(avg_cam_width) ** 2, float(tol),
                     int(maxiter), float(damping), float(cg_tol), int(cg_cap), _d(loss), _d(grad),
                     cg.ctypes.data_as(C.POINTER(C.c_int)), acc.ctypes.data_as(C.POINTER(C.c_int)))
    return dict(params=x, iter_num=it, loss=loss[:it + 1], grad_l2=grad[:it + 1], cg_iters=cg[:it], accepted=acc[:it])


def schur_dense(Cn, P, cam_idx, pt_idx, a, b, r, mu):
    """Dense reduced camera system S (8C x 8C), rhs (8C,), Vinv (P,3,3), g_p (P,3) — ba_c_schur_dense."""
    lib = load()
    cam = np.ascontiguousarray(cam_idx, dtype=np.int32)
    pt = np.ascontiguousarray(pt_idx, dtype=np.int32)
    n = 8 * Cn
    S = np.zeros((n, n))
    rhs = np.empty(n)
    Vinv = np.empty((P, 3, 3))
    gp = np.empty((P, 3))
    bad = lib.ba_c_schur_dense(Cn, P, len(cam), _i(cam), _i(pt), _d(np.ascontiguousarray(a)), _d(np.ascontiguousarray(b)),
                               _d(np.ascontiguousarray(r)), float(mu), _d(S), _d(rhs), _d(Vinv), _d(gp))
    if bad:
        raise FloatingPointError("schur_dense: a point block V_p + mu I is not positive definite")
    return S, rhs, Vinv, gp
