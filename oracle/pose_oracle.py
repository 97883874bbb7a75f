"""ORACLE (test infrastructure, NOT product code) — float64 CPU restatement of the reference's
batched single-pose refinement.  PARITY: `pose_eval` (residuals and Jacobian rows, both losses, masked rows) is PINNED to
golden vectors produced by executing the reference's own `PoseOptimization.get_residuals` source (tests/golden/
refsrc_seed0.npz, refsrc_jac_seed1.npz); the LM loop `pose_lm_batched` is UNPINNED for the same reason as
`ba_oracle.lm_solve` (jaxopt 0.7 absent; no reference test pins the path).

Follows (relative to /root/reference/src/reconstruction/bundle_adjustment/):
  PoseOptimization.get_residuals   pose_optimization.py:33-52  -> pose_eval
  cauchy_loss / l2_loss            loss.py:7-14
  JaxPoseOptimizer.create_lm_optimizer pose_optimization.py:62-71
       LevenbergMarquardt(tol=1e-7, maxiter=100, solver="cholesky"), jit(vmap(lm.run))
  jaxopt solve_cholesky: materialise J^T J (8x8), add ridge mu*I, solve(assume_a='pos')
  vmap lock-step with frozen finished lanes == independent solves per pose

Vectorised over the batch; every lane carries its own LM state.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .ba_oracle import _skew, rot_jac_factor, rot_mat_from_vec

L2, CAUCHY = 0, 1


def pose_eval(params, points, obs, cx_cy_skew, mask, w2, loss=CAUCHY, want_jac=True):
    """params (B,8)=[w,t,fx,fy]; points (B,N,3); obs (B,N,2); cx_cy_skew (B,3); mask (B,N).
    Returns r (B,N) and J (B,N,8)."""
    w = params[:, :3]
    R = rot_mat_from_vec(w)                                        # (B,3,3)
    q = np.einsum("bij,bnj->bni", R, points) + params[:, None, 3:6]
    fx = params[:, None, 6]
    fy = params[:, None, 7]
    cx = cx_cy_skew[:, None, 0]
    cy = cx_cy_skew[:, None, 1]
    s = cx_cy_skew[:, None, 2]
    iz = 1.0 / q[..., 2]
    u = (fx * q[..., 0] + s * q[..., 1]) * iz + cx
    v = fy * q[..., 1] * iz + cy
    dx = obs[..., 0] - u
    dy = obs[..., 1] - v
    if loss == CAUCHY:
        r = (np.log(1 + dx * dx) + np.log(1 + dy * dy)) * mask / w2
        ex = -2 * dx / (1 + dx * dx)       # d r / d u  (dx = obs - u)
        ey = -2 * dy / (1 + dy * dy)
    else:
        r = (dx * dx + dy * dy) * mask / w2
        ex = -2 * dx
        ey = -2 * dy
    if not want_jac:
        return r, None
    ex = ex * mask / w2
    ey = ey * mask / w2
    B, N = r.shape
    eq = np.empty((B, N, 3))
    eq[..., 0] = ex * fx * iz
    eq[..., 1] = (ex * s + ey * fy) * iz
    eq[..., 2] = -(ex * (fx * q[..., 0] + s * q[..., 1]) + ey * fy * q[..., 1]) * iz * iz
    M = rot_jac_factor(w, R)                                       # (B,3,3)
    RM = -R[:, None] @ _skew(points) @ M[:, None]                  # (B,N,3,3)
    J = np.empty((B, N, 8))
    J[..., 0:3] = np.einsum("bnj,bnjk->bnk", eq, RM)
    J[..., 3:6] = eq
    J[..., 6] = ex * q[..., 0] * iz
    J[..., 7] = ey * q[..., 1] * iz
    return r, J


@dataclass
class PoseLMResult:
    params: np.ndarray      # (B,8)
    iter_num: np.ndarray    # (B,) int
    loss: np.ndarray        # (B,)
    grad_l2: np.ndarray     # (B,)
    damping_factor: np.ndarray
    grad_history: np.ndarray = None  # (B, iterations+1): ||g|| of every lane after each of ITS updates (NaN once the lane has stopped)


def _chol_solve(A, b):
    """Batched SPD solve; a failed factorisation yields NaN like LAPACK posv through jax."""
    out = np.full_like(b, np.nan)
    for i in range(A.shape[0]):
        try:
            L = np.linalg.cholesky(A[i])
            y = np.linalg.solve(L, b[i])
            out[i] = np.linalg.solve(L.T, y)
        except np.linalg.LinAlgError:
            pass
    return out


def pose_lm_batched(params0, points, obs, cx_cy_skew, mask, avg_cam_width, loss=CAUCHY, tol=1e-7,
                    maxiter=100, damping_parameter=1e-6) -> PoseLMResult:
    w2 = float(avg_cam_width) ** 2
    x = np.array(params0, dtype=np.float64)
    B = x.shape[0]
    r, J = pose_eval(x, points, obs, cx_cy_skew, mask, w2, loss)
    JtJ = np.einsum("bni,bnj->bij", J, J)
    g = np.einsum("bni,bn->bi", J, r)
    mu = damping_parameter * np.max(np.einsum("bii->bi", JtJ), axis=1)
    nu = np.full(B, 2.0)
    cur = 0.5 * (r * r).sum(1)
    err = np.linalg.norm(g, axis=1)
    it = np.zeros(B, dtype=np.int64)
    active = np.ones(B, dtype=bool)        # first update unconditional
    eye = np.eye(8)
    hist = [err.copy()]
    while active.any():
        ia = np.flatnonzero(active)
        v = _chol_solve(JtJ[ia] + mu[ia, None, None] * eye, g[ia])
        delta = -v
        xn = x[ia] + delta
        with np.errstate(all="ignore"):
            rn, Jn = pose_eval(xn, points[ia], obs[ia], cx_cy_skew[ia], mask[ia], w2, loss)
            ln = 0.5 * (rn * rn).sum(1)
            rho = (cur[ia] - ln) / (0.5 * np.einsum("bi,bi->b", delta, mu[ia, None] * delta - g[ia]))
        ok = rho > 0.0
        acc = ia[ok]
        rej = ia[~ok]
        x[acc] = xn[ok]
        JtJ[acc] = np.einsum("bni,bnj->bij", Jn[ok], Jn[ok])
        g[acc] = np.einsum("bni,bn->bi", Jn[ok], rn[ok])
        cur[acc] = ln[ok]
        mu[acc] = mu[acc] * np.maximum(1.0 / 3.0, 1.0 - (2.0 * rho[ok] - 1.0) ** 3)
        nu[acc] = 2.0
        mu[rej] = mu[rej] * nu[rej]
        nu[rej] = 2.0 * nu[rej]
        it[ia] += 1
        err[ia] = np.linalg.norm(g[ia], axis=1)
        h = np.full(B, np.nan)
        h[ia] = err[ia]
        hist.append(h)
        active[ia] = (it[ia] < maxiter) & (err[ia] > tol)
    return PoseLMResult(x, it, cur, err, mu, np.stack(hist, axis=1))
