"""ORACLE (test infrastructure, NOT product code) — float64 CPU restatement of the reference's
bundle-adjustment hot path.

PARITY: residual side PINNED to the reference's own source, LM state machine UNPINNED.
  * pinned — `rot_mat_from_vec`, `pose_mat_to_vec`, `prepare_params`, `get_residuals` (dense-padded contract) and
    their Jacobians, against golden vectors produced by EXECUTING the reference's source files with NumPy / float64
    torch standing in for `jax.numpy` (tests/golden/make_refsrc_golden.py, make_refsrc_jac_golden.py ->
    refsrc_seed0.npz, refsrc_jac_seed1.npz); `reprojection_errors` against the reference's own
    `Camera.compute_projection_errors_alt` (make_reproj_golden.py -> reproj_small_seed0.npz); the projection against
    COLMAP's four known-answer vectors (`colmap/src/base/cost_functions_test.cc:41-69`); the pose log/exp maps against
    scipy and the reference's `test_pose.txt` fixture.
  * unpinned — `lm_solve` / `_cg_jaxopt`: the LM state machine and CG live in the un-vendored dependency
    `jaxopt==0.7` (`/root/reference/requirements.txt:30`, on `jax==0.4.8`, `:81-82`), absent from this image and from
    /opt/wheelhouse; no reference test pins them.  Restated from SURVEY.md App. A.4.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import this module.

Follows (all paths relative to /root/reference/src/reconstruction/bundle_adjustment/):
  rot_mat_from_vec        utils.py:21-28      -> rot_mat_from_vec
  pose_mat_to_vec         utils.py:31-44      -> pose_mat_to_vec
  parse_intrinsics        utils.py:47-57      -> folded into project()
  reproject_point(s)      bundle_adjustment.py:17-29
  get_residuals           bundle_adjustment.py:52-74   -> ba_residuals (sparse rows, mask==1 only)
  prepare_params          bundle_adjustment.py:93-102  -> pack_params / unpack_params
  jaxopt LevenbergMarquardt 0.7 (init_state / update / run) -> lm_solve
  jax.scipy.sparse.linalg.cg (tol=1e-5, atol=0, maxiter=10n, x0=0, no preconditioner) -> _cg_jaxopt
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import scipy.linalg
import scipy.sparse as sp


# ----------------------------------------------------------------------------- camera algebra
def rot_mat_from_vec(w):
    """utils.py:21-28: R = cos I + (1-cos) k k^T + sin [k]x, theta=|w|, k=w/theta (NaN at 0)."""
    w = np.asarray(w, dtype=np.float64)
    theta = np.linalg.norm(w, axis=-1)[..., None, None]
    with np.errstate(invalid="ignore", divide="ignore"):
        k = w / theta[..., 0]
    K = np.zeros(w.shape[:-1] + (3, 3))
    K[..., 0, 1] = -k[..., 2]
    K[..., 0, 2] = k[..., 1]
    K[..., 1, 0] = k[..., 2]
    K[..., 1, 2] = -k[..., 0]
    K[..., 2, 0] = -k[..., 1]
    K[..., 2, 1] = k[..., 0]
    kk = k[..., :, None] * k[..., None, :]
    eye = np.broadcast_to(np.eye(3), K.shape)
    return np.cos(theta) * eye + (1.0 - np.cos(theta)) * kk + np.sin(theta) * K


def pose_mat_to_vec(pose):
    """utils.py:31-44: SO(3) log via the antisymmetric part; returns [w, t] (...,6)."""
    pose = np.asarray(pose, dtype=np.float64)
    R = pose[..., :3, :3]
    n = np.zeros(R.shape[:-2] + (3,))
    angle = np.arccos(np.clip((R[..., 0, 0] + R[..., 1, 1] + R[..., 2, 2] - 1) / 2, -1, 1))
    n[..., 0] = R[..., 2, 1] - R[..., 1, 2]
    n[..., 1] = R[..., 0, 2] - R[..., 2, 0]
    n[..., 2] = R[..., 1, 0] - R[..., 0, 1]
    norms = np.linalg.norm(n, axis=-1)
    nz = norms != 0
    n[nz] *= (angle[nz] / norms[nz])[..., None]
    return np.concatenate([n, pose[..., :3, 3]], axis=-1)


def pack_params(poses0, intrinsics0, points0):
    """bundle_adjustment.py:93-102: opt_params = [C*6 (w,t) | C*2 (fx,fy) | P*3], cx_cy_skew."""
    intrinsics0 = np.asarray(intrinsics0, dtype=np.float64)
    cam = pose_mat_to_vec(np.asarray(poses0)).reshape(-1)
    return (np.concatenate([cam, intrinsics0[:, :2].reshape(-1), np.asarray(points0).reshape(-1)]),
            intrinsics0[:, 2:].copy())


def unpack_params(x, C):
    cam6 = x[:6 * C].reshape(C, 6)
    f = x[6 * C:8 * C].reshape(C, 2)
    pts = x[8 * C:].reshape(-1, 3)
    return cam6, f, pts


def _skew(v):
    S = np.zeros(v.shape[:-1] + (3, 3))
    S[..., 0, 1] = -v[..., 2]
    S[..., 0, 2] = v[..., 1]
    S[..., 1, 0] = v[..., 2]
    S[..., 1, 2] = -v[..., 0]
    S[..., 2, 0] = -v[..., 1]
    S[..., 2, 1] = v[..., 0]
    return S


def rot_jac_factor(w, R):
    """M with d(R X)/dw = -R [X]x M,  M = (w w^T + (R^T - I)[w]x)/|w|^2  (Gallego & Yezzi 2015)."""
    th2 = (w * w).sum(-1)[..., None, None]
    eye = np.eye(3)
    return (w[..., :, None] * w[..., None, :] + (np.swapaxes(R, -1, -2) - eye) @ _skew(w)) / th2


# ----------------------------------------------------------------------------- residual + Jacobian
def expand_vector2(cam_idx, pt_idx, obs):
    """Rows of the 2-vector residual model: every observation twice, (x, NaN) = its x row,
    (NaN, y) = its y row — the row contract of `b200ba_plan_set_residual_model(plan, 1)`."""
    obs = np.asarray(obs, dtype=np.float64).reshape(-1, 2)
    o2 = np.full((2 * len(obs), 2), np.nan)
    o2[0::2, 0], o2[1::2, 1] = obs[:, 0], obs[:, 1]
    return np.repeat(np.asarray(cam_idx), 2), np.repeat(np.asarray(pt_idx), 2), o2


def ba_eval(x, C, cam_idx, pt_idx, obs, cx_cy_skew, w2, want_jac=True, model="scalar"):
    """Residual r_i = ((u-x)^2+(v-y)^2)/w^2 per valid observation (bundle_adjustment.py:17-21,74)
    and the 8 camera + 3 point non-zeros of each Jacobian row (analytic).

    Returns r (N,), dxdy (N,2), a (N,8) cols [w(3), t(3), fx, fy], b (N,3)."""
    cam6, f, pts = unpack_params(x, C)
    w = cam6[:, :3]
    R = rot_mat_from_vec(w)
    Rc = R[cam_idx]
    X = pts[pt_idx]
    q = (Rc @ X[..., None])[..., 0] + cam6[cam_idx, 3:6]
    fx = f[cam_idx, 0]
    fy = f[cam_idx, 1]
    cx = cx_cy_skew[cam_idx, 0]
    cy = cx_cy_skew[cam_idx, 1]
    s = cx_cy_skew[cam_idx, 2]
    iz = 1.0 / q[:, 2]
    u = (fx * q[:, 0] + s * q[:, 1]) * iz + cx
    v = fy * q[:, 1] * iz + cy
    if model == "vector2":
        # COLMAP's BundleAdjustmentCostFunction (colmap/src/base/cost_functions.h:44-86): residual = projection - observation,
        # one row per component (rows from expand_vector2), scaled by 1/w so that both models share avg_cam_width
        xrow = np.isnan(obs[:, 1])
        iw = 1.0 / np.sqrt(w2)
        dx = np.where(xrow, u - np.where(xrow, obs[:, 0], 0.0), 0.0)
        dy = np.where(xrow, 0.0, v - np.where(xrow, 0.0, obs[:, 1]))
        r = (dx + dy) * iw
        dxdy = np.stack([dx, dy], axis=1)
        if not want_jac:
            return r, dxdy, None, None
        e_override = np.stack([np.where(xrow, iw, 0.0), np.where(xrow, 0.0, iw)], axis=1)
    else:
        dx = u - obs[:, 0]
        dy = v - obs[:, 1]
        r = (dx * dx + dy * dy) / w2
        dxdy = np.stack([dx, dy], axis=1)
        if not want_jac:
            return r, dxdy, None, None
        e_override = None
    N = len(r)
    # d(u,v)/dq
    dpi = np.zeros((N, 2, 3))
    dpi[:, 0, 0] = fx * iz
    dpi[:, 0, 1] = s * iz
    dpi[:, 0, 2] = -(fx * q[:, 0] + s * q[:, 1]) * iz * iz
    dpi[:, 1, 1] = fy * iz
    dpi[:, 1, 2] = -fy * q[:, 1] * iz * iz
    e = (2.0 / w2) * dxdy if e_override is None else e_override  # dr/d(u,v)
    eq = np.einsum("ni,nij->nj", e, dpi)      # dr/dq   (N,3)
    M = rot_jac_factor(w, R)
    dq_dw = -Rc @ _skew(X) @ M[cam_idx]       # (N,3,3)
    a = np.empty((N, 8))
    a[:, 0:3] = np.einsum("nj,njk->nk", eq, dq_dw)
    a[:, 3:6] = eq
    a[:, 6] = e[:, 0] * q[:, 0] * iz
    a[:, 7] = e[:, 1] * q[:, 1] * iz
    b = np.einsum("nj,njk->nk", eq, Rc)
    return r, dxdy, a, b


# ----------------------------------------------------------------------------- robustified 2-vector model
# The reference's COLMAP comparator wraps every 2-vector residual block in a Ceres loss function chosen at
# colmap/src/optim/bundle_adjustment.cc:53-68 (TrivialLoss / SoftLOneLoss(scale) / CauchyLoss(scale)) and Ceres applies it
# by Triggs' correction of the residual and its Jacobian.  Ceres Solver is an EXTERNAL dependency of the reference's colmap
# tree (colmap/CMakeLists.txt:91 `find_package(Ceres REQUIRED)`, no version pinned, sources not under /root/reference), so
# what follows restates its PUBLISHED algorithm (include/ceres/loss_function.h: CauchyLoss / SoftLOneLoss::Evaluate;
# internal/ceres/corrector.cc: Corrector::Corrector, CorrectResiduals, CorrectJacobian).  Parity against Ceres itself is
# UNPINNED; the restatement is anchored by `tests/test_oracle_cpu.py` (gradient of 0.5 sum rho(s) by finite differences
# equals J'^T r' of the corrected rows; a = infinity reproduces the plain vector2 model).
def expand_vector2_pairs(cam_idx, pt_idx, obs):
    """Row contract of the robust residual models (b200ba_plan_set_robust_loss): every observation as an ADJACENT pair of
    rows (x component, then y component), both carrying the full observation (x, y)."""
    obs = np.asarray(obs, dtype=np.float64).reshape(-1, 2)
    return np.repeat(np.asarray(cam_idx), 2), np.repeat(np.asarray(pt_idx), 2), np.repeat(obs, 2, axis=0)


def ceres_rho(s, loss, b):
    """rho(s), rho'(s), rho''(s) of ceres::CauchyLoss(a) / ceres::SoftLOneLoss(a), b = a^2 (loss_function.h)."""
    s = np.asarray(s, dtype=np.float64)
    c = 1.0 / b
    tot = 1.0 + s * c
    if loss in ("cauchy", "ceres_cauchy", 1):
        inv = 1.0 / tot
        return b * np.log1p(s * c), inv, -c * inv * inv  # log1p: Ceres' b log(1 + s/b) without the cancellation at s << b
    if loss in ("soft_l1", "ceres_soft_l1", 2):
        tmp = np.sqrt(tot)
        r1 = 1.0 / tmp
        return 2.0 * s / (tmp + 1.0), r1, -(c * r1) / (2.0 * tot)  # = Ceres' 2b (sqrt(1 + s/b) - 1), cancellation-free
    raise ValueError(f"unknown robust loss {loss!r}")


def ceres_corrector(s, rho1, rho2):
    """Corrector::Corrector (corrector.cc): (sqrt_rho1, residual_scaling, alpha_sq_norm) per residual block."""
    sqrt_rho1 = np.sqrt(rho1)
    simple = (s == 0.0) | (rho2 <= 0.0)
    D = 1.0 + 2.0 * s * rho2 / rho1
    alpha = 1.0 - np.sqrt(np.where(simple, 1.0, D))
    res_scale = np.where(simple, sqrt_rho1, sqrt_rho1 / (1.0 - alpha))
    alpha_sq = np.where(simple, 0.0, alpha / np.where(s == 0.0, 1.0, s))
    return sqrt_rho1, res_scale, alpha_sq


def ba_eval_robust(x, C, cam_idx, pt_idx, obs_pairs, cx_cy_skew, w2, loss, loss_scale, want_jac=True):
    """Corrected rows of the robustified 2-vector model: rows in pairs from `expand_vector2_pairs`.  Residual blocks are
    r = (projection - observation) / w (so that avg_cam_width is shared with the other models: a pixel scale `loss_scale`
    becomes b = (loss_scale / w)^2).  Returns r' (2N,), cterm (2N,: rho(s)/2 on each row of the pair, so that the LM loss
    0.5 * sum(cterm) = 0.5 * sum_obs rho(s), Ceres' cost), a' (2N,8), b' (2N,3)."""
    obs_pairs = np.asarray(obs_pairs, dtype=np.float64)
    n2 = len(obs_pairs)
    o2 = obs_pairs.copy()
    o2[0::2, 1] = np.nan  # even row = x component, odd row = y component
    o2[1::2, 0] = np.nan
    r, _, a, b = ba_eval(x, C, cam_idx, pt_idx, o2, cx_cy_skew, w2, want_jac=want_jac, model="vector2")
    rx, ry = r[0::2], r[1::2]
    s = rx * rx + ry * ry
    bb = float(loss_scale) ** 2 / w2
    rho, rho1, rho2 = ceres_rho(s, loss, bb)
    sq, rs, asq = ceres_corrector(s, rho1, rho2)
    rc = np.empty(n2)
    rc[0::2], rc[1::2] = rs * rx, rs * ry
    cterm = np.repeat(0.5 * rho, 2)
    if not want_jac:
        return rc, cterm, None, None
    # CorrectJacobian: J' = sqrt_rho1 (J - alpha_sq_norm r r^T J) on the 2 x n block
    def corr(Jm):
        Jx, Jy = Jm[0::2], Jm[1::2]
        rtJ = rx[:, None] * Jx + ry[:, None] * Jy
        out = np.empty_like(Jm)
        out[0::2] = sq[:, None] * (Jx - (asq * rx)[:, None] * rtJ)
        out[1::2] = sq[:, None] * (Jy - (asq * ry)[:, None] * rtJ)
        return out
    return rc, cterm, corr(a), corr(b)


def ba_jacobian_sparse(a, b, C, P, cam_idx, pt_idx):
    """Scatter rows into the reference's column order [C*6 | C*2 | P*3] -> CSR (N, 8C+3P)."""
    N = a.shape[0]
    rows = np.repeat(np.arange(N), 11)
    cols = np.empty((N, 11), dtype=np.int64)
    cols[:, 0:6] = 6 * cam_idx[:, None] + np.arange(6)
    cols[:, 6:8] = 6 * C + 2 * cam_idx[:, None] + np.arange(2)
    cols[:, 8:11] = 8 * C + 3 * pt_idx[:, None] + np.arange(3)
    vals = np.concatenate([a, b], axis=1)
    return sp.csr_matrix((vals.reshape(-1), (rows, cols.reshape(-1))), shape=(N, 8 * C + 3 * P))


def rms_reprojection_px(x, C, cam_idx, pt_idx, obs, cx_cy_skew):
    """sqrt(mean(dx^2+dy^2)) in pixels — the gauge-invariant quality figure (camera.py:78-85)."""
    _, dxdy, _, _ = ba_eval(x, C, cam_idx, pt_idx, obs, cx_cy_skew, 1.0, want_jac=False)
    return float(np.sqrt((dxdy ** 2).sum(1).mean()))


def reprojection_errors(x, C, cam_idx, pt_idx, obs, cx_cy_skew, loss="trivial", loss_scale=1.0):
    """Per-observation reprojection error `loss((p2d - reprojection)**2).sum(axis=0)` in pixels^2 and its
    per-camera sums — `Camera.compute_projection_errors_alt` (src/dataset/camera.py:78-85) as
    `Dataset.compute_reprojection_errors_alt` (src/dataset/dataset.py:116-123) calls it once per camera, with
    `LossFunction.TRIVIAL_LOSS` (x) or `CAUCHY_LOSS` (log(1 + x), src/dataset/loss_functions.py:7-17).
    Rows stay in the caller's order; returns (err (N,), per_camera_sum (C,), per_camera_count (C,))."""
    _, dxdy, _, _ = ba_eval(x, C, cam_idx, pt_idx, obs, cx_cy_skew, 1.0, want_jac=False)
    sq = dxdy ** 2
    if loss in ("ceres_cauchy", 2, "ceres_soft_l1", 3):
        # the COLMAP comparator's cost: rho(s), s = |projection - observation|^2 (colmap/src/base/cost_functions.h:44-86),
        # ceres::CauchyLoss(a): b log(1 + s/b); ceres::SoftLOneLoss(a): 2b (sqrt(1 + s/b) - 1); b = a^2 (Ceres Solver's
        # published loss definitions, selected at colmap/src/optim/bundle_adjustment.cc:53-68).  Ceres' cost = sum / 2.
        s, b = sq.sum(axis=1), float(loss_scale) ** 2
        err = b * np.log(1 + s / b) if loss in ("ceres_cauchy", 2) else 2 * b * (np.sqrt(1 + s / b) - 1)
    else:
        if loss in ("cauchy", 1):
            sq = np.log(1 + sq)
        elif loss not in ("trivial", 0):
            raise ValueError(f"unknown loss {loss!r}")
        err = sq.sum(axis=1)
    sums = np.zeros(C)
    for c in range(C):  # ordered per-camera sums (row order), like the reference's per-camera arrays
        sums[c] = err[cam_idx == c].sum()
    return err, sums, np.bincount(cam_idx, minlength=C)


# ----------------------------------------------------------------------------- linear solvers
def _cg_jaxopt(matvec, b, tol=1e-5, maxiter=None):
    """jax.scipy.sparse.linalg.cg as jaxopt.linear_solve.solve_cg calls it: x0=0, M=I,
    stop when r.r <= max(tol^2 b.b, atol^2) with atol=0, maxiter = 10*n."""
    n = b.shape[0]
    maxiter = 10 * n if maxiter is None else maxiter
    atol2 = tol * tol * float(b @ b)
    x = np.zeros_like(b)
    r = b.copy()
    p = r.copy()
    gamma = float(r @ r)
    k = 0
    while gamma > atol2 and k < maxiter:
        Ap = matvec(p)
        alpha = gamma / float(p @ Ap)
        x += alpha * p
        r -= alpha * Ap
        gamma_new = float(r @ r)
        p = r + (gamma_new / gamma) * p
        gamma = gamma_new
        k += 1
    return x, k


def _solve_schur_exact(a, b, g, mu, C, P, cam_idx, pt_idx):
    """Exact solve of (J^T J + mu I) v = g through the Schur complement on the camera block
     with a dense Cholesky of S — arithmetic-equivalent to a direct solve."""
    N = a.shape[0]
    # internal ordering: camera-major 8-blocks then points
    U = np.zeros((C, 8, 8))
    np.add.at(U, cam_idx, a[:, :, None] * a[:, None, :])
    V = np.zeros((P, 3, 3))
    np.add.at(V, pt_idx, b[:, :, None] * b[:, None, :])
    V += mu * np.eye(3)
    Vinv = np.linalg.inv(V)
    gc = np.empty((C, 8))
    gc[:, :6] = g[:6 * C].reshape(C, 6)
    gc[:, 6:] = g[6 * C:8 * C].reshape(C, 2)
    gp = g[8 * C:].reshape(P, 3)
    # W as sparse (8C x 3P)
    rows = (8 * cam_idx[:, None, None] + np.arange(8)[None, :, None]) + np.zeros((1, 1, 3), dtype=np.int64)
    cols = (3 * pt_idx[:, None, None] + np.arange(3)[None, None, :]) + np.zeros((1, 8, 1), dtype=np.int64)
    vals = a[:, :, None] * b[:, None, :]
    W = sp.csr_matrix((vals.reshape(-1), (rows.reshape(-1), cols.reshape(-1))), shape=(8 * C, 3 * P))
    Vinv_bd = sp.block_diag([sp.csr_matrix(Vinv[p]) for p in range(P)], format="csr") if P <= 2000 else \
        _block_diag3(Vinv)
    WVi = W @ Vinv_bd
    S = -(WVi @ W.T).toarray()
    for c in range(C):
        S[8 * c:8 * c + 8, 8 * c:8 * c + 8] += U[c] + mu * np.eye(8)
    rhs = gc.reshape(-1) - WVi @ gp.reshape(-1)
    cho = scipy.linalg.cho_factor(S, lower=True, check_finite=False)
    vc = scipy.linalg.cho_solve(cho, rhs, check_finite=False)
    vp = np.einsum("pij,pj->pi", Vinv, gp - (W.T @ vc).reshape(P, 3))
    v = np.empty_like(g)
    vc8 = vc.reshape(C, 8)
    v[:6 * C] = vc8[:, :6].reshape(-1)
    v[6 * C:8 * C] = vc8[:, 6:].reshape(-1)
    v[8 * C:] = vp.reshape(-1)
    return v


def _block_diag3(B):
    P = B.shape[0]
    rows = (3 * np.arange(P)[:, None, None] + np.arange(3)[None, :, None]) + np.zeros((1, 1, 3), dtype=np.int64)
    cols = (3 * np.arange(P)[:, None, None] + np.arange(3)[None, None, :]) + np.zeros((1, 3, 1), dtype=np.int64)
    return sp.csr_matrix((B.reshape(-1), (rows.reshape(-1), cols.reshape(-1))), shape=(3 * P, 3 * P))


# ----------------------------------------------------------------------------- LM state machine
@dataclass
class LMTrace:
    loss: list = field(default_factory=list)        # loss after each update (index 0 = init)
    grad_l2: list = field(default_factory=list)
    mu: list = field(default_factory=list)
    rho: list = field(default_factory=list)
    accepted: list = field(default_factory=list)
    inner_iters: list = field(default_factory=list)
    n_eval_jac: int = 0
    n_eval_res: int = 0


@dataclass
class LMResult:
    params: np.ndarray
    iter_num: int
    loss: float
    grad_l2: float
    damping_factor: float
    increase_factor: float
    trace: LMTrace


def lm_solve(x0, C, P, cam_idx, pt_idx, obs, cx_cy_skew, avg_cam_width, tol=1e-6, maxiter=100,
             damping_parameter=1e-6, solver="exact", round_jac_f32=False, cg_tol=1e-5, model="scalar", robust=None) -> LMResult:
    """jaxopt 0.7 LevenbergMarquardt.run as the reference configures it
    (bundle_adjustment.py:83-91: tol=1e-6, maxiter=100, default solver, 'grad-l2-norm' stop).

    solver: 'exact'     — Schur + dense Cholesky (the parity target)
            'jaxopt_cg' — the reference's actual inner solver: unpreconditioned CG on the full
                          normal equations, rel. tol 1e-5, maxiter 10n (`cg_tol` to tighten it)
    round_jac_f32: emulate fp32 storage of Jacobian rows (design study; gradient stays exact).
    model: 'scalar' (the reference) or 'vector2' (rows from expand_vector2: the COLMAP comparator's residual).
    robust: None or (loss, scale) with model='vector2' and rows from expand_vector2_pairs: Ceres' CauchyLoss / SoftLOneLoss
            applied by its corrector (ba_eval_robust); the LM loss is 0.5 * sum rho(s).
    Accept/reject uses a select instead of jaxopt's arithmetic blend."""
    cam_idx = np.asarray(cam_idx, dtype=np.int64)
    pt_idx = np.asarray(pt_idx, dtype=np.int64)
    w2 = float(avg_cam_width) ** 2
    x = np.array(x0, dtype=np.float64)
    tr = LMTrace()

    if robust is not None and model != "vector2":
        raise ValueError("robust losses belong to the vector2 model")

    def residual(xx, want_jac):
        """rows r, a, b and the loss 0.5 * sum of the rows' cost terms"""
        if robust is None:
            r, _, a, b = ba_eval(xx, C, cam_idx, pt_idx, obs, cx_cy_skew, w2, want_jac=want_jac, model=model)
            return r, a, b, 0.5 * float(r @ r)
        r, ct, a, b = ba_eval_robust(xx, C, cam_idx, pt_idx, obs, cx_cy_skew, w2, robust[0], robust[1], want_jac=want_jac)
        return r, a, b, 0.5 * float(ct.sum())

    cur = {}

    def evaluate(xx):
        r, a, b, cur["loss"] = residual(xx, True)
        J = ba_jacobian_sparse(a, b, C, P, cam_idx, pt_idx)
        g = J.T @ r
        if round_jac_f32:
            a = a.astype(np.float32).astype(np.float64)
            b = b.astype(np.float32).astype(np.float64)
            J = ba_jacobian_sparse(a, b, C, P, cam_idx, pt_idx)
        tr.n_eval_jac += 1
        return r, a, b, J, g

    # init_state
    r, a, b, J, g = evaluate(x)
    diag = np.asarray(J.multiply(J).sum(axis=0)).reshape(-1)
    mu = damping_parameter * float(diag.max())
    nu = 2.0
    loss = cur["loss"]
    err = float(np.linalg.norm(g))
    it = 0
    tr.loss.append(loss)
    tr.grad_l2.append(err)
    tr.mu.append(mu)

    while True:
        # ---- update()
        if solver == "exact":
            v = _solve_schur_exact(a, b, g, mu, C, P, cam_idx, pt_idx)
            inner = 0
        elif solver == "jaxopt_cg":
            v, inner = _cg_jaxopt(lambda p: J.T @ (J @ p) + mu * p, g, tol=cg_tol)
        else:
            raise ValueError(solver)
        delta = -v
        x_next = x + delta
        _, _, _, loss_next = residual(x_next, False)
        tr.n_eval_res += 1
        rho = (loss - loss_next) / (0.5 * float(delta @ (mu * delta - g)))
        ok = bool(rho > 0.0)
        if ok:
            x = x_next
            r, a, b, J, g = evaluate(x)
            mu = mu * max(1.0 / 3.0, 1.0 - (2.0 * rho - 1.0) ** 3)
            nu = 2.0
        else:
            mu = mu * nu
            nu = 2.0 * nu
        it += 1
        loss = cur["loss"]
        err = float(np.linalg.norm(g))
        tr.loss.append(loss)
        tr.grad_l2.append(err)
        tr.mu.append(mu)
        tr.rho.append(rho)
        tr.accepted.append(ok)
        tr.inner_iters.append(inner)
        # ---- run(): first update unconditional, then while iter_num < maxiter and error > tol
        if not (it < maxiter and err > tol):
            break
    return LMResult(x, it, loss, err, mu, nu, tr)


def lm_solve_dense_c(x0, C, P, cam_idx, pt_idx, obs, cx_cy_skew, avg_cam_width, tol=1e-6, maxiter=100,
                     damping_parameter=1e-6, verbose=False) -> LMResult:
    """`lm_solve(solver="exact")` for shapes scipy.sparse cannot reach (Venice-1778: 5 M rows): the same LM state machine,
    rows from the C restatement (`ba_c_eval`), the reduced camera system formed densely in C (`ba_c_schur_dense`,
    S is 14 224^2 fp64 = 1.6 GB at Venice) and factored by LAPACK Cholesky, points back-substituted in NumPy.
    Same mathematics as `_solve_schur_exact`; `tests/test_oracle_cpu.py` holds the two equal on the small shapes."""
    from . import c_binding as cb
    cam_idx = np.ascontiguousarray(cam_idx, dtype=np.int32)
    pt_idx = np.ascontiguousarray(pt_idx, dtype=np.int32)
    obs = np.ascontiguousarray(obs, dtype=np.float64)
    k3 = np.ascontiguousarray(cx_cy_skew, dtype=np.float64)
    w2 = float(avg_cam_width) ** 2
    x = np.array(x0, dtype=np.float64)
    tr = LMTrace()

    def seg(idx, w, n):
        return np.stack([np.bincount(idx, weights=w[:, k], minlength=n) for k in range(w.shape[1])], axis=1)

    def evaluate(xx):
        r, a, b, cost = cb.ba_eval(xx, C, P, cam_idx, pt_idx, obs, k3, w2, want_jac=True)
        gc = seg(cam_idx, a * r[:, None], C)
        gp = seg(pt_idx, b * r[:, None], P)
        tr.n_eval_jac += 1
        return r, a, b, gc, gp

    def pack(vc8, vp):
        return np.concatenate([vc8[:, :6].reshape(-1), vc8[:, 6:].reshape(-1), vp.reshape(-1)])

    r, a, b, gc, gp = evaluate(x)
    maxdiag = max(seg(cam_idx, a * a, C).max(), seg(pt_idx, b * b, P).max())
    mu, nu = damping_parameter * float(maxdiag), 2.0
    loss = 0.5 * float(r @ r)
    err = float(np.sqrt((gc ** 2).sum() + (gp ** 2).sum()))
    it = 0
    tr.loss.append(loss); tr.grad_l2.append(err); tr.mu.append(mu)
    while True:
        S, rhs, Vinv, gp_c = cb.schur_dense(C, P, cam_idx, pt_idx, a, b, r, mu)
        cho = scipy.linalg.cho_factor(S, lower=True, overwrite_a=True, check_finite=False)
        vc = scipy.linalg.cho_solve(cho, rhs, check_finite=False).reshape(C, 8)
        del S, cho
        sig = np.einsum("nk,nk->n", a, vc[cam_idx])
        t = seg(pt_idx, b * sig[:, None], P)
        vp = np.einsum("pij,pj->pi", Vinv, gp_c - t)
        g = pack(gc, gp)
        v = pack(vc, vp)
        x_next = x - v
        r_next, _, _, _ = cb.ba_eval(x_next, C, P, cam_idx, pt_idx, obs, k3, w2, want_jac=False)
        tr.n_eval_res += 1
        loss_next = 0.5 * float(r_next @ r_next)
        rho = (loss - loss_next) / (0.5 * float(mu * (v @ v) + v @ g))  # delta = -v
        ok = bool(rho > 0.0)
        if ok:
            x = x_next
            r, a, b, gc, gp = evaluate(x)
            mu = mu * max(1.0 / 3.0, 1.0 - (2.0 * rho - 1.0) ** 3)
            nu = 2.0
        else:
            mu = mu * nu
            nu = 2.0 * nu
        it += 1
        loss = 0.5 * float(r @ r)
        err = float(np.sqrt((gc ** 2).sum() + (gp ** 2).sum()))
        tr.loss.append(loss); tr.grad_l2.append(err); tr.mu.append(mu); tr.rho.append(rho); tr.accepted.append(ok)
        tr.inner_iters.append(0)
        if verbose:
            print(f"[oracle dense] it {it} loss {loss:.12e} |g| {err:.3e} mu {mu:.3e} rho {rho:.3e} acc {ok}", flush=True)
        if not (it < maxiter and err > tol):
            break
    return LMResult(x, it, loss, err, mu, nu, tr)
