"""Design study (CPU, numpy): inner-solver variants on the reduced camera system along the exact LM
trajectory — standard block-Jacobi PCG, Chronopoulos-Gear, pipelined PCG (Ghysels-Vanroose) and a
warm start from the previous LM step.  Reports iterations to rtol and the TRUE relative residual.
Not a test (no test_ prefix); run: python tests/studies/pcg_study.py ladybug49"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import scipy.sparse as sp

import b200ba
from oracle import ba_oracle as bo


def schur_system(a, b, g, mu, C, P, cam_idx, pt_idx):
    U = np.zeros((C, 8, 8)); np.add.at(U, cam_idx, a[:, :, None] * a[:, None, :])
    V = np.zeros((P, 3, 3)); np.add.at(V, pt_idx, b[:, :, None] * b[:, None, :])
    V += mu * np.eye(3)
    Vinv = np.linalg.inv(V)
    gc = np.empty((C, 8)); gc[:, :6] = g[:6 * C].reshape(C, 6); gc[:, 6:] = g[6 * C:8 * C].reshape(C, 2)
    gp = g[8 * C:].reshape(P, 3)
    rows = (8 * cam_idx[:, None, None] + np.arange(8)[None, :, None]) + np.zeros((1, 1, 3), dtype=np.int64)
    cols = (3 * pt_idx[:, None, None] + np.arange(3)[None, None, :]) + np.zeros((1, 8, 1), dtype=np.int64)
    W = sp.csr_matrix(((a[:, :, None] * b[:, None, :]).reshape(-1), (rows.reshape(-1), cols.reshape(-1))), shape=(8 * C, 3 * P))
    Vb = bo._block_diag3(Vinv)
    Ud = sp.block_diag([U[c] + mu * np.eye(8) for c in range(C)], format="csr")
    WV = (W @ Vb).tocsr()
    Wt = W.T.tocsr()
    mv = lambda x: Ud @ x - WV @ (Wt @ x)
    rhs = gc.reshape(-1) - WV @ gp.reshape(-1)
    # block-Jacobi blocks of S
    beta = np.einsum("ni,nij,nj->n", b, Vinv[pt_idx], b)
    D = U + mu * np.eye(8)
    np.add.at(D, cam_idx, -(a * beta[:, None])[:, :, None] * a[:, None, :])
    Minv = np.linalg.inv(D)
    prec = lambda r: np.einsum("cij,cj->ci", Minv, r.reshape(C, 8)).reshape(-1)
    return mv, prec, rhs


def pcg(mv, prec, b, rtol, x0=None, maxiter=1000):
    x = np.zeros_like(b) if x0 is None else x0.copy()
    r = b - mv(x) if x0 is not None else b.copy()
    z = prec(r); p = z.copy(); rz = r @ z; bb = b @ b
    k = 0
    while r @ r > rtol * rtol * bb and k < maxiter:
        Ap = mv(p); al = rz / (p @ Ap); x += al * p; r -= al * Ap
        z = prec(r); rz2 = r @ z; p = z + (rz2 / rz) * p; rz = rz2; k += 1
    return x, k


def pipelined_pcg(mv, prec, b, rtol, x0=None, maxiter=1000):
    """Ghysels & Vanroose 2014, Alg. 3 (preconditioned pipelined CG)."""
    x = np.zeros_like(b) if x0 is None else x0.copy()
    r = b - mv(x) if x0 is not None else b.copy()
    u = prec(r); w = mv(u)
    z = q = s = p = None
    bb = b @ b; k = 0; gam_old = al_old = None
    while True:
        gam = r @ u; dl = w @ u; rr = r @ r
        if rr <= rtol * rtol * bb or k >= maxiter:
            break
        m = prec(w); n = mv(m)
        if k == 0:
            be = 0.0; al = gam / dl
            z, q, s, p = n.copy(), m.copy(), w.copy(), u.copy()
        else:
            be = gam / gam_old; al = gam / (dl - be * gam / al_old)
            z = n + be * z; q = m + be * q; s = w + be * s; p = u + be * p
        x += al * p; r -= al * s; u -= al * q; w -= al * z
        gam_old, al_old = gam, al; k += 1
    return x, k


def main(shape="ladybug49", rtol=1e-10):
    sc = b200ba.synth.make_ba_scene(shape, seed=0)
    x, k3 = bo.pack_params(sc.poses0, sc.intrinsics0, sc.points0)
    C, P = sc.C, sc.P
    ci, pi = sc.cam_idx.astype(np.int64), sc.pt_idx.astype(np.int64)
    w2 = sc.avg_cam_width ** 2
    r, _, a, b = bo.ba_eval(x, C, ci, pi, sc.obs, k3, w2)
    J = bo.ba_jacobian_sparse(a, b, C, P, ci, pi); g = J.T @ r
    mu = 1e-6 * float(np.asarray(J.multiply(J).sum(axis=0)).max()); nu = 2.0
    loss = 0.5 * r @ r; prev = None
    for it in range(12):
        a32 = a.astype(np.float32).astype(np.float64); b32 = b.astype(np.float32).astype(np.float64)
        mv, prec, rhs = schur_system(a32, b32, g, mu, C, P, ci, pi)
        xs, k_std = pcg(mv, prec, rhs, rtol)
        xp, k_pipe = pipelined_pcg(mv, prec, rhs, rtol)
        tr = lambda v: np.linalg.norm(rhs - mv(v)) / np.linalg.norm(rhs)
        line = f"it {it} mu {mu:.2e} std {k_std} (true {tr(xs):.1e}) pipe {k_pipe} (true {tr(xp):.1e}) |xp-xs|/|xs| {np.linalg.norm(xp-xs)/np.linalg.norm(xs):.1e}"
        if prev is not None:
            for sc_ in (0.5, 1.0):
                xw, k_w = pcg(mv, prec, rhs, rtol, x0=sc_ * prev)
                line += f" warm{sc_} {k_w}+1 (r0 {tr(sc_*prev):.1e})"
            al = (prev @ rhs) / (prev @ mv(prev))
            xw, k_w = pcg(mv, prec, rhs, rtol, x0=al * prev)
            line += f" warm-opt({al:.2f}) {k_w}+2"
        print(line, flush=True)
        prev = xs
        v = bo._solve_schur_exact(a, b, g, mu, C, P, ci, pi)
        xn = x - v
        rn, _, _, _ = bo.ba_eval(xn, C, ci, pi, sc.obs, k3, w2, want_jac=False)
        ln = 0.5 * rn @ rn
        rho = (loss - ln) / (0.5 * (-v) @ (mu * (-v) - g))
        if rho > 0:
            x = xn; r, _, a, b = bo.ba_eval(x, C, ci, pi, sc.obs, k3, w2)
            J = bo.ba_jacobian_sparse(a, b, C, P, ci, pi); g = J.T @ r
            mu *= max(1 / 3, 1 - (2 * rho - 1) ** 3); nu = 2.0; loss = ln
        else:
            mu *= nu; nu *= 2
        if np.linalg.norm(g) <= 1e-6:
            break


if __name__ == "__main__":
    main(*(sys.argv[1:2] or ["ladybug49"]))
