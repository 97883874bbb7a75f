"""The per-row arithmetic of the CUDA source (`csrc/common.cuh`: camera table, projection, Jacobian row, evaluator
losses, 8x8 Cholesky) compiled for the host with g++ (tests/host_math/device_math_host.cpp) and checked against the
oracle — so a change to those expressions is caught by `-m "not gpu"` before any GPU time is spent.  Test scaffolding:
the product never loads this library and has no CPU path (test_abi_cpu.py::test_no_cpu_fallback)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import b200ba
from oracle import ba_oracle as bo
from oracle import pose_oracle as po

HERE = os.path.dirname(os.path.abspath(__file__))
dp = C.POINTER(C.c_double)


def _p(a):
    return a.ctypes.data_as(dp)


@pytest.fixture(scope="module")
def hm(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("host_math") / "libdevice_math_host.so")
    src = os.path.join(HERE, "host_math", "device_math_host.cpp")
    r = subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-o", so, src], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    lib = C.CDLL(so)
    lib.hm_cam_table.argtypes = [dp, dp, dp]
    lib.hm_ba_row.argtypes = [dp, dp, dp, C.c_double, dp, dp, dp, dp]
    lib.hm_pose_row.argtypes = [dp, dp, dp, C.c_double, C.c_double, C.c_int, dp, dp]
    lib.hm_ba_row_v2.argtypes = [dp, dp, dp, C.c_double, dp, dp, dp]
    lib.hm_ba_row_robust.argtypes = [dp, dp, dp, C.c_double, C.c_double, C.c_int, C.c_int, dp, dp, dp, dp]
    lib.hm_reproj_loss.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double]
    lib.hm_reproj_loss.restype = C.c_double
    lib.hm_chol8_solve.argtypes = [dp, dp]
    assert lib.hm_tab_stride() == 32
    return lib


def _tables(lib, cam8, k3):
    T = np.zeros((len(cam8), 32))
    for c in range(len(cam8)):
        lib.hm_cam_table(_p(np.ascontiguousarray(cam8[c])), _p(np.ascontiguousarray(k3[c])), _p(T[c]))
    return T


def test_ba_rows_of_the_cuda_source_vs_oracle(hm):
    sc = b200ba.synth.make_ba_scene("small", seed=0)
    x0, k3 = bo.pack_params(sc.poses0, sc.intrinsics0, sc.points0)
    C_ = sc.C
    cam8 = np.concatenate([x0[:6 * C_].reshape(C_, 6), x0[6 * C_:8 * C_].reshape(C_, 2)], axis=1)
    T = _tables(hm, cam8, k3)
    np.testing.assert_allclose(T[:, :9].reshape(C_, 3, 3), bo.rot_mat_from_vec(cam8[:, :3]), atol=1e-15)
    np.testing.assert_allclose(T[:, 12:21].reshape(C_, 3, 3), bo.rot_jac_factor(cam8[:, :3], bo.rot_mat_from_vec(cam8[:, :3])), atol=1e-13)
    w2 = sc.avg_cam_width ** 2
    ro, dxdy_o, ao, bo_ = bo.ba_eval(x0, C_, sc.cam_idx, sc.pt_idx, sc.obs, k3, w2)
    r, dxdy, a, b = np.zeros(sc.N), np.zeros((sc.N, 2)), np.zeros((sc.N, 8)), np.zeros((sc.N, 3))
    for i in range(sc.N):
        X = np.ascontiguousarray(sc.points0[sc.pt_idx[i]])
        hm.hm_ba_row(_p(T[sc.cam_idx[i]]), _p(X), _p(np.ascontiguousarray(sc.obs[i])), 1.0 / w2, _p(r[i:i + 1]), _p(dxdy[i]), _p(a[i]), _p(b[i]))
    np.testing.assert_allclose(r, ro, rtol=1e-11)
    np.testing.assert_allclose(dxdy, dxdy_o, rtol=1e-9, atol=1e-10)
    np.testing.assert_allclose(a, ao, rtol=1e-9, atol=1e-12 * np.abs(ao).max())
    np.testing.assert_allclose(b, bo_, rtol=1e-9, atol=1e-12 * np.abs(bo_).max())
    # the evaluator's per-row values from the same dx, dy, against the vectors computed by the REFERENCE's own function
    gold = np.load(os.path.join(HERE, "golden", "reproj_small_seed0.npz"))
    for code, name in ((0, "trivial"), (1, "cauchy")):
        e = np.array([hm.hm_reproj_loss(code, dx, dy, 1.0) for dx, dy in dxdy])
        np.testing.assert_allclose(e, gold[name], rtol=1e-11)
    for code, name, scale in ((2, "ceres_cauchy", 1.0), (2, "ceres_cauchy", 3.5), (3, "ceres_soft_l1", 2.0)):
        e = np.array([hm.hm_reproj_loss(code, dx, dy, scale * scale) for dx, dy in dxdy])
        eo, _, _ = bo.reprojection_errors(x0, C_, sc.cam_idx, sc.pt_idx, sc.obs, k3, name, scale)
        np.testing.assert_allclose(e, eo, rtol=1e-11)


@pytest.mark.parametrize("loss", [po.L2, po.CAUCHY])
def test_pose_rows_of_the_cuda_source_vs_oracle(hm, loss):
    pb = b200ba.synth.make_pose_batch(B=3, npts=16, seed=4, ragged=True)
    q0 = np.concatenate([bo.pose_mat_to_vec(pb.poses0), pb.intrinsics0[:, :2]], axis=1)
    k3 = np.ascontiguousarray(pb.intrinsics0[:, 2:])
    w2 = pb.avg_cam_width ** 2
    ro, Jo = po.pose_eval(q0, pb.points, pb.obs, k3, pb.mask, w2, loss)
    T = _tables(hm, q0, k3)
    for bi in range(pb.B):
        for n in range(pb.points.shape[1]):
            r, a = np.zeros(1), np.zeros(8)
            hm.hm_pose_row(_p(T[bi]), _p(np.ascontiguousarray(pb.points[bi, n])), _p(np.ascontiguousarray(pb.obs[bi, n])),
                           float(pb.mask[bi, n]), 1.0 / w2, int(loss), _p(r), _p(a))
            assert r[0] == pytest.approx(ro[bi, n], rel=1e-11, abs=1e-22)
            np.testing.assert_allclose(a, Jo[bi, n], rtol=1e-9, atol=1e-12 * np.abs(Jo[bi]).max())


def test_chol8_of_the_cuda_source(hm):
    rng = np.random.default_rng(0)
    for _ in range(20):
        M = rng.normal(size=(8, 8))
        A = M @ M.T + 1e-3 * np.eye(8)
        rhs = rng.normal(size=8)
        Aw, x = A.copy(), rhs.copy()
        hm.hm_chol8_solve(_p(Aw), _p(x))
        np.testing.assert_allclose(x, np.linalg.solve(A, rhs), rtol=1e-9)
    Aw, x = -np.eye(8), np.ones(8)   # non-positive pivot: NaN, like LAPACK posv through jax
    hm.hm_chol8_solve(_p(Aw), _p(x))
    assert np.isnan(x).all()


def test_vector2_rows_of_the_cuda_source_vs_oracle(hm):
    """k_resjac_v2's row arithmetic on the (x, NaN) / (NaN, y) row contract."""
    sc = b200ba.synth.make_ba_scene("small", seed=0)
    x0, k3 = bo.pack_params(sc.poses0, sc.intrinsics0, sc.points0)
    C_ = sc.C
    cam8 = np.concatenate([x0[:6 * C_].reshape(C_, 6), x0[6 * C_:8 * C_].reshape(C_, 2)], axis=1)
    T = _tables(hm, cam8, k3)
    c2, p2, o2 = bo.expand_vector2(sc.cam_idx, sc.pt_idx, sc.obs)
    w = sc.avg_cam_width
    ro, _, ao, bo_ = bo.ba_eval(x0, C_, c2, p2, o2, k3, w * w, model="vector2")
    n = len(c2)
    r, a, b = np.zeros(n), np.zeros((n, 8)), np.zeros((n, 3))
    for i in range(n):
        hm.hm_ba_row_v2(_p(T[c2[i]]), _p(np.ascontiguousarray(sc.points0[p2[i]])), _p(np.ascontiguousarray(o2[i])), 1.0 / w,
                        _p(r[i:i + 1]), _p(a[i]), _p(b[i]))
    assert np.isfinite(r).all() and np.isfinite(a).all() and np.isfinite(b).all()
    np.testing.assert_allclose(r, ro, rtol=1e-9, atol=1e-15)
    np.testing.assert_allclose(a, ao, rtol=1e-9, atol=1e-12 * np.abs(ao).max())
    np.testing.assert_allclose(b, bo_, rtol=1e-9, atol=1e-12 * np.abs(bo_).max())


@pytest.mark.parametrize("loss,code", [("cauchy", 1), ("soft_l1", 2)])
def test_robust_rows_of_the_cuda_source_vs_oracle(hm, loss, code):
    """`robust_residual<2 | 3>` (common.cuh: the 2-vector model under Ceres' Cauchy / soft-L1 loss as k_resjac<JT, 2 | 3, .>
    evaluates it) against the oracle's restatement of Ceres' loss functions + corrector (`ba_eval_robust`): corrected residuals,
    the rows' cost terms and the corrected Jacobian rows, on the paired-row contract (component = parity of the row)."""
    sc = b200ba.synth.make_ba_scene("small", seed=0)
    x0, k3 = bo.pack_params(sc.poses0, sc.intrinsics0, sc.points0)
    C_ = sc.C
    cam8 = np.concatenate([x0[:6 * C_].reshape(C_, 6), x0[6 * C_:8 * C_].reshape(C_, 2)], axis=1)
    T = _tables(hm, cam8, k3)
    c2, p2, o2 = bo.expand_vector2_pairs(sc.cam_idx, sc.pt_idx, sc.obs)
    w, scale = sc.avg_cam_width, 2.0
    ro, cto, ao, bo_ = bo.ba_eval_robust(x0, C_, c2, p2, o2, k3, w * w, loss, scale)
    n = len(c2)
    r, ct, a, b = np.zeros(n), np.zeros(n), np.zeros((n, 8)), np.zeros((n, 3))
    for i in range(n):
        hm.hm_ba_row_robust(_p(T[c2[i]]), _p(np.ascontiguousarray(sc.points0[p2[i]])), _p(np.ascontiguousarray(o2[i])), 1.0 / w,
                            (scale / w) ** 2, code, i & 1, _p(r[i:i + 1]), _p(ct[i:i + 1]), _p(a[i]), _p(b[i]))
    np.testing.assert_allclose(r, ro, rtol=1e-9, atol=1e-15)
    np.testing.assert_allclose(ct, cto, rtol=1e-12)
    np.testing.assert_allclose(a, ao, rtol=1e-9, atol=1e-12 * np.abs(ao).max())
    np.testing.assert_allclose(b, bo_, rtol=1e-9, atol=1e-12 * np.abs(bo_).max())
