"""K3e — dense reduced-camera path (ba_dense.cuh): pair-list assembly of S, DMMA tile Cholesky, triangular solves, selected
for 8C <= 2560 on one rank.  Parity against the oracle's exact Schur/Cholesky solve (`_solve_schur_exact`, 1e-9) and the
LM trajectory goldens; the PCG path must give the same trajectory."""
import ctypes
import os

import numpy as np
import pytest
import torch

import b200ba
from oracle import ba_oracle as bo

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _plan_info(prob):
    info = (ctypes.c_int * 8)()
    prob.lib.b200ba_plan_info(prob.plan, info)
    return list(info)


@pytest.mark.parametrize("jac_f32", [False, True])
@pytest.mark.parametrize("shape", ["tiny", "small", "ladybug49", "trafalgar257"])
def test_dense_step_vs_exact_oracle(monkeypatch, shape, jac_f32):
    """One damped solve: v from the dense GPU path against the oracle's dense Cholesky on the SAME rows."""
    monkeypatch.setenv("B200BA_DENSE", "1")
    sc = b200ba.synth.make_ba_scene(shape, seed=0)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, jac_f32=jac_f32)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
    assert _plan_info(prob)[5] == 1, "dense path not selected"
    K = b200ba.BAKernels(prob, sc.avg_cam_width)
    xd = torch.from_numpy(x0).cuda()
    cam8 = torch.cat([xd[:6 * sc.C].view(sc.C, 6), xd[6 * sc.C:8 * sc.C].view(sc.C, 2)], 1).contiguous()
    U, gc, V, gp, (loss, g2, maxdiag) = K.linearize(cam8, torch.from_numpy(k3).cuda(), xd[8 * sc.C:].view(-1, 3).contiguous())
    ci, pi = sc.cam_idx.astype(np.int64), sc.pt_idx.astype(np.int64)
    r, _, a, b = bo.ba_eval(x0, sc.C, ci, pi, sc.obs, k3, sc.avg_cam_width ** 2)
    if jac_f32:  # the operator is the one of the STORED rows; the gradient too (k_resjac)
        a, b = a.astype(np.float32).astype(np.float64), b.astype(np.float32).astype(np.float64)
    gco = np.zeros((sc.C, 8)); np.add.at(gco, ci, a * r[:, None])
    gpo = np.zeros((sc.P, 3)); np.add.at(gpo, pi, b * r[:, None])
    mu = 1e-6 * maxdiag
    g = np.concatenate([gco[:, :6].reshape(-1), gco[:, 6:].reshape(-1), gpo.reshape(-1)])
    v = bo._solve_schur_exact(a, b, g, mu, sc.C, sc.P, ci, pi)
    K.schur_prepare(mu)
    vc, vp, iters = K.schur_solve()
    assert iters == 0  # direct solve
    vco = np.concatenate([v[:6 * sc.C].reshape(sc.C, 6), v[6 * sc.C:8 * sc.C].reshape(sc.C, 2)], 1)
    vpo = v[8 * sc.C:].reshape(sc.P, 3)
    # two fp64 Cholesky solves of the same system: ~1e-10 relative at cond(S) ~ 1e5 (small ... trafalgar), 5e-7 on the
    # 3-camera `tiny` scene whose reduced system is far worse conditioned
    tol = 2e-6 if shape == "tiny" else 1e-8
    assert np.linalg.norm(vc.cpu().numpy() - vco) <= tol * np.linalg.norm(vco)
    assert np.linalg.norm(vp.cpu().numpy() - vpo) <= tol * np.linalg.norm(vpo)


@pytest.mark.parametrize("shape", ["small", "ladybug49", "trafalgar257"])
def test_dense_lm_trajectory_golden_and_equals_pcg_path(monkeypatch, shape):
    gold = np.load(os.path.join(GOLD, f"ba_{shape}_seed0.npz"))
    sc = b200ba.synth.make_ba_scene(shape, seed=0)
    out = []
    for dense in ("1", "0"):
        monkeypatch.setenv("B200BA_DENSE", dense)
        ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
        x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
        prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
        assert _plan_info(prob)[5] == int(dense)
        params, st = ba.optimize(x0, prob, None, k3, None)
        assert st.iter_num == int(gold["iter_num"]) and st.status == 0
        np.testing.assert_allclose(st.loss_trajectory, gold["loss"], rtol=1e-5)
        rms = bo.rms_reprojection_px(params, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3)
        assert rms == pytest.approx(float(gold["rms_px"]), rel=1e-4)
        if dense == "1":
            assert st.pcg_iters.sum() == 0
        out.append(st.loss_trajectory)
    np.testing.assert_allclose(out[0], out[1], rtol=1e-5)


def test_dense_reference_benchmark_size_15x400(monkeypatch):
    """The reference's own BA benchmark size (benchmark_impl_bundle_adjustment.py:28-32: 15 cameras, <= 400 points) through
    the dense-padded facade, against the exact oracle run on the same scene."""
    monkeypatch.setenv("B200BA_DENSE", "1")
    sc = b200ba.synth.make_ba_scene(C=15, P=400, N=2000, seed=4)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    p2d, idx, mask = sc.to_dense_padded()
    params, st = ba.optimize(x0, p2d, idx, k3, mask)
    ref = bo.lm_solve(x0, sc.C, sc.P, sc.cam_idx, sc.pt_idx, sc.obs, k3, sc.avg_cam_width)
    assert st.iter_num == ref.iter_num
    np.testing.assert_allclose(st.loss_trajectory, np.array(ref.trace.loss), rtol=1e-5)


def test_dense_with_a_camera_seeing_a_point_twice(monkeypatch):
    """Duplicate (camera, point) rows (real COLMAP models have them): both orders of such a pair
    belong to the diagonal block of S."""
    monkeypatch.setenv("B200BA_DENSE", "1")
    sc = b200ba.synth.make_ba_scene("small", seed=0)
    dup = np.arange(0, sc.N, 37)
    rng = np.random.default_rng(0)
    ci = np.concatenate([sc.cam_idx, sc.cam_idx[dup]])
    pi = np.concatenate([sc.pt_idx, sc.pt_idx[dup]])
    ob = np.concatenate([sc.obs, (sc.obs[dup] + rng.normal(size=(len(dup), 2))).astype(np.float32).astype(np.float64)])
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, jac_f32=False)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    params, st = ba.optimize(x0, ba.build_problem_sparse(ci, pi, ob, sc.P), None, k3, None)
    ref = bo.lm_solve(x0, sc.C, sc.P, ci, pi, ob, k3, sc.avg_cam_width)
    assert st.iter_num == ref.iter_num
    np.testing.assert_allclose(st.loss_trajectory, np.array(ref.trace.loss), rtol=1e-5)


@pytest.mark.parametrize("shape", ["small", "ladybug49"])
def test_device_resident_lm_loop_equals_host_driven_loop(monkeypatch, shape):
    """The dense path runs the whole LM loop as ONE CUDA graph (WHILE / IF conditional nodes, conditions set by kernels; no
    host synchronisation inside the loop).  Same kernels, same order: it must reproduce the host-driven loop bit for bit —
    iteration count, trajectories, accept pattern, parameters — and solving twice from the same plan must replay cleanly."""
    monkeypatch.setenv("B200BA_DENSE", "1")
    sc = b200ba.synth.make_ba_scene(shape, seed=0)
    out = []
    for graph in ("1", "0"):
        monkeypatch.setenv("B200BA_GRAPH", graph)
        ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
        x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
        prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
        n0 = b200ba.launch_count()
        params, st = ba.optimize(x0, prob, None, k3, None)
        launches = b200ba.launch_count() - n0
        params2, st2 = ba.optimize(x0, prob, None, k3, None)   # replay of the same graph / same plan
        np.testing.assert_array_equal(params, params2)
        assert st.iter_num == st2.iter_num
        out.append((params, st, launches))
    (pa, sa, la), (pb, sb, lb) = out
    assert sa.iter_num == sb.iter_num and sa.status == sb.status == 0
    np.testing.assert_array_equal(sa.loss_trajectory, sb.loss_trajectory)
    np.testing.assert_array_equal(sa.grad_trajectory, sb.grad_trajectory)
    np.testing.assert_array_equal(sa.mu_trajectory, sb.mu_trajectory)
    np.testing.assert_array_equal(sa.accepted, sb.accepted)
    np.testing.assert_array_equal(pa, pb)
    assert sa.n_jac_evals == sb.n_jac_evals and sa.n_res_evals == sb.n_res_evals
    assert la - lb == 1 + sa.iter_num   # the same kernels, plus the graph's bookkeeping: 1 at init, 1 per iteration (the WHILE condition; the IF condition rides on the accept copy)


def test_device_resident_lm_loop_respects_maxiter(monkeypatch):
    monkeypatch.setenv("B200BA_DENSE", "1")
    sc = b200ba.synth.make_ba_scene("small", seed=0)
    for mi in (1, 3):
        ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, maxiter=mi)
        x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
        params, st = ba.optimize(x0, ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None)
        assert st.iter_num == mi and (st.status & 1) and len(st.loss_trajectory) == mi + 1
