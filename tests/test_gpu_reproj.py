"""GPU parity of the reprojection-error evaluator (`b200ba_reprojection_errors`) against the
oracle and against tests/golden/reproj_small_seed0.npz — vectors computed by the reference's own
`Camera.compute_projection_errors_alt` (tests/golden/make_reproj_golden.py).  fp64 throughout: 1e-9 relative (observed on the host-compiled CUDA arithmetic: ~1e-13)."""
import os

import numpy as np
import pytest
import torch

import b200ba
from oracle import ba_oracle as bo

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reproj_small_seed0.npz")


def _dev_params(x0, k3, C):
    xd = torch.from_numpy(x0).cuda()
    cam8 = torch.cat([xd[:6 * C].view(C, 6), xd[6 * C:8 * C].view(C, 2)], 1).contiguous()
    return cam8, torch.from_numpy(k3).cuda(), xd[8 * C:].view(-1, 3).contiguous()


@pytest.mark.parametrize("loss", ["trivial", "cauchy"])
def test_small_scene_vs_reference_golden_and_oracle(loss):
    sc = b200ba.synth.make_ba_scene("small", seed=0)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
    code = b200ba.LossFunction.CAUCHY if loss == "cauchy" else b200ba.LossFunction.TRIVIAL
    err, sums = ba.compute_reprojection_errors_alt(x0, k3, prob, code)
    gold = np.load(GOLD)[loss]
    np.testing.assert_allclose(err, gold, rtol=1e-9)  # the reference itself, caller's row order
    eo, so, _ = bo.reprojection_errors(x0, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3, loss)
    np.testing.assert_allclose(err, eo, rtol=1e-9)
    np.testing.assert_allclose(sums, so, rtol=1e-9)


@pytest.mark.parametrize("shape", ["tiny", "ladybug49"])
def test_shapes_and_row_order_independence(shape):
    sc = b200ba.synth.make_ba_scene(shape, seed=1)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    cam8, k3d, pts = _dev_params(x0, k3, sc.C)
    prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
    K = b200ba.BAKernels(prob, sc.avg_cam_width)
    for loss, code, scale in (("trivial", 0, 1.0), ("cauchy", 1, 1.0), ("ceres_cauchy", 2, 1.0), ("ceres_cauchy", 2, 3.5),
                              ("ceres_soft_l1", 3, 2.0)):
        err, sums = K.reprojection_errors(cam8, k3d, pts, code, scale)
        eo, so, _ = bo.reprojection_errors(x0, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3, loss, scale)
        np.testing.assert_allclose(err.cpu().numpy(), eo, rtol=1e-9)
        np.testing.assert_allclose(sums.cpu().numpy(), so, rtol=1e-9)
    # trivial loss is the BA residual times w^2 (bundle_adjustment.py:74 divides the same squared error by w^2)
    r, _, _, _ = K.residual_jacobian(cam8, k3d, pts)
    e0, s0 = K.reprojection_errors(cam8, k3d, pts, 0)
    np.testing.assert_allclose(e0.cpu().numpy(), r.cpu().numpy() * sc.avg_cam_width ** 2, rtol=1e-12)
    # a permutation of the caller's rows permutes the per-row output and leaves the per-camera sums bit-identical
    perm = np.random.default_rng(3).permutation(sc.N)
    prob2 = ba.build_problem_sparse(sc.cam_idx[perm], sc.pt_idx[perm], sc.obs[perm], sc.P)
    e2, s2 = b200ba.BAKernels(prob2, sc.avg_cam_width).reprojection_errors(cam8, k3d, pts, 0)
    assert torch.equal(e2.cpu(), e0.cpu()[torch.from_numpy(perm)])
    assert torch.equal(s2, s0)


def test_errors_drop_after_solve_and_null_outputs():
    sc = b200ba.synth.make_ba_scene("small", seed=0)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
    e0, _ = ba.compute_reprojection_errors_alt(x0, k3, prob)
    p2d, idx, mask = sc.to_dense_padded()
    x1, st = ba.optimize(x0, p2d, idx, k3, mask)
    e1, s1 = ba.compute_reprojection_errors_alt(x1, k3, prob)
    assert e1.mean() < 0.01 * e0.mean()
    rms = bo.rms_reprojection_px(x1, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3)
    assert np.sqrt(e1.mean()) == pytest.approx(rms, rel=1e-9)
    assert s1.sum() == pytest.approx(e1.sum(), rel=1e-12)
    # either output may be NULL; a bad loss code is an error, not a silent default
    cam8, k3d, pts = _dev_params(x1, k3, sc.C)
    lib = prob.lib
    sums = torch.zeros(sc.C, dtype=torch.float64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    assert lib.b200ba_reprojection_errors(prob.plan, cam8.data_ptr(), k3d.data_ptr(), pts.data_ptr(), 0, 1.0, None, sums.data_ptr(), s) == 0
    torch.cuda.synchronize()
    np.testing.assert_array_equal(sums.cpu().numpy(), s1)
    assert lib.b200ba_reprojection_errors(prob.plan, cam8.data_ptr(), k3d.data_ptr(), pts.data_ptr(), 7, 1.0, None, sums.data_ptr(), s) != 0
    assert lib.b200ba_reprojection_errors(prob.plan, cam8.data_ptr(), k3d.data_ptr(), pts.data_ptr(), 2, 0.0, None, sums.data_ptr(), s) != 0


def test_colmap_model_file_to_solution_file(tmp_path):
    """File -> arrays -> solve -> evaluate -> export -> file (SURVEY.md §8 f1-f3 chained): the solve from a COLMAP model on
    disk equals the solve from the in-memory scene it was written from, and the exported model reads back as the solution."""
    from b200ba import model_io as mio
    sc = b200ba.synth.make_ba_scene("small", seed=0)
    mio.write_colmap_model(mio.model_from_ba_arrays(sc.poses0, sc.intrinsics0, sc.points0, sc.cam_idx, sc.pt_idx, sc.obs), tmp_path / "in")
    m = mio.read_colmap_model(tmp_path / "in")
    a = m.to_ba_arrays()
    # PINHOLE cameras carry no skew: compare with the scene at skew 0
    intr = sc.intrinsics0.copy()
    intr[:, 4] = 0.0
    ba = b200ba.B200BundleAdjustment(sc.C, a["avg_cam_width"])
    x0, k3 = ba.prepare_params(a["poses0"], a["intrinsics0"], a["points0"])
    prob = ba.build_problem_sparse(a["cam_idx"], a["pt_idx"], a["obs"], len(a["points0"]))
    x, st = ba.optimize(x0, prob, cx_cy_skew=k3)
    xr0, k3r = ba.prepare_params(sc.poses0, intr, sc.points0)
    ref = bo.lm_solve(xr0, sc.C, sc.P, sc.cam_idx, sc.pt_idx, sc.obs, k3r, sc.avg_cam_width)
    assert st.iter_num == ref.iter_num
    np.testing.assert_allclose(st.loss_trajectory, np.array(ref.trace.loss), rtol=1e-5)
    err, per_cam = ba.compute_reprojection_errors_alt(x, k3, prob)
    assert np.sqrt(err.mean()) == pytest.approx(bo.rms_reprojection_px(ref.params, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3r), rel=1e-4)
    C = sc.C
    poses, intr_out, pts_out = b200ba.unpack_params(x, k3, C)
    f = intr_out[:, :2]
    out = m.with_solution(poses, intr_out, pts_out, a["point_ids"])
    mio.write_colmap_model(out, tmp_path / "out")
    b = mio.read_colmap_model(tmp_path / "out").to_ba_arrays()
    np.testing.assert_allclose(b["poses0"], poses, atol=1e-12)
    np.testing.assert_array_equal(b["points0"], x[8 * C:].reshape(-1, 3))
    np.testing.assert_allclose(b["intrinsics0"][:, :2], f)
