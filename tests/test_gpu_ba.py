"""GPU parity tests proper: CUDA path (through the C-ABI) vs the float64 oracle on the same
seeded inputs.  Tolerances: per-iteration cost 1e-5 rel, final RMS 1e-4 rel (north_star)."""
import numpy as np
import pytest
import torch

import b200ba
from oracle import ba_oracle as bo

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _pcg_path(monkeypatch):
    """These tests exercise the PCG machinery (implicit Schur matvec, ring / warp-task / chunk / tile kernels) on shapes
    small enough for the dense reduced-camera path, which would otherwise be selected (tests/test_gpu_dense.py covers it)."""
    monkeypatch.setenv("B200BA_DENSE", "0")


def _setup(shape, seed=0, jac_f32=False, **kw):  # building-block tests use fp64 rows (tight row tolerances)
    sc = b200ba.synth.make_ba_scene(shape, seed=seed, **kw)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, jac_f32=jac_f32)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    return sc, ba, x0, k3


def _dev_params(x0, k3, C):
    xd = torch.from_numpy(x0).cuda()
    cam8 = torch.cat([xd[:6 * C].view(C, 6), xd[6 * C:8 * C].view(C, 2)], 1).contiguous()
    pts = xd[8 * C:].view(-1, 3).contiguous()
    return cam8, torch.from_numpy(k3).cuda(), pts


@pytest.mark.parametrize("shape", ["tiny", "small", "ladybug49"])
def test_residual_jacobian_rows(shape):
    sc, ba, x0, k3 = _setup(shape)
    prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
    K = b200ba.BAKernels(prob, sc.avg_cam_width)
    cam8, k3d, pts = _dev_params(x0, k3, sc.C)
    r, Jc, Jp, cost = K.residual_jacobian(cam8, k3d, pts)
    ro, _, a, b = bo.ba_eval(x0, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3, sc.avg_cam_width ** 2)
    np.testing.assert_allclose(r.cpu().numpy(), ro, rtol=1e-10, atol=1e-18)
    np.testing.assert_allclose(Jc.cpu().numpy(), a, rtol=1e-9, atol=1e-12 * np.abs(a).max())
    np.testing.assert_allclose(Jp.cpu().numpy(), b, rtol=1e-9, atol=1e-12 * np.abs(b).max())
    assert cost == pytest.approx(0.5 * float(ro @ ro), rel=1e-12)


@pytest.mark.parametrize("shape", ["tiny", "small", "ladybug49"])
def test_schur_blocks_and_step(shape):
    sc, ba, x0, k3 = _setup(shape)
    prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
    K = b200ba.BAKernels(prob, sc.avg_cam_width)
    cam8, k3d, pts = _dev_params(x0, k3, sc.C)
    U, gc, V, gp, (loss, g2, maxdiag) = K.linearize(cam8, k3d, pts)
    ci, pi = sc.cam_idx.astype(np.int64), sc.pt_idx.astype(np.int64)
    r, _, a, b = bo.ba_eval(x0, sc.C, ci, pi, sc.obs, k3, sc.avg_cam_width ** 2)
    Uo = np.zeros((sc.C, 8, 8)); np.add.at(Uo, ci, a[:, :, None] * a[:, None, :])
    Vo = np.zeros((sc.P, 3, 3)); np.add.at(Vo, pi, b[:, :, None] * b[:, None, :])
    gco = np.zeros((sc.C, 8)); np.add.at(gco, ci, a * r[:, None])
    gpo = np.zeros((sc.P, 3)); np.add.at(gpo, pi, b * r[:, None])
    np.testing.assert_allclose(U.cpu().numpy(), Uo, rtol=1e-9, atol=1e-12 * np.abs(Uo).max())
    Vf = V.cpu().numpy()
    np.testing.assert_allclose(Vf, Vo[:, [0, 0, 0, 1, 1, 2], [0, 1, 2, 1, 2, 2]], rtol=1e-9, atol=1e-12 * np.abs(Vo).max())
    np.testing.assert_allclose(gc.cpu().numpy(), gco, rtol=1e-9, atol=1e-12 * np.abs(gco).max())
    np.testing.assert_allclose(gp.cpu().numpy(), gpo, rtol=1e-9, atol=1e-12 * np.abs(gpo).max())
    md = max(np.einsum("cii->ci", Uo).max(), np.einsum("pii->pi", Vo).max())
    assert maxdiag == pytest.approx(md, rel=1e-12)
    assert g2 == pytest.approx((gco ** 2).sum() + (gpo ** 2).sum(), rel=1e-10)
    # one damped solve vs the exact Schur/Cholesky oracle solve
    mu = 1e-6 * md
    g = np.concatenate([gco[:, :6].reshape(-1), gco[:, 6:].reshape(-1), gpo.reshape(-1)])
    v = bo._solve_schur_exact(a, b, g, mu, sc.C, sc.P, ci, pi)
    K.schur_prepare(mu)
    vc, vp, iters = K.schur_solve(rtol=1e-13, maxiter=2000)
    vco = np.concatenate([v[:6 * sc.C].reshape(sc.C, 6), v[6 * sc.C:8 * sc.C].reshape(sc.C, 2)], 1)
    vpo = v[8 * sc.C:].reshape(sc.P, 3)
    assert iters < 2000
    assert np.linalg.norm(vc.cpu().numpy() - vco) <= 1e-5 * np.linalg.norm(vco)
    assert np.linalg.norm(vp.cpu().numpy() - vpo) <= 1e-5 * np.linalg.norm(vpo)


@pytest.mark.parametrize("jac_f32", [True, False])
@pytest.mark.parametrize("shape", ["small", "ladybug49"])
def test_lm_trajectory_vs_exact_oracle(shape, jac_f32):
    """default build (fp32-stored rows, fp64 accumulation) and the fp64-row build, both within the
    north_star tolerances: per-iteration cost 1e-5 rel, final RMS 1e-4 rel."""
    sc, ba, x0, k3 = _setup(shape, jac_f32=jac_f32)
    params, st = ba.optimize(x0, ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None)
    ref = bo.lm_solve(x0, sc.C, sc.P, sc.cam_idx, sc.pt_idx, sc.obs, k3, sc.avg_cam_width)
    assert st.iter_num == ref.iter_num
    np.testing.assert_allclose(st.loss_trajectory, np.array(ref.trace.loss), rtol=1e-5)
    assert list(st.accepted.astype(bool)) == ref.trace.accepted
    rms = bo.rms_reprojection_px(params, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3)
    rms_ref = bo.rms_reprojection_px(ref.params, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3)
    assert rms == pytest.approx(rms_ref, rel=1e-4)
    assert not np.isnan(params).any()


def test_dense_padded_contract_equals_sparse():
    sc, ba, x0, k3 = _setup("small")
    p2d, idx, mask = sc.to_dense_padded()
    pa, sa = ba.optimize(x0, p2d, idx, k3, mask)
    pb, sb = ba.optimize(x0, ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None)
    assert sa.iter_num == sb.iter_num
    np.testing.assert_array_equal(pa, pb)  # row order must not matter: bit-identical
    np.testing.assert_array_equal(sa.loss_trajectory, sb.loss_trajectory)


def test_trafalgar_config_trajectory_golden():
    """BASELINE config 3 (Trafalgar-257 shape) against the frozen exact-oracle trajectory
    (tests/golden/ba_trafalgar257_seed0.npz; the oracle solve takes ~10 s so it is not re-run here)."""
    import os
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "ba_trafalgar257_seed0.npz"))
    sc = b200ba.synth.make_ba_scene("trafalgar257", seed=0)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    np.testing.assert_array_equal(x0[:64], gold["x0"])
    params, st = ba.optimize(x0, ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None)
    assert st.iter_num == int(gold["iter_num"])
    np.testing.assert_allclose(st.loss_trajectory, gold["loss"], rtol=1e-5)
    rms = bo.rms_reprojection_px(params, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3)
    assert rms == pytest.approx(float(gold["rms_px"]), rel=1e-4)


def test_venice_config_trajectory_golden():
    """The BENCHED config (Venice-1778: 1 778 cams / 993 923 pts / 5 001 946 obs) against the frozen EXACT-oracle
    trajectory (tests/golden/ba_venice1778_seed0.npz: LM state machine of oracle/ba_oracle.py, reduced camera system
    formed densely in C and factored by LAPACK — tests/golden/make_venice_golden.py, ~3 min of host time, not re-run
    here).  north_star: per-iteration cost within 1e-5, final RMS reprojection error within 1e-4."""
    import os
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "ba_venice1778_seed0.npz"))
    sc = b200ba.synth.make_ba_scene("venice1778", seed=0)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    np.testing.assert_array_equal(x0[:64], gold["x0"])
    params, st = ba.optimize(x0, ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None)
    assert st.iter_num == int(gold["iter_num"])
    np.testing.assert_allclose(st.loss_trajectory, gold["loss"], rtol=1e-5)
    assert list(st.accepted.astype(bool)) == list(gold["accepted"])
    rms = bo.rms_reprojection_px(params, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3)
    assert rms == pytest.approx(float(gold["rms_px"]), rel=1e-4)


def test_tiled_and_chunk_point_kernels_agree(monkeypatch):
    """The point side of the Schur matvec exists twice (ba_ps_tile.cuh: exchange through shared memory in tile-local B
    order; ba_pcg.cuh: chunk kernel with gathers through the L2 window).  Same row arithmetic in the same order: the
    trajectories must agree with each other far inside the parity bar, and with the exact oracle to 1e-5.  1 024-row tiles put
    many tiles (and tile-straddling camera runs) into a Trafalgar-sized problem."""
    import os
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "ba_trafalgar257_seed0.npz"))
    sc = b200ba.synth.make_ba_scene("trafalgar257", seed=0)
    out = []
    for tile in ("0", "1"):
        monkeypatch.setenv("B200BA_PS_TILE", tile)
        monkeypatch.setenv("B200BA_TILE_ROWS", "1024")
        ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
        info_expected = tile == "1" 
        x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
        prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
        info = (__import__("ctypes").c_int * 8)()
        prob.lib.b200ba_plan_info(prob.plan, info)
        assert bool(info[0]) == info_expected and (not info_expected or info[2] > 100)   # really the tiled kernel, many tiles
        params, st = ba.optimize(x0, prob, None, k3, None)
        assert st.iter_num == int(gold["iter_num"])
        np.testing.assert_allclose(st.loss_trajectory, gold["loss"], rtol=1e-5)
        out.append((params, st.loss_trajectory))
    # same row arithmetic in the same order; the only difference is the PCG scalar reduction over the cameras (128 threads in
    # the chunk kernel, 256 in the tiled one), i.e. last-bit differences of alpha / beta: measured 4e-7 on the trajectory
    np.testing.assert_allclose(out[0][1], out[1][1], rtol=2e-6)
    np.testing.assert_allclose(out[0][0], out[1][0], rtol=0, atol=1e-6 * np.abs(out[0][0]).max())


def test_fp32_rows_track_fp64_rows_at_scale():
    """Size-independent check at a shape the exact oracle cannot reach in seconds: the default
    fp32-row build follows the (oracle-validated) fp64-row build within 1e-5 per iteration."""
    sc = b200ba.synth.make_ba_scene(C=600, P=150000, N=750000, seed=5)
    out = []
    for f32 in (True, False):
        ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, jac_f32=f32, pcg_rtol=1e-10 if f32 else 1e-12)
        x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
        out.append(ba.optimize(x0, ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None))
    (pa, sa), (pb, sb) = out
    assert sa.iter_num == sb.iter_num
    np.testing.assert_allclose(sa.loss_trajectory, sb.loss_trajectory, rtol=1e-5)
    ra = bo.rms_reprojection_px(pa, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3)
    rb = bo.rms_reprojection_px(pb, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3)
    assert ra == pytest.approx(rb, rel=1e-4)


@pytest.mark.parametrize("jac_f32", [True, False])
@pytest.mark.parametrize("cpn", [(3, 4000, 11500), (2, 6000, 12000)])
def test_heavy_cameras_take_the_tail_and_fallback_paths(cpn, jac_f32):
    """Cameras with more rows than a thread's register stash (3072 fp32 / 1536 fp64 rows: ring tail path,
    rows re-read in pass A) and more than the whole ring holds (5376 / 2688 rows: generic camera kernel)
    must give the same trajectory as the exact oracle."""
    C, P, N = cpn
    sc = b200ba.synth.make_ba_scene(C=C, P=P, N=N, seed=2)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, jac_f32=jac_f32)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    params, st = ba.optimize(x0, ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None)
    ref = bo.lm_solve(x0, sc.C, sc.P, sc.cam_idx, sc.pt_idx, sc.obs, k3, sc.avg_cam_width)
    assert st.iter_num == ref.iter_num
    np.testing.assert_allclose(st.loss_trajectory, np.array(ref.trace.loss), rtol=1e-5)
    rms = bo.rms_reprojection_px(params, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3)
    rms_ref = bo.rms_reprojection_px(ref.params, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3)
    assert rms == pytest.approx(rms_ref, rel=1e-4)


def test_point_block_size_does_not_matter(monkeypatch):
    """The point-block size only changes WHERE rows live (run boundaries, padding) and which thread slot sums
    which row of a camera; 1 block and ~16 blocks must agree at the level a PCG solve to rtol 1e-10 pins the
    step (measured 2.5e-7), far inside the 1e-5 parity bar."""
    sc, ba, x0, k3 = _setup("ladybug49", jac_f32=True)
    out = []
    for rows in ("1000000", "2048"):
        monkeypatch.setenv("B200BA_BLOCK_ROWS", rows)
        out.append(ba.optimize(x0, ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None))
    (pa, sa), (pb, sb) = out
    assert sa.iter_num == sb.iter_num
    np.testing.assert_allclose(sa.loss_trajectory, sb.loss_trajectory, rtol=2e-6)
    np.testing.assert_allclose(pa, pb, rtol=0, atol=1e-6 * np.abs(pa).max())


@pytest.mark.parametrize("ringx_d", ["1", "3", "8"])
def test_fused_multi_rank_ring_kernel_on_one_rank(monkeypatch, ringx_d):
    """ba_cs_ringx.cuh (the multi-rank camera kernel: partial sums exchanged inside the kernel, D cameras in flight per
    CTA, pass A from the rows retained in shared memory) forced onto one rank: same trajectory as the single-rank ring
    kernel, for pipeline depths 1, 3 and the maximum, on a shape with several slots per camera (ladybug49: ~650 rows)."""
    sc, ba, x0, k3 = _setup("ladybug49", jac_f32=True)
    out = []
    for split in ("0", "1"):
        monkeypatch.setenv("B200BA_FORCE_SPLIT", split)
        monkeypatch.setenv("B200BA_RINGX", "1")
        monkeypatch.setenv("B200BA_RINGX_D", ringx_d)
        prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
        info = (__import__("ctypes").c_int * 8)()
        prob.lib.b200ba_plan_info(prob.plan, info)
        assert (info[7] & 1) == int(split) and (info[7] & 2)
        out.append(ba.optimize(x0, prob, None, k3, None))
    (pa, sa), (pb, sb) = out
    assert sa.iter_num == sb.iter_num
    np.testing.assert_allclose(sa.loss_trajectory, sb.loss_trajectory, rtol=2e-6)
    np.testing.assert_allclose(pa, pb, rtol=0, atol=1e-6 * np.abs(pa).max())


@pytest.mark.parametrize("cw_u", ["3", "4", "6"])
@pytest.mark.parametrize("shape", ["ladybug49", "trafalgar257"])
def test_warp_per_camera_fused_multi_rank_kernel_on_one_rank(monkeypatch, shape, cw_u):
    """ba_cs_warpx.cuh (one cooperative launch per PCG iteration on several ranks: a persistent warp per camera does pass C,
    the packet exchange, the camera-local update and pass A) forced onto one rank: same trajectory as the single-rank ring
    kernel, all three unroll instantiations; trafalgar257 has more cameras than one CTA wave of warps on a small grid."""
    sc, ba, x0, k3 = _setup(shape, jac_f32=True)
    out = []
    for split in ("0", "1"):
        monkeypatch.setenv("B200BA_FORCE_SPLIT", split)
        monkeypatch.setenv("B200BA_WARPX", "1")
        monkeypatch.setenv("B200BA_CW_U", cw_u)
        prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
        info = (__import__("ctypes").c_int * 8)()
        prob.lib.b200ba_plan_info(prob.plan, info)
        assert (info[7] & 1) == int(split) and (info[7] & 4)
        out.append(ba.optimize(x0, prob, None, k3, None))
    (pa, sa), (pb, sb) = out
    assert sa.iter_num == sb.iter_num and sa.pcg_iters.sum() > 0
    np.testing.assert_allclose(sa.loss_trajectory, sb.loss_trajectory, rtol=2e-6)
    np.testing.assert_allclose(pa, pb, rtol=0, atol=1e-6 * np.abs(pa).max())


@pytest.mark.parametrize("cw_u", ["3", "4", "6"])
@pytest.mark.parametrize("task_rows", ["64", "100000"])
def test_split_camera_passes_match_the_fused_iteration(monkeypatch, task_rows, cw_u):
    """The multi-rank kernel sequence (warp-task partial sums | all-reduce point | camera-local vector step |
    warp-task sigma pass) forced onto one rank must reproduce the fused ring-kernel iteration; 64-row tasks cut
    every camera into many parts (ordered last-arriver sum), 100000-row tasks keep cameras whole."""
    sc, ba, x0, k3 = _setup("ladybug49", jac_f32=True)
    out = []
    for split in ("0", "1"):
        monkeypatch.setenv("B200BA_FORCE_SPLIT", split)
        monkeypatch.setenv("B200BA_RINGX", "0")  # the three-launch sequence (fallback when peer memory cannot be mapped)
        monkeypatch.setenv("B200BA_TASK_ROWS", task_rows)
        monkeypatch.setenv("B200BA_CW_U", cw_u)  # all three unroll instantiations of k_cs_warp
        out.append(ba.optimize(x0, ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None))
    (pa, sa), (pb, sb) = out
    assert sa.iter_num == sb.iter_num
    np.testing.assert_allclose(sa.loss_trajectory, sb.loss_trajectory, rtol=2e-6)
    np.testing.assert_allclose(pa, pb, rtol=0, atol=1e-6 * np.abs(pa).max())


def _drop_camera_rows(sc, cam, keep_point):
    """Remove the observations of camera `cam` on the points where keep_point is False, as long as the point
    keeps >= 2 observations.  Returns filtered (cam_idx, pt_idx, obs)."""
    deg = np.bincount(sc.pt_idx, minlength=sc.P)
    drop = (sc.cam_idx == cam) & ~keep_point[sc.pt_idx] & (deg[sc.pt_idx] > 2)
    return sc.cam_idx[~drop], sc.pt_idx[~drop], sc.obs[~drop]


@pytest.mark.parametrize("split", ["0", "1"])
def test_camera_with_very_few_rows(monkeypatch, split):
    """One camera keeps only a handful of rows (empty runs in most point blocks, a nearly empty warp task / ring
    step): same trajectory as the exact oracle, on the fused and on the split camera path."""
    monkeypatch.setenv("B200BA_FORCE_SPLIT", split)
    monkeypatch.setenv("B200BA_BLOCK_ROWS", "2048")
    sc, ba, x0, k3 = _setup("ladybug49", jac_f32=True)
    keep = np.zeros(sc.P, dtype=bool)
    keep[:40] = True
    ci, pi, ob = _drop_camera_rows(sc, 3, keep)
    assert 0 < (ci == 3).sum() < 60
    params, st = ba.optimize(x0, ba.build_problem_sparse(ci, pi, ob, sc.P), None, k3, None)
    ref = bo.lm_solve(x0, sc.C, sc.P, ci, pi, ob, k3, sc.avg_cam_width)
    assert st.iter_num == ref.iter_num
    np.testing.assert_allclose(st.loss_trajectory, np.array(ref.trace.loss), rtol=1e-5)


@pytest.mark.parametrize("jac_f32", [False, True])
def test_point_with_more_rows_than_a_chunk_holds(jac_f32):
    """One point with 603 observations (its three cameras see it 201 times each: duplicates are legal rows) — more than the
    512 rows of a point chunk: a chunk of its own from the device-side chunk builder, the CTA-wide branch of the point-side
    kernel in the matvec and in the back-substitution, a 603-row run in the point-block kernels.  PCG path against the exact
    oracle on the same rows."""
    sc, ba, x0, k3 = _setup("ladybug49", jac_f32=jac_f32)
    rows0 = np.flatnonzero(sc.pt_idx == 0)
    rng = np.random.default_rng(0)
    rep = np.repeat(rows0, 600 // len(rows0) + 1)[:600]
    cam_idx = np.concatenate([sc.cam_idx, sc.cam_idx[rep]]).astype(np.int32)
    pt_idx = np.concatenate([sc.pt_idx, sc.pt_idx[rep]]).astype(np.int32)
    obs = np.concatenate([sc.obs, sc.obs[rep] + rng.normal(0.0, 0.5, (len(rep), 2))])
    perm = rng.permutation(len(cam_idx))
    cam_idx, pt_idx, obs = cam_idx[perm], pt_idx[perm], np.ascontiguousarray(obs[perm])
    assert np.bincount(pt_idx, minlength=sc.P)[0] > 512
    prob = ba.build_problem_sparse(cam_idx, pt_idx, obs, sc.P)
    info = (__import__("ctypes").c_int * 8)()
    prob.lib.b200ba_plan_info(prob.plan, info)
    assert not info[5]  # PCG path (the fixture switches the dense path off)
    params, st = ba.optimize(x0, prob, None, k3, None)
    ref = bo.lm_solve(x0, sc.C, sc.P, cam_idx, pt_idx, obs, k3, sc.avg_cam_width)
    assert st.iter_num == ref.iter_num and st.pcg_iters.sum() > 0
    np.testing.assert_allclose(st.loss_trajectory, np.array(ref.trace.loss), rtol=1e-5)
    np.testing.assert_array_equal(st.accepted.astype(bool), np.array(ref.trace.accepted))
    rms = bo.rms_reprojection_px(params, sc.C, cam_idx, pt_idx, obs, k3)
    assert rms == pytest.approx(bo.rms_reprojection_px(ref.params, sc.C, cam_idx, pt_idx, obs, k3), rel=1e-4)


@pytest.mark.parametrize("dense", ["0", "1"])
def test_points_without_observations_stay_put(monkeypatch, dense):
    """Three points (the first, one in the middle, the last) lose all their observations: empty A runs, V_p = 0 so that
    (V_p + mu I)^-1 = I / mu and g_p = 0.  Both solver paths follow the exact oracle and leave those points untouched."""
    monkeypatch.setenv("B200BA_DENSE", dense)
    sc, ba, x0, k3 = _setup("small", jac_f32=True)
    gone = [0, 7, sc.P - 1]
    keep = ~np.isin(sc.pt_idx, gone)
    ci, pi, ob = sc.cam_idx[keep], sc.pt_idx[keep], np.ascontiguousarray(sc.obs[keep])
    params, st = ba.optimize(x0, ba.build_problem_sparse(ci, pi, ob, sc.P), None, k3, None)
    ref = bo.lm_solve(x0, sc.C, sc.P, ci, pi, ob, k3, sc.avg_cam_width)
    assert st.iter_num == ref.iter_num
    np.testing.assert_allclose(st.loss_trajectory, np.array(ref.trace.loss), rtol=1e-5)
    np.testing.assert_array_equal(params[8 * sc.C:].reshape(-1, 3)[gone], sc.points0[gone])
    assert np.isfinite(params).all()


def test_schur_matvec_matches_dense_schur_complement():
    """b200ba_schur_matvec (camera head -> point kernel -> camera tail) against S formed densely by the oracle's
    formulas: S = U + mu I - W (V + mu I)^-1 W^T."""
    sc, ba, x0, k3 = _setup("small")
    prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
    K = b200ba.BAKernels(prob, sc.avg_cam_width)
    cam8, k3d, pts = _dev_params(x0, k3, sc.C)
    K.linearize(cam8, k3d, pts)
    ci, pi = sc.cam_idx.astype(np.int64), sc.pt_idx.astype(np.int64)
    r, _, a, b = bo.ba_eval(x0, sc.C, ci, pi, sc.obs, k3, sc.avg_cam_width ** 2)
    U = np.zeros((sc.C, 8, 8)); np.add.at(U, ci, a[:, :, None] * a[:, None, :])
    V = np.zeros((sc.P, 3, 3)); np.add.at(V, pi, b[:, :, None] * b[:, None, :])
    mu = 1e-6 * max(np.einsum("cii->ci", U).max(), np.einsum("pii->pi", V).max())
    Vinv = np.linalg.inv(V + mu * np.eye(3))
    rng = np.random.default_rng(0)
    x = rng.standard_normal((sc.C, 8))
    s = np.einsum("nk,nk->n", a, x[ci])                      # a_i . x_c
    t = np.zeros((sc.P, 3)); np.add.at(t, pi, b * s[:, None])  # sum b sigma
    z = np.einsum("pij,pj->pi", Vinv, t)
    w = np.einsum("nk,nk->n", b, z[pi])
    y = np.einsum("cij,cj->ci", U + mu * np.eye(8), x)
    np.add.at(y, ci, -a * w[:, None])
    K.schur_prepare(mu)
    yg = K.schur_matvec(torch.from_numpy(x).cuda()).cpu().numpy()
    np.testing.assert_allclose(yg, y, rtol=1e-9, atol=1e-12 * np.abs(y).max())
