"""CPU tests: the C-ABI library loads and exports every symbol include/b200ba.h declares; host logic
(packing, dense->sparse contract, point sharding incl. a world_size-2 gloo run). No compute calls."""
import os
import re
import socket

import numpy as np
import pytest
import torch

import b200ba

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "b200ba.h")).read()
    declared = set(re.findall(r"\b(b200ba_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    lib = b200ba.load()
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert set(b200ba.SYMBOLS) <= declared
    assert lib.b200ba_version() >= 100
    assert lib.b200ba_launch_count() == 0


def test_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    ba = b200ba.B200BundleAdjustment(3, 1000.0)
    sc = b200ba.synth.make_ba_scene("tiny", seed=1)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    with pytest.raises(b200ba.B200BAError):
        ba.optimize(x0, *sc.to_dense_padded()[:2], k3, sc.to_dense_padded()[2])
    with pytest.raises(b200ba.B200BAError):
        b200ba.B200PoseOptimizer(1000.0).optimize(np.zeros((1, 8)), np.zeros((1, 4, 3)), np.zeros((1, 4, 2)),
                                                   np.zeros((1, 3)), np.zeros((1, 4)))


def test_product_code_never_imports_oracle():
    pkg = os.path.join(ROOT, "jaxopt-3d-reconstruction_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "oracle" not in src.replace("oracle and the CUDA", ""), f


def test_prepare_params_layout():
    """[C*6 | C*2 | P*3] and (cx,cy,skew) as bundle_adjustment.py:93-102 / helpers.py:67-74."""
    sc = b200ba.synth.make_ba_scene("tiny", seed=2)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    assert x0.shape == (8 * sc.C + 3 * sc.P,) and k3.shape == (sc.C, 3)
    np.testing.assert_array_equal(x0[6 * sc.C:8 * sc.C].reshape(sc.C, 2), sc.intrinsics0[:, :2])
    np.testing.assert_array_equal(x0[8 * sc.C:].reshape(-1, 3), sc.points0)
    np.testing.assert_array_equal(x0[3:6], sc.poses0[0, :, 3])
    np.testing.assert_array_equal(k3, sc.intrinsics0[:, 2:])
    po = b200ba.B200PoseOptimizer(1000.0)
    q, c3 = po.prepare_params(sc.poses0, sc.intrinsics0)
    assert q.shape == (sc.C, 8) and c3.shape == (sc.C, 3)
    np.testing.assert_array_equal(q[:, 6:], sc.intrinsics0[:, :2])


def test_dense_padded_layout_matches_reference_adapter():
    """pad = the camera's first observation, mask 0 (benchmark_bundle_adjustment.py:81-84)."""
    sc = b200ba.synth.make_ba_scene("tiny", seed=3)
    p2d, idx, mask = sc.to_dense_padded()
    assert mask.sum() == sc.N
    for c in range(sc.C):
        n = int(mask[c].sum())
        assert np.all(mask[c, :n] == 1) and np.all(mask[c, n:] == 0)
        assert np.all(idx[c, n:] == idx[c, 0]) and np.all(p2d[c, n:] == p2d[c, 0])
    rows = {(c, int(i)) for c in range(sc.C) for i in idx[c, :int(mask[c].sum())]}
    assert rows == set(zip(sc.cam_idx.tolist(), sc.pt_idx.tolist()))


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_shard_points_partitions_and_balances(world):
    sc = b200ba.synth.make_ba_scene("ladybug49", seed=0)
    b = b200ba.shard_points(sc.pt_idx, sc.P, world)
    assert b[0] == 0 and b[-1] == sc.P and np.all(np.diff(b) >= 0)
    cnt = [int(((sc.pt_idx >= b[r]) & (sc.pt_idx < b[r + 1])).sum()) for r in range(world)]
    assert sum(cnt) == sc.N
    assert max(cnt) - min(cnt) <= 0.02 * sc.N / world + 64


def test_robust_loss_arguments_are_validated_on_the_host():
    """`robust_loss` belongs to the 2-vector model (the reference's own cost has no robustifier); unknown names are refused
    before anything touches a device."""
    with pytest.raises(ValueError):
        b200ba.B200BundleAdjustment(3, 1000.0, device="cpu", robust_loss="cauchy")
    with pytest.raises(ValueError):
        b200ba.B200BundleAdjustment(3, 1000.0, device="cpu", residual_model="vector2", robust_loss="huber")
    with pytest.raises(ValueError):
        b200ba.B200BundleAdjustment(3, 1000.0, device="cpu", residual_model="vector3")
    ba = b200ba.B200BundleAdjustment(3, 1000.0, device="cpu", residual_model="vector2", robust_loss="soft_l1", loss_scale=2)
    assert ba.robust_loss == "soft_l1" and ba.loss_scale == 2.0 and ba.jac_f32 is False
    assert b200ba.B200BundleAdjustment(3, 1000.0, device="cpu", residual_model="vector2", robust_loss="trivial").robust_loss is None
    hdr = open(os.path.join(ROOT, "include", "b200ba.h")).read()
    assert "B200BA_ROBUST_CAUCHY 1" in hdr and "B200BA_ROBUST_SOFT_L1 2" in hdr  # the codes api.py passes


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = b200ba.synth.make_ba_scene("small", seed=0)
    # multi-rank mode is opt-in: an initialised process group alone must NOT switch it on
    assert b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, device="cpu").world == 1
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, device="cpu", distributed=True)
    assert ba.world == world and ba.rank == rank
    b = b200ba.shard_points(sc.pt_idx, sc.P, world)
    np.testing.assert_array_equal(b, b200ba.shard_points(torch.from_numpy(sc.pt_idx), sc.P, world))  # tensor path = host path
    keep = (sc.pt_idx >= b[rank]) & (sc.pt_idx < b[rank + 1])
    ci, pi, ob, (lo, hi), b2 = b200ba.shard_rows(sc.cam_idx, sc.pt_idx, sc.obs, sc.P, world, rank)
    assert (lo, hi) == (b[rank], b[rank + 1]) and np.array_equal(b2, b)
    assert np.array_equal(ci, sc.cam_idx[keep]) and np.array_equal(pi + lo, sc.pt_idx[keep]) and np.array_equal(ob, sc.obs[keep])
    # the sum of per-rank partial camera reductions equals the global one (what NCCL sums on GPU)
    part = torch.zeros(sc.C, dtype=torch.float64)
    part.index_add_(0, torch.from_numpy(sc.cam_idx[keep].astype(np.int64)), torch.from_numpy(sc.obs[keep, 0]))
    dist.all_reduce(part)
    full = torch.zeros(sc.C, dtype=torch.float64)
    full.index_add_(0, torch.from_numpy(sc.cam_idx.astype(np.int64)), torch.from_numpy(sc.obs[:, 0]))
    # the full point array is assembled by one all-reduce of zero-filled buffers holding disjoint ranges (api._solve)
    pts = torch.zeros(sc.P, 3, dtype=torch.float64)
    pts[lo:hi] = torch.from_numpy(sc.points0[lo:hi])
    dist.all_reduce(pts)
    q.put((rank, bool(torch.allclose(part, full, rtol=1e-12)), bool(torch.equal(pts, torch.from_numpy(sc.points0)))))
    dist.destroy_process_group()


def test_point_sharding_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    [p.join(120) for p in ps]
    out = sorted(q.get(timeout=10) for _ in range(2))
    assert out == [(0, True, True), (1, True, True)]


def test_unpack_params_inverts_prepare_params_and_matches_scipy():
    """`unpack_params` = array part of `_parse_output_params_bundle` (helpers.py:59-115, Rotation.from_rotvec)."""
    import b200ba
    from scipy.spatial.transform import Rotation
    sc = b200ba.synth.make_ba_scene("small", seed=0)
    ba = b200ba.B200BundleAdjustment.__new__(b200ba.B200BundleAdjustment)
    ba.cam_num = sc.C
    x0, k3 = b200ba.B200BundleAdjustment.prepare_params(ba, sc.poses0, sc.intrinsics0, sc.points0)
    poses, intr, pts = b200ba.unpack_params(x0, k3, sc.C)
    np.testing.assert_allclose(poses, sc.poses0, atol=1e-13)
    np.testing.assert_array_equal(intr, sc.intrinsics0)
    np.testing.assert_array_equal(pts, sc.points0)
    w = np.random.default_rng(0).normal(size=(50, 3)) * np.array([1e-9, 1e-3, 1.0, 3.0, 0.0])[np.arange(50) % 5][:, None]
    np.testing.assert_allclose(b200ba.rot_vec_to_mat(w), Rotation.from_rotvec(w).as_matrix(), atol=1e-14)
    bad = x0.copy()
    bad[3] = np.nan
    with pytest.raises(ValueError):
        b200ba.unpack_params(bad, k3, sc.C)


def test_synthetic_noise_model_vs_reference_operations():
    """tests/golden/noise_ops.npz = the reference's own CameraPose / CameraIntrinsics / Dataset._random_direction operations
    (tests/golden/make_noise_golden.py) that `Dataset.with_noise` (dataset.py:43-63) is made of."""
    import b200ba
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "noise_ops.npz"))
    # the reference's apply_move is a no-op on W2C poses (camera_pose.py:101-107) ...
    np.testing.assert_allclose(g["after_move"], g["poses"], atol=1e-13)
    # ... so its pose noise is the rotation step alone = apply_pose_noise with no displacement
    np.testing.assert_allclose(b200ba.synth.apply_pose_noise(g["poses"], np.zeros((6, 3)), g["rotvec"]), g["after_both"], atol=1e-13)
    # with a displacement the synthetic scenes move the camera CENTRE by exactly `move` (the operation's stated intent)
    # (the rotation step Rn [R|t] leaves the centre -R^T t where it is)
    q = b200ba.synth.apply_pose_noise(g["poses"], g["move"], g["rotvec"])
    centre = lambda T: -(np.swapaxes(T[:, :, :3], 1, 2) @ T[:, :, 3:])[..., 0]  # noqa: E731
    np.testing.assert_allclose(centre(q) - centre(g["poses"]), g["move"], atol=1e-13)
    # intrinsics noise: fx, fy, cx, cy, skew each get their own entry of a U[0,10)^(3x3) matrix (camera.py:30-47)
    nm = g["noise_matrix"]
    np.testing.assert_allclose(g["intr1"], [1200 + nm[0, 0], 1200 + nm[1, 1], 500 + nm[0, 2], 500 + nm[1, 2], nm[0, 1]])
    # random directions are unit vectors in the POSITIVE orthant (rand, not randn: dataset.py:32-41) — like synth._pos_dir
    for d in (g["dirs3"], g["dirs2"], b200ba.synth._pos_dir(np.random.default_rng(0), 200, 3)):
        assert np.all(d >= 0) and np.allclose(np.linalg.norm(d, axis=1), 1.0)
