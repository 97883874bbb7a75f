"""f3 — device-side `Dataset.with_noise` (noise_kernels.cuh, `b200ba.with_noise`) against the reference's OWN operations
(tests/golden/noise_ops.npz, made by tests/golden/make_noise_golden.py from CameraPose / CameraIntrinsics / Dataset)."""
import os

import numpy as np
import pytest

import b200ba

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "noise_ops.npz"))


def test_pose_and_intrinsics_operations_equal_the_reference():
    poses, move, rotvec = G["poses"], G["move"], G["rotvec"]
    C = len(poses)
    intr = np.tile(np.array([1200.0, 1200.0, 500.0, 500.0, 0.0]), (C, 1))
    nm = G["noise_matrix"]
    add = np.tile(np.array([nm[0, 0], nm[1, 1], nm[0, 2], nm[1, 2], nm[0, 1]]), (C, 1))
    ex = {"move_vec": move, "rot_vec": rotvec, "intr_add": add, "pts_vec": np.zeros((1, 3)), "obs_vec": np.zeros((1, 2))}
    # the reference as it behaves: apply_move is a no-op on W2C poses, then Rn [R|t]
    p, k, _, _ = b200ba.with_noise(poses, intr, np.zeros((1, 3)), np.zeros((1, 2)), explicit=ex, move_cameras=False)
    np.testing.assert_allclose(p, G["after_both"], atol=1e-13)
    np.testing.assert_allclose(k[0], G["intr1"], atol=1e-12)
    # the intended displacement of the camera centre = the host generator of the synthetic workloads
    p2, _, _, _ = b200ba.with_noise(poses, intr, np.zeros((1, 3)), np.zeros((1, 2)), explicit=ex, move_cameras=True)
    np.testing.assert_allclose(p2, b200ba.synth.apply_pose_noise(poses, move, rotvec), atol=1e-13)


def test_generated_perturbations_follow_the_reference_distributions():
    rng = np.random.default_rng(0)
    C, P, N = 300, 5000, 20000
    poses = b200ba.synth._look_at_cameras(rng, C)
    intr = np.tile(np.array([1200.0, 1200.0, 500.0, 500.0, 0.0]), (C, 1))
    pts, obs = rng.uniform(-1, 1, (P, 3)), rng.uniform(0, 1000, (N, 2))
    a = b200ba.with_noise(poses, intr, pts, obs, seed=7)
    b = b200ba.with_noise(poses, intr, pts, obs, seed=7)
    c = b200ba.with_noise(poses, intr, pts, obs, seed=8)
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)          # a function of the seed only
    assert not np.array_equal(a[2], c[2])
    dp, do, dk = a[2] - pts, a[3] - obs, a[1] - intr
    # magnitude x unit direction in the POSITIVE orthant (dataset.py:32-41 uses rand, not randn)
    np.testing.assert_allclose(np.linalg.norm(dp, axis=1), 3e-2, rtol=1e-9)
    np.testing.assert_allclose(np.linalg.norm(do, axis=1), 1.0, rtol=1e-9)
    assert (dp >= 0).all() and (do >= 0).all()
    # directions = normalised U[0,1)^d like np.random.rand: compare the component means with the NumPy model
    ref3 = b200ba.synth._pos_dir(np.random.default_rng(1), 200000, 3).mean(axis=0)
    np.testing.assert_allclose(dp.mean(axis=0) / 3e-2, ref3, atol=0.02)
    # intrinsics: += U[0, 10) per entry (camera.py:30-47)
    assert (dk >= 0).all() and (dk < 10).all() and abs(dk.mean() - 5.0) < 0.3
    # rotation noise: R' = Rn R with |rotvec| = 5e-2; the camera centre does not move (the reference's behaviour)
    Rn = a[0][:, :, :3] @ np.swapaxes(poses[:, :, :3], 1, 2)
    ang = np.arccos(np.clip((np.trace(Rn, axis1=1, axis2=2) - 1) / 2, -1, 1))
    np.testing.assert_allclose(ang, 5e-2, rtol=1e-7)
    centre = lambda T: -(np.swapaxes(T[:, :, :3], 1, 2) @ T[:, :, 3:])[..., 0]  # noqa: E731
    np.testing.assert_allclose(centre(a[0]), centre(poses), atol=1e-12)
    m = b200ba.with_noise(poses, intr, pts, obs, seed=7, move_cameras=True)
    np.testing.assert_allclose(np.linalg.norm(centre(m[0]) - centre(poses), axis=1), 5e-2, rtol=1e-9)
