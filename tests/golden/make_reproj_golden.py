"""Golden vectors for the reprojection-error evaluator computed by the REFERENCE ITSELF:
`Camera.compute_projection_errors_alt` (src/dataset/camera.py:78-85) imported from /root/reference in the build
container (numpy + scipy only), fed the `small` synthetic scene at its noisy start.  Run from the repo root:
    python tests/golden/make_reproj_golden.py
Writes tests/golden/reproj_small_seed0.npz: per-row errors (caller's row order) for both loss functions."""
import os
import sys

import numpy as np
from scipy.spatial.transform import Rotation

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(1, "/root/reference")
import b200ba  # noqa: E402
from src.dataset.camera import Camera, CameraIntrinsics  # noqa: E402  (the reference)
from src.dataset.camera_pose.camera_pose import CameraPose  # noqa: E402
from src.dataset.camera_pose.enums_and_types import CoordinateSystem, TransformationDirection  # noqa: E402
from src.dataset.loss_functions import LossFunction  # noqa: E402

sc = b200ba.synth.make_ba_scene("small", seed=0)
out = {}
for name, lf in (("trivial", LossFunction.TRIVIAL_LOSS), ("cauchy", LossFunction.CAUCHY_LOSS)):
    err = np.full(sc.N, np.nan)
    for c in range(sc.C):
        fx, fy, cx, cy, sk = sc.intrinsics0[c]
        K = np.array([[fx, sk, cx], [0, fy, cy], [0, 0, 1.0]])
        pose = CameraPose(rotation=Rotation.from_matrix(sc.poses0[c, :, :3]), translation=sc.poses0[c, :, 3].copy(),
                          coordinate_system=CoordinateSystem.COLMAP, direction=TransformationDirection.W2C)
        cam = Camera(camera_pose=pose, camera_intrinsics=CameraIntrinsics(K, (fx, fy), sk, cx, cy), width=1000, height=1000)
        rows = np.flatnonzero(sc.cam_idx == c)
        e = cam.compute_projection_errors_alt(p2d=list(sc.obs[rows]), p3d=list(sc.points0[sc.pt_idx[rows]]), loss_function=lf)
        err[rows] = e
    assert np.isfinite(err).all()
    out[name] = err
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "reproj_small_seed0.npz"), **out)
print({k: (v.sum(), v.max()) for k, v in out.items()})
