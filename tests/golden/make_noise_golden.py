"""What the reference's own pose / point / intrinsics noise operations do (`Dataset.with_noise`, src/dataset/dataset.py:43-63),
measured by executing `CameraPose.apply_move` + `apply_transform_3d` (camera_pose.py:93-107), `CameraIntrinsics.apply_noise`
(camera.py:30-47) and `Dataset._random_direction*` (dataset.py:32-41) from /root/reference.  Pins
`b200ba.synth.apply_pose_noise` and records that `apply_move` is a no-op on W2C poses.
Run from the repo root:  python tests/golden/make_noise_golden.py  -> tests/golden/noise_ops.npz"""
import os
import sys

import numpy as np
from scipy.spatial.transform import Rotation

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(1, "/root/reference")
import b200ba  # noqa: E402
from src.dataset.camera import CameraIntrinsics  # noqa: E402
from src.dataset.camera_pose.camera_pose import CameraPose  # noqa: E402
from src.dataset.camera_pose.enums_and_types import CoordinateSystem, TransformationDirection  # noqa: E402
from src.dataset.dataset import Dataset  # noqa: E402

rng = np.random.default_rng(0)
poses = b200ba.synth._look_at_cameras(rng, 6)
move = rng.random((6, 3)) * 5e-2
rotvec = rng.random((6, 3)) * 5e-2
after_move, after_both = [], []
for c in range(6):
    p = CameraPose(rotation=Rotation.from_matrix(poses[c, :, :3]), translation=poses[c, :, 3].copy(),
                   coordinate_system=CoordinateSystem.COLMAP, direction=TransformationDirection.W2C)
    p.apply_move(move[c])
    after_move.append(p.rotation_translation_matrix)
    p.apply_transform_3d(Rotation.from_rotvec(rotvec[c]).as_matrix())
    after_both.append(p.rotation_translation_matrix)
K = np.array([[1200.0, 0.0, 500.0], [0.0, 1200.0, 500.0], [0.0, 0.0, 1.0]])
ci = CameraIntrinsics(K.copy(), (1200.0, 1200.0), 0.0, 500.0, 500.0)
nm = rng.random((3, 3)) * 10
ci.apply_noise(nm)
dirs3 = np.stack([Dataset._random_direction() for _ in range(200)])
dirs2 = np.stack([Dataset._random_direction_2d() for _ in range(200)])
np.savez_compressed(os.path.join(HERE, "noise_ops.npz"), poses=poses, move=move, rotvec=rotvec, after_move=np.stack(after_move),
                    after_both=np.stack(after_both), K0=K, noise_matrix=nm, K1=ci.camera_intrinsics_matrix,
                    intr1=np.array([ci.focal_x, ci.focal_y, ci.center_x, ci.center_y, ci.skew_factor]), dirs3=dirs3, dirs2=dirs2)
print("apply_move changed W2C poses by", np.abs(np.stack(after_move) - poses).max())
