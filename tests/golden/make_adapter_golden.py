"""Golden arrays from the reference's OWN dataset adapter, executed in the build container: a reference `Dataset` is
built from tests/golden/colmap_small/ (through the reference's readers `read_cameras_bin` / `read_images_bin` /
`read_points3d_bin` and its `_parse_points` / `get_intrinsics`; image files are not needed because widths come from
cameras.bin), then `JaxoptBundleAdjustmentBenchmark._prepare_dataset` (benchmark_bundle_adjustment.py:45-106, which
calls `Dataset.make_reduced_dataset`, dataset.py:172-201) runs for a full and a reduced (camera_limit, points_limit)
configuration; and `JaxoptSinglePoseBenchmarkBatched.setup/_prepare_points/_get_mask`
(benchmark_pose_optimization.py:32-130) for the single-pose batch.  jax is replaced by tests/golden/jax_numpy_shim.py (only imports are touched on this code path).
Run from the repo root:  python tests/golden/make_adapter_golden.py  -> tests/golden/adapter_colmap_small.npz"""
import os
import sys
from pathlib import Path

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import jax_numpy_shim  # noqa: E402

jax_numpy_shim.install()
import types  # noqa: E402
# imported at module level by helpers.py / benchmark.py (plots, Se3 output parsing); absent here, unused on this path
for _m, _names in {"matplotlib": ("pyplot",), "matplotlib.pyplot": (), "triangulation_relaxations": (),
                   "triangulation_relaxations.se3": ("Se3",), "triangulation_relaxations.so3": ("rotvec_to_r",)}.items():
    _mod = sys.modules.setdefault(_m, types.ModuleType(_m))
    for _n in _names:
        setattr(_mod, _n, None)
sys.path.insert(1, "/root/reference")
from src.benchmark.jaxopt_benchmark.benchmark_bundle_adjustment import JaxoptBundleAdjustmentBenchmark  # noqa: E402
from src.benchmark.jaxopt_benchmark.benchmark_pose_optimization import JaxoptSinglePoseBenchmarkBatched  # noqa: E402
from src.dataset.camera import Camera  # noqa: E402
from src.dataset.camera_pose.camera_pose import CameraPose  # noqa: E402
from src.dataset.camera_pose.enums_and_types import CoordinateSystem, TransformationDirection  # noqa: E402
from src.dataset.dataset import Dataset  # noqa: E402
from src.dataset.datasetEntry import DatasetEntry  # noqa: E402
from src.dataset.imageMetadata import ImageMetadata  # noqa: E402
from src.dataset.loaders.colmap_dataset_loader.cameras import read_cameras_bin  # noqa: E402
from src.dataset.loaders.colmap_dataset_loader.images import read_images_bin  # noqa: E402
from src.dataset.loaders.colmap_dataset_loader.loader import _parse_points, get_intrinsics, params_to_intrinsics  # noqa: E402
from src.dataset.loaders.colmap_dataset_loader.points import read_points3d_bin  # noqa: E402
from src.dataset.point import Point2D  # noqa: E402

D = os.path.join(HERE, "colmap_small")
cameras, images, points = read_cameras_bin(f"{D}/cameras.bin"), read_images_bin(f"{D}/images.bin"), read_points3d_bin(f"{D}/points3D.bin")
entries = []
for im in images.values():  # loader.py:60-93 without the PIL size lookup
    info = im.image_information
    cam = cameras.get(info.camera_id)
    pose = CameraPose.from_string_wxyz_quaternion_translation(
        f"{info.qw!r} {info.qx!r} {info.qy!r} {info.qz!r} {info.tx!r} {info.ty!r} {info.tz!r}", identifier=Path(info.name).name,
        coordinate_system=CoordinateSystem.COLMAP, direction=TransformationDirection.W2C)
    # the reference's get_intrinsics (loader.py:33-43) knows SIMPLE_RADIAL and 4/5-parameter (fx, fy, cx, cy[, s]) models;
    # SIMPLE_PINHOLE (f, cx, cy) makes it raise, so that one camera is expanded to (f, f, cx, cy) here
    intr = (params_to_intrinsics(cam.params[0], cam.params[0], cam.params[1], cam.params[2])
            if cam.camera_model_type.name == "SIMPLE_PINHOLE" else get_intrinsics(cam))
    entries.append(DatasetEntry(ImageMetadata(info.name, None, cam.width, cam.height),
                                [Point2D(p.id, p.x, p.y, p.point3d_id, {}) for p in im.point2d_entries],
                                Camera(camera_pose=pose, camera_intrinsics=intr, width=cam.width, height=cam.height)))
ds = Dataset(_parse_points(points), entries, name="colmap_small")

out = {}
for tag, (cl, pl) in {"full": (None, None), "reduced": (4, 12)}.items():
    b = JaxoptBundleAdjustmentBenchmark(ds)
    b.camera_limit, b.points_limit = cl, pl
    p2d, p3d, idx, mask, poses, K, idx2id, w = b._prepare_dataset()
    out.update({f"{tag}_p2d": p2d, f"{tag}_p3d": p3d, f"{tag}_idx": idx, f"{tag}_mask": mask, f"{tag}_poses": poses, f"{tag}_K": K,
                f"{tag}_point_ids": np.array([idx2id[i] for i in range(len(p3d))]), f"{tag}_avg_cam_width": w})
    print(tag, p2d.shape, p3d.shape, int(mask.sum()), w)
# ---- single-pose adapter (benchmark_pose_optimization.py:32-130): per-image correspondences, zero-padded to the longest
pb = JaxoptSinglePoseBenchmarkBatched(ds)
pb.setup()
B = len(pb)
pts, obs = pb._prepare_points(0, batch_size=B)
out.update(pose_poses=pb.cam_poses, pose_K=pb.intrinsics, pose_points=np.asarray(pts), pose_obs=np.asarray(obs),
           pose_mask=np.asarray(pb._get_mask(0, batch_size=B)), pose_avg_cam_width=pb.avg_cam_width)
print("pose", np.asarray(pts).shape, pb.avg_cam_width)
np.savez_compressed(os.path.join(HERE, "adapter_colmap_small.npz"), **out)
