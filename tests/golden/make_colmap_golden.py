"""Writes the COLMAP sparse-model fixtures under tests/golden/colmap_small/ with the REFERENCE's own writer
(`src/dataset/loaders/colmap_dataset_loader/read_write_model.py`, imported from /root/reference in the build
container — it needs numpy only) plus `expected.npz`, the arrays that were fed to it.  Run from the repo root:
    python tests/golden/make_colmap_golden.py
The fixtures pin `b200ba.model_io`: its readers must recover `expected.npz` from both the .bin
and the .txt files, and its binary writer must reproduce the .bin files byte for byte.
"""
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import b200ba  # noqa: E402

REF = "/root/reference/src/dataset/loaders/colmap_dataset_loader/read_write_model.py"
spec = importlib.util.spec_from_file_location("ref_rwm", REF)
rwm = importlib.util.module_from_spec(spec)
spec.loader.exec_module(rwm)

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "colmap_small")
os.makedirs(OUT, exist_ok=True)

sc = b200ba.synth.make_ba_scene(C=5, P=40, N=130, seed=7)
rng = np.random.default_rng(11)
img_ids = np.array([3, 1, 7, 12, 5])          # unordered, non-contiguous image ids
cam_ids = np.array([2, 2, 9, 4, 4])           # shared cameras
pt_ids = np.sort(rng.choice(np.arange(1, 400), sc.P, replace=False)).astype(np.int64)[rng.permutation(sc.P)]
models = {2: "PINHOLE", 9: "SIMPLE_RADIAL", 4: "SIMPLE_PINHOLE"}

cameras, images, points = {}, {}, {}
for cid, model in models.items():
    c = int(np.flatnonzero(cam_ids == cid)[0])
    fx, fy, cx, cy, _ = sc.intrinsics0[c]
    params = {"PINHOLE": [fx, fy, cx, cy], "SIMPLE_RADIAL": [fx, cx, cy, 0.0125], "SIMPLE_PINHOLE": [fx, cx, cy]}[model]
    cameras[cid] = rwm.Camera(id=cid, model=model, width=1000 + 10 * cid, height=800, params=np.array(params))

exp = dict(img_ids=img_ids, cam_ids=cam_ids, pt_ids=pt_ids, xyz=sc.points0)
tracks = {int(p): ([], []) for p in pt_ids}
for c in range(sc.C):
    rows = np.flatnonzero(sc.cam_idx == c)
    n_extra = 2 + c                               # 2D points without a 3D point
    xys = np.concatenate([sc.obs[rows], rng.uniform(0, 1000, (n_extra, 2))])
    ids = np.concatenate([pt_ids[sc.pt_idx[rows]], np.full(n_extra, -1)])
    perm = rng.permutation(len(ids))
    xys, ids = xys[perm], ids[perm]
    R = sc.poses0[c, :, :3]
    q = b200ba.model_io.matrix_to_quat_wxyz(R)
    images[int(img_ids[c])] = rwm.Image(id=int(img_ids[c]), qvec=q, tvec=sc.poses0[c, :, 3].copy(), camera_id=int(cam_ids[c]),
                                        name=f"dir {c}/img_{c}.jpg" if c == 2 else f"img_{c}.jpg", xys=xys, point3D_ids=ids)
    for k, pid in enumerate(ids):
        if pid >= 0:
            tracks[int(pid)][0].append(int(img_ids[c]))
            tracks[int(pid)][1].append(k)
    exp[f"xys_{c}"], exp[f"p3d_{c}"], exp[f"q_{c}"], exp[f"t_{c}"] = xys, ids, q, sc.poses0[c, :, 3]
rgb = rng.integers(0, 256, (sc.P, 3)).astype(np.uint8)
err = rng.uniform(0, 2, sc.P)
for i, pid in enumerate(pt_ids):
    points[int(pid)] = rwm.Point3D(id=int(pid), xyz=sc.points0[i], rgb=rgb[i], error=float(err[i]),
                                   image_ids=np.array(tracks[int(pid)][0]), point2D_idxs=np.array(tracks[int(pid)][1]))
exp.update(rgb=rgb, err=err)

rwm.write_model(cameras, images, points, OUT, ext=".bin")
rwm.write_model(cameras, images, points, OUT, ext=".txt")
# what the reference's own readers return from those files (cross-check of the fixture itself)
c2, i2, p2 = rwm.read_model(OUT, ext=".bin")
assert set(c2) == set(cameras) and set(i2) == set(images) and set(p2) == set(points)
np.savez_compressed(os.path.join(OUT, "expected.npz"), **exp)
print("wrote", sorted(os.listdir(OUT)), {f: os.path.getsize(os.path.join(OUT, f)) for f in os.listdir(OUT)})
