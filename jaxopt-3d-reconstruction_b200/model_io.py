"""Real-data ingest / egress for the solver: COLMAP sparse models (binary and
text) and BAL problem files, straight into the sparse row arrays `B200BundleAdjustment.build_problem_sparse`
takes — no `C x Nmax` dense-padded intermediate (row f1).

Reference interfaces this mirrors (formats only; host code, nothing here touches the GPU):
  * `read_cameras_bin`   src/dataset/loaders/colmap_dataset_loader/cameras.py:49-59
  * `read_images_bin`    src/dataset/loaders/colmap_dataset_loader/images.py:38-55
  * `read_points3d_bin`  src/dataset/loaders/colmap_dataset_loader/points.py:35-49
  * `load_colmap_dataset` / `export_in_colmap_format`  .../loader.py:137-212
  * the array contract of `_prepare_dataset`  src/benchmark/jaxopt_benchmark/benchmark_bundle_adjustment.py:45-106
    (cameras in image order, points = the observed ones sorted by identifier, W2C `[R|t]`,
    `avg_cam_width` = mean image width) — `SparseModel.to_ba_arrays`.

Records are parsed with numpy views of the file buffer (one `frombuffer` per image / per file section), not one
`struct.unpack` per 2D point: a Final-13682-sized model (29 M observations) loads in seconds.
Tests: tests/test_model_io_cpu.py — fixtures under tests/golden/colmap_small/ were written by the reference's own
`read_write_model.py`; binary output here is byte-identical to it.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np

# model id -> (name, number of parameters); cameras.py:21-46
CAMERA_MODELS = {
    0: ("SIMPLE_PINHOLE", 3), 1: ("PINHOLE", 4), 2: ("SIMPLE_RADIAL", 4), 3: ("RADIAL", 5), 4: ("OPENCV", 8),
    5: ("OPENCV_FISHEYE", 8), 6: ("FULL_OPENCV", 12), 7: ("FOV", 5), 8: ("SIMPLE_RADIAL_FISHEYE", 4),
    9: ("RADIAL_FISHEYE", 5), 10: ("THIN_PRISM_FISHEYE", 12),
}
_MODEL_ID = {name: (mid, n) for mid, (name, n) in CAMERA_MODELS.items()}


class ModelFormatError(ValueError):
    pass


@dataclass
class CameraRec:
    camera_id: int
    model: str
    width: int
    height: int
    params: np.ndarray  # float64

    def fx_fy_cx_cy_skew(self) -> np.ndarray:
        """[fx, fy, cx, cy, skew] as `get_intrinsics` builds K (loader.py:33-43): SIMPLE_* models share f, every
        other model is read as (fx, fy, cx, cy[, skew]) and distortion terms are ignored by the BA path."""
        p = self.params
        # (the reference itself only handles SIMPLE_RADIAL and 4/5-parameter models: SIMPLE_PINHOLE makes its
        # get_intrinsics raise and RADIAL's (f, cx, cy, k1, k2) would be read as (fx, fy, cx, cy, s); here every
        # single-focal model is expanded to (f, f, cx, cy))
        if self.model in ("SIMPLE_PINHOLE", "SIMPLE_RADIAL", "SIMPLE_RADIAL_FISHEYE", "RADIAL", "RADIAL_FISHEYE"):
            return np.array([p[0], p[0], p[1], p[2], 0.0])
        return np.array([p[0], p[1], p[2], p[3], 0.0])


@dataclass
class ImageRec:
    image_id: int
    qvec: np.ndarray  # (4,) w, x, y, z  world-to-camera
    tvec: np.ndarray  # (3,)
    camera_id: int
    name: str
    xys: np.ndarray  # (n, 2) float64
    point3D_ids: np.ndarray  # (n,) int64, -1 = no 3D point


@dataclass
class SparseModel:
    """A COLMAP sparse model in arrays.  Points are struct-of-arrays; tracks are CSR."""
    cameras: Dict[int, CameraRec]
    images: Dict[int, ImageRec]
    point_ids: np.ndarray  # (P,) int64 (file order)
    xyz: np.ndarray  # (P, 3) float64
    rgb: np.ndarray  # (P, 3) uint8
    error: np.ndarray  # (P,) float64
    track_ptr: np.ndarray = field(default_factory=lambda: np.zeros(1, np.int64))  # (P+1,)
    track_image_ids: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))
    track_point2D_idxs: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))

    # ------------------------------------------------------------------ solver arrays
    def to_ba_arrays(self) -> dict:
        """The sparse row arrays of the BA problem, with the semantics of `_prepare_dataset`
        (benchmark_bundle_adjustment.py:45-106): camera c = c-th image in file order; a 2D point is an
        observation iff its 3D identifier is truthy (`points_with_3d`, datasetEntry.py:19-20 — identifier 0
        is dropped like None) and names a point of the model; points = observed identifiers, ascending.
        Returns poses0 (C,3,4), intrinsics0 (C,5), points0 (P,3), cam_idx/pt_idx (N,) int32, obs (N,2),
        point2D_idx (N,), point_ids (P,) [index -> identifier], image_ids (C,), avg_cam_width."""
        imgs = list(self.images.values())
        C = len(imgs)
        if C == 0:
            raise ModelFormatError("model has no images")
        poses = np.empty((C, 3, 4))
        intr = np.empty((C, 5))
        widths = np.empty(C)
        cam_l, id_l, xy_l, p2_l = [], [], [], []
        known = np.sort(self.point_ids)
        for c, im in enumerate(imgs):
            cam = self.cameras.get(im.camera_id)
            if cam is None:
                raise ModelFormatError(f"image {im.image_id} names camera {im.camera_id}, which is not in the model")
            poses[c, :, :3] = quat_wxyz_to_matrix(im.qvec)
            poses[c, :, 3] = im.tvec
            intr[c] = cam.fx_fy_cx_cy_skew()
            widths[c] = cam.width
            ids = im.point3D_ids
            keep = ids > 0
            if keep.any() and len(known):
                pos = np.searchsorted(known, ids[keep])
                pos[pos == len(known)] = 0
                ok = known[pos] == ids[keep]
                sel = np.flatnonzero(keep)[ok]
            else:
                sel = np.zeros(0, np.int64)
            cam_l.append(np.full(len(sel), c, np.int32))
            id_l.append(ids[sel])
            xy_l.append(im.xys[sel])
            p2_l.append(sel.astype(np.int32))
        ids = np.concatenate(id_l) if id_l else np.zeros(0, np.int64)
        point_ids, pt_idx = np.unique(ids, return_inverse=True)
        lut = np.argsort(self.point_ids, kind="stable")
        src = lut[np.searchsorted(self.point_ids[lut], point_ids)]
        return dict(poses0=poses, intrinsics0=intr, points0=self.xyz[src].astype(np.float64),
                    cam_idx=np.concatenate(cam_l), pt_idx=pt_idx.astype(np.int32),
                    obs=np.concatenate(xy_l).reshape(-1, 2).astype(np.float64), point2D_idx=np.concatenate(p2_l),
                    point_ids=point_ids, image_ids=np.array([im.image_id for im in imgs], np.int64),
                    avg_cam_width=float(widths.mean()))

    def to_pose_batch(self) -> dict:
        """The batched single-pose inputs as `JaxoptSinglePoseBenchmarkBatched` builds them
        (benchmark_pose_optimization.py:55-130): one lane per image with its 2D-3D correspondences in 2D-point order,
        zero-padded to the longest lane, mask 1 on real rows.  Returns poses0 (B,3,4), intrinsics0 (B,5),
        points (B,Nmax,3), observations (B,Nmax,2), mask (B,Nmax), avg_cam_width."""
        a = self.to_ba_arrays()
        B = len(a["poses0"])
        counts = np.bincount(a["cam_idx"], minlength=B)
        nmax = int(counts.max()) if B else 0
        slot = np.arange(len(a["cam_idx"])) - np.repeat(np.cumsum(counts) - counts, counts)  # rows are camera-major
        pts, obs, mask = np.zeros((B, nmax, 3)), np.zeros((B, nmax, 2)), np.zeros((B, nmax))
        pts[a["cam_idx"], slot] = a["points0"][a["pt_idx"]]
        obs[a["cam_idx"], slot] = a["obs"]
        mask[a["cam_idx"], slot] = 1.0
        return dict(poses0=a["poses0"], intrinsics0=a["intrinsics0"], points=pts, observations=obs, mask=mask,
                    avg_cam_width=a["avg_cam_width"])

    def reduced(self, camera_limit: Optional[int], points_limit: Optional[int]) -> "SparseModel":
        """`Dataset.make_reduced_dataset(camera_limit, points_limit)` (src/dataset/dataset.py:172-201), the step
        `_prepare_dataset` starts with: the first `camera_limit` images; per image only its first `points_limit`
        2D points that have a 3D point count; 3D points seen fewer than twice among those are dropped; every other
        2D point loses its 3D link.  None = no limit (points seen once are still dropped, as in the reference)."""
        imgs = list(self.images.values())[:camera_limit]
        firsts = []
        for im in imgs:
            firsts.append(np.flatnonzero(im.point3D_ids > 0)[:points_limit])
        seen = np.concatenate([im.point3D_ids[f] for im, f in zip(imgs, firsts)]) if imgs else np.zeros(0, np.int64)
        ids, cnt = np.unique(seen, return_counts=True)
        ids = ids[cnt >= 2]
        ids = ids[np.isin(ids, self.point_ids)]
        new_imgs = {}
        for im, f in zip(imgs, firsts):
            keep = f[np.isin(im.point3D_ids[f], ids)]
            p3 = np.full(len(im.point3D_ids), -1, np.int64)
            p3[keep] = im.point3D_ids[keep]
            new_imgs[im.image_id] = ImageRec(im.image_id, im.qvec, im.tvec, im.camera_id, im.name, im.xys, p3)
        sel = np.flatnonzero(np.isin(self.point_ids, ids))
        cams = {im.camera_id: self.cameras[im.camera_id] for im in imgs if im.camera_id in self.cameras}
        m = SparseModel(cams, new_imgs, self.point_ids[sel], self.xyz[sel], self.rgb[sel], self.error[sel])
        m.rebuild_tracks()
        return m

    def with_solution(self, poses, intrinsics, points, point_ids) -> "SparseModel":
        """A copy carrying an optimised solution (the egress half of `_parse_output_params_bundle`,
        helpers.py:67-115 + `export_in_colmap_format`, loader.py:153-212): poses (C,3,4) W2C in image order,
        intrinsics (C,5), points (P',3) for identifiers `point_ids`.  Cameras are written as PINHOLE, one per
        image, ids 1..C in image order — what the reference's exporter does (loader.py:163-170)."""
        imgs = list(self.images.values())
        if len(poses) != len(imgs):
            raise ModelFormatError("one pose per image expected")
        cams, new_imgs = {}, {}
        for k, (im, T, ki) in enumerate(zip(imgs, np.asarray(poses), np.asarray(intrinsics)), start=1):
            old = self.cameras[im.camera_id]
            cams[k] = CameraRec(k, "PINHOLE", old.width, old.height, np.array(ki[:4], np.float64))
            new_imgs[im.image_id] = ImageRec(im.image_id, matrix_to_quat_wxyz(T[:, :3]), np.array(T[:, 3], np.float64),
                                             k, im.name, im.xys, im.point3D_ids)
        xyz = self.xyz.copy()
        lut = np.argsort(self.point_ids, kind="stable")
        pos = lut[np.searchsorted(self.point_ids[lut], np.asarray(point_ids))]
        xyz[pos] = np.asarray(points)
        return SparseModel(cams, new_imgs, self.point_ids, xyz, self.rgb, self.error, self.track_ptr,
                           self.track_image_ids, self.track_point2D_idxs)

    def rebuild_tracks(self) -> None:
        """Recompute the point tracks from the images' 2D->3D links (loader.py:187-192 does the same before export)."""
        img_l, p2_l, id_l = [], [], []
        for im in self.images.values():
            sel = np.flatnonzero(im.point3D_ids >= 0)
            img_l.append(np.full(len(sel), im.image_id, np.int32))
            p2_l.append(sel.astype(np.int32))
            id_l.append(im.point3D_ids[sel])
        ids = np.concatenate(id_l) if id_l else np.zeros(0, np.int64)
        lut = np.argsort(self.point_ids, kind="stable")
        pos = np.searchsorted(self.point_ids[lut], ids)
        pos[pos == len(lut)] = 0
        ok = self.point_ids[lut][pos] == ids if len(lut) else np.zeros(len(ids), bool)
        row = lut[pos[ok]] if len(lut) else np.zeros(0, np.int64)
        order = np.argsort(row, kind="stable")
        self.track_ptr = np.concatenate([[0], np.cumsum(np.bincount(row, minlength=len(self.point_ids)))]).astype(np.int64)
        self.track_image_ids = np.concatenate(img_l)[ok][order] if len(ids) else np.zeros(0, np.int32)
        self.track_point2D_idxs = np.concatenate(p2_l)[ok][order] if len(ids) else np.zeros(0, np.int32)


def to_dense_padded(cam_idx, pt_idx, obs, C: int):
    """Sparse rows -> the reference's dense contract (benchmark_bundle_adjustment.py:72-87): `(C, Nmax, 2)` 2D points,
    `(C, Nmax)` point indices, `(C, Nmax)` mask; rows keep their order inside a camera, the padding repeats the
    camera's FIRST observation with mask 0.  For comparisons with the reference only — the solver takes the rows."""
    cam_idx, pt_idx, obs = np.asarray(cam_idx), np.asarray(pt_idx), np.asarray(obs, np.float64).reshape(-1, 2)
    order = np.argsort(cam_idx, kind="stable")
    cam_s = cam_idx[order]
    counts = np.bincount(cam_s, minlength=C)
    if counts.min() == 0:
        raise ModelFormatError("a camera without observations has no first observation to pad with")
    nmax = int(counts.max())
    start = np.concatenate([[0], np.cumsum(counts)[:-1]])
    slot = np.arange(len(order)) - start[cam_s]
    first = order[start]
    p2d = np.repeat(obs[first][:, None, :], nmax, axis=1)
    idx = np.repeat(pt_idx[first].astype(np.int64)[:, None], nmax, axis=1)
    mask = np.zeros((C, nmax))
    p2d[cam_s, slot], idx[cam_s, slot], mask[cam_s, slot] = obs[order], pt_idx[order], 1.0
    return p2d, idx, mask


# ---------------------------------------------------------------------- rotations
def quat_wxyz_to_matrix(q) -> np.ndarray:
    """Unit-normalised (w, x, y, z) -> R, the convention of `CameraPose.from_string_wxyz_quaternion_translation`
    (camera_pose.py:203-214, scipy `Rotation.from_quat` on the re-ordered xyzw)."""
    w, x, y, z = np.asarray(q, np.float64) / np.linalg.norm(q)
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def matrix_to_quat_wxyz(R) -> np.ndarray:
    """R -> (w, x, y, z), w >= 0 (largest-component branch for numerical safety)."""
    R = np.asarray(R, np.float64)
    t = np.trace(R)
    if t > 0:
        s = np.sqrt(t + 1.0) * 2
        q = np.array([0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s])
    else:
        i = int(np.argmax(np.diag(R)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = np.sqrt(1.0 + R[i, i] - R[j, j] - R[k, k]) * 2
        q = np.empty(4)
        q[0] = (R[k, j] - R[j, k]) / s
        q[1 + i] = 0.25 * s
        q[1 + j] = (R[j, i] + R[i, j]) / s
        q[1 + k] = (R[k, i] + R[i, k]) / s
    return q if q[0] >= 0 else -q


# ---------------------------------------------------------------------- COLMAP binary
_CAM_HEAD = np.dtype([("id", "<u4"), ("model", "<i4"), ("w", "<u8"), ("h", "<u8")])  # "<IiQQ", cameras.py:54
_IMG_HEAD = np.dtype([("id", "<u4"), ("q", "<f8", 4), ("t", "<f8", 3), ("cam", "<u4")])  # "<I7dI", images.py:43
_P2D = np.dtype([("xy", "<f8", 2), ("p3d", "<u8")])  # "<2dQ", images.py:48
_P3D_HEAD = np.dtype([("id", "<u8"), ("xyz", "<f8", 3), ("rgb", "u1", 3), ("err", "<f8")])  # "<Q3d3Bd", points.py:40
_TRACK = np.dtype([("img", "<u4"), ("p2d", "<u4")])  # "<2I", points.py:45
assert _CAM_HEAD.itemsize == 24 and _IMG_HEAD.itemsize == 64 and _P2D.itemsize == 24
assert _P3D_HEAD.itemsize == 43 and _TRACK.itemsize == 8


def _need(buf, off, n, what):
    if off + n > len(buf):
        raise ModelFormatError(f"{what}: file truncated at byte {off} (need {n}, have {len(buf) - off})")


def read_cameras_bin(path) -> Dict[int, CameraRec]:
    buf = open(path, "rb").read()
    _need(buf, 0, 8, "cameras.bin")
    n = int(np.frombuffer(buf, "<u8", 1, 0)[0])
    off, out = 8, {}
    for _ in range(n):
        _need(buf, off, 24, "cameras.bin")
        h = np.frombuffer(buf, _CAM_HEAD, 1, off)[0]
        off += 24
        if int(h["model"]) not in CAMERA_MODELS:
            raise ModelFormatError(f"cameras.bin: unknown camera model id {int(h['model'])}")
        name, npar = CAMERA_MODELS[int(h["model"])]
        _need(buf, off, 8 * npar, "cameras.bin")
        out[int(h["id"])] = CameraRec(int(h["id"]), name, int(h["w"]), int(h["h"]),
                                      np.frombuffer(buf, "<f8", npar, off).copy())
        off += 8 * npar
    return out


def read_images_bin(path) -> Dict[int, ImageRec]:
    buf = open(path, "rb").read()
    _need(buf, 0, 8, "images.bin")
    n = int(np.frombuffer(buf, "<u8", 1, 0)[0])
    off, out = 8, {}
    for _ in range(n):
        _need(buf, off, 64, "images.bin")
        h = np.frombuffer(buf, _IMG_HEAD, 1, off)[0]
        off += 64
        end = buf.find(b"\x00", off)
        if end < 0:
            raise ModelFormatError("images.bin: unterminated image name")
        name = buf[off:end].decode("utf-8")
        off = end + 1
        _need(buf, off, 8, "images.bin")
        m = int(np.frombuffer(buf, "<u8", 1, off)[0])
        off += 8
        _need(buf, off, 24 * m, "images.bin")
        rec = np.frombuffer(buf, _P2D, m, off)
        off += 24 * m
        ids = rec["p3d"].astype(np.int64)  # 2^64-1 wraps to -1 == "no 3D point"
        out[int(h["id"])] = ImageRec(int(h["id"]), h["q"].copy(), h["t"].copy(), int(h["cam"]), name,
                                     rec["xy"].copy().reshape(m, 2), ids)
    return out


def read_points3d_bin(path):
    """-> (point_ids, xyz, rgb, error, track_ptr, track_image_ids, track_point2D_idxs)."""
    buf = open(path, "rb").read()
    _need(buf, 0, 8, "points3D.bin")
    n = int(np.frombuffer(buf, "<u8", 1, 0)[0])
    ids, xyz, rgb, err = np.empty(n, np.int64), np.empty((n, 3)), np.empty((n, 3), np.uint8), np.empty(n)
    ptr = np.zeros(n + 1, np.int64)
    starts = np.empty(n, np.int64)
    off = 8
    u8 = np.frombuffer(buf, np.uint8)
    for i in range(n):  # record length depends on the track length: offsets are inherently sequential
        _need(buf, off, 51, "points3D.bin")
        tl = int.from_bytes(buf[off + 43:off + 51], "little")
        starts[i] = off
        ptr[i + 1] = ptr[i] + tl
        off += 51 + 8 * tl
    _need(buf, off, 0, "points3D.bin")
    if n:
        head = u8[(starts[:, None] + np.arange(43)[None, :])].reshape(-1).view(_P3D_HEAD)
        ids[:], xyz[:], rgb[:], err[:] = head["id"].astype(np.int64), head["xyz"], head["rgb"], head["err"]
    T = int(ptr[-1])
    t_img, t_p2 = np.empty(T, np.int32), np.empty(T, np.int32)
    if T:
        # byte index of every track entry: record start + 51 + 8*(k - ptr[i])
        rec = np.repeat(np.arange(n), np.diff(ptr))
        base = starts[rec] + 51 + 8 * (np.arange(T) - ptr[rec])
        tr = u8[(base[:, None] + np.arange(8)[None, :])].reshape(-1).view(_TRACK)
        t_img[:], t_p2[:] = tr["img"].astype(np.int32), tr["p2d"].astype(np.int32)
    return ids, xyz, rgb, err, ptr, t_img, t_p2


def write_cameras_bin(cameras: Dict[int, CameraRec], path) -> None:
    with open(path, "wb") as f:
        f.write(np.uint64(len(cameras)).tobytes())
        for c in cameras.values():
            mid, npar = _MODEL_ID[c.model]
            if len(c.params) != npar:
                raise ModelFormatError(f"camera {c.camera_id}: model {c.model} takes {npar} parameters")
            h = np.zeros(1, _CAM_HEAD)
            h["id"], h["model"], h["w"], h["h"] = c.camera_id, mid, c.width, c.height
            f.write(h.tobytes())
            f.write(np.asarray(c.params, "<f8").tobytes())


def write_images_bin(images: Dict[int, ImageRec], path) -> None:
    with open(path, "wb") as f:
        f.write(np.uint64(len(images)).tobytes())
        for im in images.values():
            h = np.zeros(1, _IMG_HEAD)
            h["id"], h["q"], h["t"], h["cam"] = im.image_id, im.qvec, im.tvec, im.camera_id
            f.write(h.tobytes())
            f.write(im.name.encode("utf-8") + b"\x00")
            m = len(im.point3D_ids)
            f.write(np.uint64(m).tobytes())
            rec = np.zeros(m, _P2D)
            rec["xy"] = np.asarray(im.xys, np.float64).reshape(m, 2)
            rec["p3d"] = np.asarray(im.point3D_ids, np.int64).view(np.uint64)  # -1 -> 2^64-1
            f.write(rec.tobytes())


def write_points3d_bin(model: SparseModel, path) -> None:
    n = len(model.point_ids)
    ptr = model.track_ptr
    if len(ptr) != n + 1:
        raise ModelFormatError("track_ptr must have P+1 entries (call rebuild_tracks())")
    tl = np.diff(ptr)
    starts = 8 + 51 * np.arange(n, dtype=np.int64) + 8 * ptr[:-1]
    out = np.zeros(8 + 51 * n + 8 * int(ptr[-1]), np.uint8)
    out[:8] = np.frombuffer(np.uint64(n).tobytes(), np.uint8)
    if n:
        head = np.zeros(n, _P3D_HEAD)
        head["id"], head["xyz"], head["rgb"], head["err"] = model.point_ids, model.xyz, model.rgb, model.error
        out[starts[:, None] + np.arange(43)[None, :]] = head.view(np.uint8).reshape(n, 43)
        out[(starts + 43)[:, None] + np.arange(8)[None, :]] = tl.astype("<u8").view(np.uint8).reshape(n, 8)
    T = int(ptr[-1])
    if T:
        tr = np.zeros(T, _TRACK)
        tr["img"], tr["p2d"] = model.track_image_ids, model.track_point2D_idxs
        rec = np.repeat(np.arange(n), tl)
        base = starts[rec] + 51 + 8 * (np.arange(T) - ptr[rec])
        out[base[:, None] + np.arange(8)[None, :]] = tr.view(np.uint8).reshape(T, 8)
    with open(path, "wb") as f:
        f.write(out.tobytes())


# ---------------------------------------------------------------------- COLMAP text
def _data_lines(path, keep_empty=False):
    with open(path, "r") as f:
        for line in f:
            s = line.strip()
            if s.startswith("#"):
                continue
            if s or keep_empty:
                yield s


def read_cameras_txt(path) -> Dict[int, CameraRec]:
    out = {}
    for s in _data_lines(path):
        t = s.split()
        if len(t) < 4 or t[1] not in _MODEL_ID:
            raise ModelFormatError(f"cameras.txt: bad line {s[:60]!r}")
        if len(t) - 4 != _MODEL_ID[t[1]][1]:
            raise ModelFormatError(f"cameras.txt: model {t[1]} takes {_MODEL_ID[t[1]][1]} parameters")
        out[int(t[0])] = CameraRec(int(t[0]), t[1], int(t[2]), int(t[3]), np.array(t[4:], np.float64))
    return out


def read_images_txt(path) -> Dict[int, ImageRec]:
    out = {}
    with open(path, "r") as f:
        lines = [ln.rstrip("\n") for ln in f]
    i = 0
    while i < len(lines):
        s = lines[i].strip()
        i += 1
        if not s or s.startswith("#"):
            continue
        t = s.split()
        if len(t) < 10:
            raise ModelFormatError(f"images.txt: bad image line {s[:60]!r}")
        pts = lines[i].split() if i < len(lines) else []  # the second line may be empty (image without points)
        i += 1
        if len(pts) % 3:
            raise ModelFormatError(f"images.txt: image {t[0]}: 2D point list is not a multiple of 3 values")
        a = np.array(pts, np.float64).reshape(-1, 3) if pts else np.zeros((0, 3))
        ids = np.array(pts[2::3], np.int64) if pts else np.zeros(0, np.int64)
        out[int(t[0])] = ImageRec(int(t[0]), np.array(t[1:5], np.float64), np.array(t[5:8], np.float64), int(t[8]),
                                  " ".join(t[9:]), a[:, :2].copy(), ids)
    return out


def read_points3d_txt(path):
    ids, xyz, rgb, err, tl, ti, tp = [], [], [], [], [], [], []
    for s in _data_lines(path):
        t = s.split()
        if len(t) < 8 or (len(t) - 8) % 2:
            raise ModelFormatError(f"points3D.txt: bad line {s[:60]!r}")
        ids.append(int(t[0]))
        xyz.append([float(v) for v in t[1:4]])
        rgb.append([int(v) for v in t[4:7]])
        err.append(float(t[7]))
        tr = np.array(t[8:], np.int64).reshape(-1, 2)
        tl.append(len(tr))
        ti.append(tr[:, 0])
        tp.append(tr[:, 1])
    n = len(ids)
    ptr = np.concatenate([[0], np.cumsum(tl)]).astype(np.int64)
    return (np.array(ids, np.int64), np.array(xyz, np.float64).reshape(n, 3), np.array(rgb, np.uint8).reshape(n, 3),
            np.array(err, np.float64), ptr,
            (np.concatenate(ti) if n else np.zeros(0)).astype(np.int32),
            (np.concatenate(tp) if n else np.zeros(0)).astype(np.int32))


def write_cameras_txt(cameras, path) -> None:
    with open(path, "w") as f:
        f.write("# Camera list with one line of data per camera:\n#   CAMERA_ID, MODEL, WIDTH, HEIGHT, PARAMS[]\n"
                f"# Number of cameras: {len(cameras)}\n")
        for c in cameras.values():
            f.write(" ".join([str(c.camera_id), c.model, str(c.width), str(c.height)] + [repr(float(p)) for p in c.params]) + "\n")


def write_images_txt(images, path) -> None:
    mean = (sum(len(im.point3D_ids) for im in images.values()) / len(images)) if images else 0
    with open(path, "w") as f:
        f.write("# Image list with two lines of data per image:\n#   IMAGE_ID, QW, QX, QY, QZ, TX, TY, TZ, CAMERA_ID, NAME\n"
                f"#   POINTS2D[] as (X, Y, POINT3D_ID)\n# Number of images: {len(images)}, mean observations per image: {mean}\n")
        for im in images.values():
            f.write(" ".join([str(im.image_id)] + [repr(float(v)) for v in (*im.qvec, *im.tvec)] + [str(im.camera_id), im.name]) + "\n")
            f.write(" ".join(f"{float(x)!r} {float(y)!r} {int(p)}" for (x, y), p in zip(im.xys, im.point3D_ids)) + "\n")


def write_points3d_txt(model: SparseModel, path) -> None:
    n = len(model.point_ids)
    mean = (int(model.track_ptr[-1]) / n) if n else 0
    with open(path, "w") as f:
        f.write("# 3D point list with one line of data per point:\n"
                "#   POINT3D_ID, X, Y, Z, R, G, B, ERROR, TRACK[] as (IMAGE_ID, POINT2D_IDX)\n"
                f"# Number of points: {n}, mean track length: {mean}\n")
        for i in range(n):
            a, b = int(model.track_ptr[i]), int(model.track_ptr[i + 1])
            head = [str(int(model.point_ids[i]))] + [repr(float(v)) for v in model.xyz[i]] + \
                   [str(int(v)) for v in model.rgb[i]] + [repr(float(model.error[i]))]
            tr = [f"{int(g)} {int(p)}" for g, p in zip(model.track_image_ids[a:b], model.track_point2D_idxs[a:b])]
            f.write(" ".join(head + tr) + "\n")


# ---------------------------------------------------------------------- whole models
def read_colmap_model(sparse_dir, binary: Optional[bool] = None) -> SparseModel:
    """`load_colmap_dataset` (loader.py:137-150) without the image files: widths/heights come from cameras.*
    (the reference opens every image with PIL only to read its size, loader.py:46-49,67-68).
    binary=None: use the .bin files when all three exist, else .txt."""
    j = lambda n: os.path.join(sparse_dir, n)  # noqa: E731
    if binary is None:
        binary = all(os.path.isfile(j(n)) for n in ("cameras.bin", "images.bin", "points3D.bin"))
    if binary:
        cams, imgs, pts = read_cameras_bin(j("cameras.bin")), read_images_bin(j("images.bin")), read_points3d_bin(j("points3D.bin"))
    else:
        cams, imgs, pts = read_cameras_txt(j("cameras.txt")), read_images_txt(j("images.txt")), read_points3d_txt(j("points3D.txt"))
    return SparseModel(cams, imgs, *pts)


def write_colmap_model(model: SparseModel, sparse_dir, binary=True) -> None:
    """`export_in_colmap_format` (loader.py:153-212)."""
    os.makedirs(sparse_dir, exist_ok=True)
    j = lambda n: os.path.join(sparse_dir, n)  # noqa: E731
    if binary:
        write_cameras_bin(model.cameras, j("cameras.bin"))
        write_images_bin(model.images, j("images.bin"))
        write_points3d_bin(model, j("points3D.bin"))
    else:
        write_cameras_txt(model.cameras, j("cameras.txt"))
        write_images_txt(model.images, j("images.txt"))
        write_points3d_txt(model, j("points3D.txt"))


def model_from_ba_arrays(poses, intrinsics, points, cam_idx, pt_idx, obs, width=1000, height=1000,
                         names: Optional[List[str]] = None) -> SparseModel:
    """A COLMAP model from solver arrays (synthetic scenes, BAL files): image/camera ids 1..C, point ids 1..P,
    2D points in row order per camera, PINHOLE cameras."""
    poses, intrinsics, points = np.asarray(poses), np.asarray(intrinsics), np.asarray(points)
    cam_idx, pt_idx, obs = np.asarray(cam_idx), np.asarray(pt_idx), np.asarray(obs, np.float64).reshape(-1, 2)
    C, P = len(poses), len(points)
    order = np.argsort(cam_idx, kind="stable")
    ptr = np.concatenate([[0], np.cumsum(np.bincount(cam_idx, minlength=C))])
    cams, imgs = {}, {}
    for c in range(C):
        rows = order[ptr[c]:ptr[c + 1]]
        cams[c + 1] = CameraRec(c + 1, "PINHOLE", int(width), int(height), np.array(intrinsics[c, :4], np.float64))
        imgs[c + 1] = ImageRec(c + 1, matrix_to_quat_wxyz(poses[c, :, :3]), np.array(poses[c, :, 3], np.float64), c + 1,
                               names[c] if names else f"image_{c + 1:06d}.jpg", obs[rows], pt_idx[rows].astype(np.int64) + 1)
    m = SparseModel(cams, imgs, np.arange(1, P + 1, dtype=np.int64), np.array(points, np.float64),
                    np.full((P, 3), 255, np.uint8), np.full(P, 999.0))  # defaults of loader.py:198-199
    m.rebuild_tracks()
    return m


# ---------------------------------------------------------------------- BAL ("Bundle Adjustment in the Large")
def read_bal(path, avg_cam_width=1000.0) -> dict:
    """BAL problem file -> solver arrays in the reference's camera model.  File: `C P N`, N lines
    `cam pt x y`, then 9 values per camera (Rodrigues w, t, f, k1, k2), then 3 per point.  BAL projects
    `p = -Pc.xy/Pc.z`, `u = f (1 + k1|p|^2 + k2|p|^4) p`; the reference's pinhole looks down +z without
    distortion (bundle_adjustment.py:17-21).  Mapped exactly for k1 = k2 = 0 by `[R|t] <- diag(1,-1,-1)[R|t]`,
    `y_obs <- -y_obs`, fx = fy = f, cx = cy = skew = 0; k1, k2 are returned separately and not modelled."""
    with open(path, "r") as f:
        tok = f.read().split()
    if len(tok) < 3:
        raise ModelFormatError("BAL: missing header")
    C, P, N = int(tok[0]), int(tok[1]), int(tok[2])
    need = 3 + 4 * N + 9 * C + 3 * P
    if len(tok) != need:
        raise ModelFormatError(f"BAL: expected {need} values for C={C} P={P} N={N}, file has {len(tok)}")
    o = np.array(tok[3:3 + 4 * N], np.float64).reshape(N, 4)
    cam = np.array(tok[3 + 4 * N:3 + 4 * N + 9 * C], np.float64).reshape(C, 9)
    pts = np.array(tok[3 + 4 * N + 9 * C:], np.float64).reshape(P, 3)
    cam_idx, pt_idx = o[:, 0].astype(np.int32), o[:, 1].astype(np.int32)
    if N and (cam_idx.min() < 0 or cam_idx.max() >= C or pt_idx.min() < 0 or pt_idx.max() >= P):
        raise ModelFormatError("BAL: observation index out of range")
    S = np.diag([1.0, -1.0, -1.0])
    poses = np.empty((C, 3, 4))
    for c in range(C):
        poses[c, :, :3] = S @ _rodrigues(cam[c, :3])
        poses[c, :, 3] = S @ cam[c, 3:6]
    intr = np.zeros((C, 5))
    intr[:, 0] = intr[:, 1] = cam[:, 6]
    obs = o[:, 2:4].copy()
    obs[:, 1] = -obs[:, 1]
    return dict(poses0=poses, intrinsics0=intr, points0=pts, cam_idx=cam_idx, pt_idx=pt_idx, obs=obs,
                distortion=cam[:, 7:9].copy(), avg_cam_width=float(avg_cam_width))


def write_bal(path, poses, intrinsics, points, cam_idx, pt_idx, obs, distortion=None) -> None:
    """Inverse of `read_bal` (fx is written as f; cx, cy, skew must be 0 for an exact round trip)."""
    poses, intrinsics, points = np.asarray(poses), np.asarray(intrinsics), np.asarray(points)
    C, P, N = len(poses), len(points), len(cam_idx)
    S = np.diag([1.0, -1.0, -1.0])
    d = np.zeros((C, 2)) if distortion is None else np.asarray(distortion)
    obs = np.asarray(obs, np.float64).reshape(N, 2)
    with open(path, "w") as f:
        f.write(f"{C} {P} {N}\n")
        for c, p, (x, y) in zip(cam_idx, pt_idx, obs):
            f.write(f"{int(c)} {int(p)} {float(x)!r} {float(-y)!r}\n")
        for c in range(C):
            w = _log_so3(S @ poses[c, :, :3])
            t = S @ poses[c, :, 3]
            for v in (*w, *t, intrinsics[c, 0], *d[c]):
                f.write(f"{float(v)!r}\n")
        for X in points:
            for v in X:
                f.write(f"{float(v)!r}\n")


def _rodrigues(w):
    th = np.linalg.norm(w)
    if th < 1e-300:
        return np.eye(3)
    k = w / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.cos(th) * np.eye(3) + (1 - np.cos(th)) * np.outer(k, k) + np.sin(th) * K


def _log_so3(R):
    q = matrix_to_quat_wxyz(R)
    s = np.linalg.norm(q[1:])
    if s < 1e-300:
        return np.zeros(3)
    return q[1:] / s * (2.0 * np.arctan2(s, q[0]))
