"""CPU tests of the real-data ingest/egress: COLMAP sparse models and BAL files.
Fixtures in tests/golden/colmap_small/ were written by the reference's own `read_write_model.py`
(tests/golden/make_colmap_golden.py); nothing here needs a GPU or /root/reference."""
import filecmp
import os

import numpy as np
import pytest

import b200ba
from b200ba import model_io as mio

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "colmap_small")


@pytest.fixture(scope="module")
def exp():
    return np.load(os.path.join(GOLD, "expected.npz"))


def _check_model(m, exp, exact):
    cmp = np.testing.assert_array_equal if exact else (lambda a, b: np.testing.assert_allclose(a, b, rtol=1e-15, atol=0))
    assert list(m.images) == list(exp["img_ids"])  # file order kept
    assert {k: c.model for k, c in m.cameras.items()} == {2: "PINHOLE", 9: "SIMPLE_RADIAL", 4: "SIMPLE_PINHOLE"}
    assert m.cameras[9].width == 1090 and m.cameras[9].height == 800 and len(m.cameras[9].params) == 4
    for c, iid in enumerate(exp["img_ids"]):
        im = m.images[int(iid)]
        assert im.camera_id == exp["cam_ids"][c]
        assert im.name == (f"dir {c}/img_{c}.jpg" if c == 2 else f"img_{c}.jpg")
        cmp(im.xys, exp[f"xys_{c}"])
        np.testing.assert_array_equal(im.point3D_ids, exp[f"p3d_{c}"])
        cmp(im.qvec, exp[f"q_{c}"])
        cmp(im.tvec, exp[f"t_{c}"])
    np.testing.assert_array_equal(m.point_ids, exp["pt_ids"])
    cmp(m.xyz, exp["xyz"])
    np.testing.assert_array_equal(m.rgb, exp["rgb"])
    cmp(m.error, exp["err"])
    # tracks agree with the images' 2D->3D links
    assert m.track_ptr[-1] == sum(int((exp[f"p3d_{c}"] >= 0).sum()) for c in range(5))
    for i, pid in enumerate(m.point_ids):
        for g, k in zip(m.track_image_ids[m.track_ptr[i]:m.track_ptr[i + 1]], m.track_point2D_idxs[m.track_ptr[i]:m.track_ptr[i + 1]]):
            assert m.images[int(g)].point3D_ids[int(k)] == pid


def test_read_reference_written_binary_model(exp):
    _check_model(mio.read_colmap_model(GOLD, binary=True), exp, exact=True)
    assert mio.read_colmap_model(GOLD).point_ids.shape == (40,)  # auto-detects .bin


def test_read_reference_written_text_model(exp):
    _check_model(mio.read_colmap_model(GOLD, binary=False), exp, exact=True)  # str(float) round-trips exactly


def test_binary_writer_is_byte_identical_to_reference_writer(tmp_path):
    m = mio.read_colmap_model(GOLD, binary=True)
    mio.write_colmap_model(m, tmp_path, binary=True)
    for n in ("cameras.bin", "images.bin", "points3D.bin"):
        assert filecmp.cmp(os.path.join(GOLD, n), tmp_path / n, shallow=False), n


def test_text_writer_roundtrip(tmp_path, exp):
    m = mio.read_colmap_model(GOLD, binary=True)
    mio.write_colmap_model(m, tmp_path, binary=False)
    _check_model(mio.read_colmap_model(tmp_path, binary=False), exp, exact=True)


def test_rebuild_tracks_matches_file_tracks():
    m = mio.read_colmap_model(GOLD, binary=True)
    ref = [sorted(zip(m.track_image_ids[a:b].tolist(), m.track_point2D_idxs[a:b].tolist()))
           for a, b in zip(m.track_ptr[:-1], m.track_ptr[1:])]
    m.rebuild_tracks()
    got = [sorted(zip(m.track_image_ids[a:b].tolist(), m.track_point2D_idxs[a:b].tolist()))
           for a, b in zip(m.track_ptr[:-1], m.track_ptr[1:])]
    assert got == ref


def test_to_ba_arrays_follows_reference_adapter(exp):
    """`_prepare_dataset` semantics (benchmark_bundle_adjustment.py:45-106): camera c = c-th image, points = observed
    identifiers ascending, W2C [R|t], K from get_intrinsics (loader.py:33-43), avg width = mean image width."""
    m = mio.read_colmap_model(GOLD)
    a = m.to_ba_arrays()
    C = 5
    n_obs = [int((exp[f"p3d_{c}"] >= 0).sum()) for c in range(C)]
    assert a["poses0"].shape == (C, 3, 4) and a["intrinsics0"].shape == (C, 5)
    assert len(a["cam_idx"]) == sum(n_obs) == 130 and a["cam_idx"].dtype == np.int32 and a["pt_idx"].dtype == np.int32
    np.testing.assert_array_equal(np.bincount(a["cam_idx"], minlength=C), n_obs)
    assert np.all(np.diff(a["point_ids"]) > 0) and set(a["point_ids"]) == set(exp["pt_ids"])
    order = np.argsort(exp["pt_ids"])
    np.testing.assert_array_equal(a["points0"], exp["xyz"][order])
    for c in range(C):
        rows = a["cam_idx"] == c
        keep = exp[f"p3d_{c}"] >= 0
        np.testing.assert_array_equal(a["obs"][rows], exp[f"xys_{c}"][keep])
        np.testing.assert_array_equal(a["point_ids"][a["pt_idx"][rows]], exp[f"p3d_{c}"][keep])
        np.testing.assert_array_equal(a["point2D_idx"][rows], np.flatnonzero(keep))
        np.testing.assert_allclose(a["poses0"][c, :, 3], exp[f"t_{c}"])
        R = a["poses0"][c, :, :3]
        np.testing.assert_allclose(R @ R.T, np.eye(3), atol=1e-14)
        np.testing.assert_allclose(mio.matrix_to_quat_wxyz(R), exp[f"q_{c}"], atol=1e-14)
    # SIMPLE_RADIAL / SIMPLE_PINHOLE share f; skew 0
    assert a["intrinsics0"][2, 0] == a["intrinsics0"][2, 1] and np.all(a["intrinsics0"][:, 4] == 0)
    widths = [m.cameras[int(k)].width for k in exp["cam_ids"]]
    assert a["avg_cam_width"] == pytest.approx(np.mean(widths))
    # the arrays are a valid BA problem for the host facade's packing
    ba_pack = b200ba.B200BundleAdjustment.__new__(b200ba.B200BundleAdjustment)
    ba_pack.cam_num = C
    x0, k3 = b200ba.B200BundleAdjustment.prepare_params(ba_pack, a["poses0"], a["intrinsics0"], a["points0"])
    assert x0.shape == (8 * C + 3 * len(a["points0"]),) and k3.shape == (C, 3)


def test_quaternion_matrix_roundtrip_against_scipy():
    from scipy.spatial.transform import Rotation
    rng = np.random.default_rng(0)
    for _ in range(50):
        q = rng.normal(size=4)
        R = mio.quat_wxyz_to_matrix(q)
        Rs = Rotation.from_quat([q[1], q[2], q[3], q[0]]).as_matrix()  # camera_pose.py:211
        np.testing.assert_allclose(R, Rs, atol=1e-14)
        q2 = mio.matrix_to_quat_wxyz(R)
        qn = q / np.linalg.norm(q) * (1 if q[0] >= 0 else -1)
        np.testing.assert_allclose(q2, qn, atol=1e-13)
    # near-180-degree rotations take the trace<=0 branch
    for ax in np.eye(3):
        R = Rotation.from_rotvec(ax * (np.pi - 1e-9)).as_matrix()
        np.testing.assert_allclose(mio.quat_wxyz_to_matrix(mio.matrix_to_quat_wxyz(R)), R, atol=1e-12)


def test_solution_export_roundtrip(tmp_path):
    m = mio.read_colmap_model(GOLD)
    a = m.to_ba_arrays()
    rng = np.random.default_rng(1)
    pts = a["points0"] + rng.normal(size=a["points0"].shape) * 1e-3
    poses = a["poses0"].copy()
    poses[:, :, 3] += 0.01
    intr = a["intrinsics0"] + 1.0
    out = m.with_solution(poses, intr, pts, a["point_ids"])
    mio.write_colmap_model(out, tmp_path, binary=True)
    b = mio.read_colmap_model(tmp_path).to_ba_arrays()
    np.testing.assert_allclose(b["points0"], pts, rtol=0, atol=0)
    np.testing.assert_allclose(b["poses0"], poses, atol=1e-14)
    np.testing.assert_allclose(b["intrinsics0"][:, :4], intr[:, :4])
    np.testing.assert_array_equal(b["cam_idx"], a["cam_idx"])
    np.testing.assert_array_equal(b["pt_idx"], a["pt_idx"])
    assert all(c.model == "PINHOLE" for c in out.cameras.values()) and list(out.cameras) == [1, 2, 3, 4, 5]


def test_model_from_synthetic_scene_roundtrip(tmp_path):
    sc = b200ba.synth.make_ba_scene("small", seed=0)
    m = mio.model_from_ba_arrays(sc.poses0, sc.intrinsics0, sc.points0, sc.cam_idx, sc.pt_idx, sc.obs)
    for binary in (True, False):
        d = tmp_path / ("bin" if binary else "txt")
        mio.write_colmap_model(m, d, binary=binary)
        a = mio.read_colmap_model(d, binary=binary).to_ba_arrays()
        assert len(a["cam_idx"]) == sc.N and a["avg_cam_width"] == 1000.0
        np.testing.assert_array_equal(a["points0"], sc.points0)
        np.testing.assert_allclose(a["poses0"], sc.poses0, atol=1e-13)
        # same multiset of rows (the model stores them camera-major)
        key = lambda c, p, o: sorted(zip(c.tolist(), p.tolist(), o[:, 0].tolist(), o[:, 1].tolist()))  # noqa: E731
        assert key(a["cam_idx"], a["pt_idx"], a["obs"]) == key(sc.cam_idx, sc.pt_idx, sc.obs)


def test_truncated_and_malformed_files_raise(tmp_path):
    for n in ("cameras.bin", "images.bin", "points3D.bin"):
        raw = open(os.path.join(GOLD, n), "rb").read()
        (tmp_path / n).write_bytes(raw[:len(raw) - 5])
    with pytest.raises(mio.ModelFormatError):
        mio.read_cameras_bin(tmp_path / "cameras.bin")
    with pytest.raises(mio.ModelFormatError):
        mio.read_images_bin(tmp_path / "images.bin")
    with pytest.raises(mio.ModelFormatError):
        mio.read_points3d_bin(tmp_path / "points3D.bin")
    (tmp_path / "cameras.txt").write_text("1 NOT_A_MODEL 10 10 1 2 3\n")
    with pytest.raises(mio.ModelFormatError):
        mio.read_cameras_txt(tmp_path / "cameras.txt")
    # empty model
    empty = mio.SparseModel({}, {}, np.zeros(0, np.int64), np.zeros((0, 3)), np.zeros((0, 3), np.uint8), np.zeros(0))
    mio.write_colmap_model(empty, tmp_path / "e", binary=True)
    e = mio.read_colmap_model(tmp_path / "e")
    assert len(e.cameras) == 0 and len(e.images) == 0 and len(e.point_ids) == 0
    with pytest.raises(mio.ModelFormatError):
        e.to_ba_arrays()


def _bal_project(cam9, X):
    """BAL's published camera model (Agarwal et al., 'Bundle Adjustment in the Large')."""
    R = mio._rodrigues(cam9[:3])
    P = R @ X + cam9[3:6]
    p = -P[:2] / P[2]
    n2 = p @ p
    return cam9[6] * (1 + cam9[7] * n2 + cam9[8] * n2 * n2) * p


def test_bal_roundtrip_and_camera_model(tmp_path):
    rng = np.random.default_rng(5)
    C, P = 4, 30
    cam9 = np.concatenate([rng.normal(size=(C, 3)) * 0.1, rng.normal(size=(C, 3)) * 0.1 + [0, 0, -8.0],
                           rng.uniform(500, 900, (C, 1)), np.zeros((C, 2))], axis=1)
    X = rng.uniform(-1, 1, (P, 3))
    cam_idx = np.repeat(np.arange(C), P).astype(np.int32)
    pt_idx = np.tile(np.arange(P), C).astype(np.int32)
    uv = np.array([_bal_project(cam9[c], X[p]) for c, p in zip(cam_idx, pt_idx)])
    path = tmp_path / "problem.txt"
    with open(path, "w") as f:
        f.write(f"{C} {P} {len(cam_idx)}\n")
        for c, p, (x, y) in zip(cam_idx, pt_idx, uv):
            f.write(f"{c} {p}     {float(x)!r} {float(y)!r}\n")
        f.write("\n".join(repr(float(v)) for v in cam9.reshape(-1)) + "\n")
        f.write("\n".join(repr(float(v)) for v in X.reshape(-1)) + "\n")
    a = mio.read_bal(path)
    # exact observations of the BAL model are exact observations of the reference's pinhole after the mapping
    proj = b200ba.synth.project_np(a["poses0"][a["cam_idx"]], a["intrinsics0"][a["cam_idx"]], a["points0"][a["pt_idx"]])
    np.testing.assert_allclose(proj, a["obs"], atol=1e-9)
    assert np.all((a["poses0"][a["cam_idx"], 2, :3] * a["points0"][a["pt_idx"]]).sum(1) + a["poses0"][a["cam_idx"], 2, 3] > 0)
    out = tmp_path / "again.txt"
    mio.write_bal(out, a["poses0"], a["intrinsics0"], a["points0"], a["cam_idx"], a["pt_idx"], a["obs"], a["distortion"])
    b = mio.read_bal(out)
    np.testing.assert_allclose(b["poses0"], a["poses0"], atol=1e-13)
    np.testing.assert_array_equal(b["obs"], a["obs"])
    np.testing.assert_array_equal(b["points0"], a["points0"])
    with pytest.raises(mio.ModelFormatError):
        path.write_text("2 2 1\n0 0 1.0 2.0\n")
        mio.read_bal(path)


@pytest.mark.parametrize("tag,limits", [("full", (None, None)), ("reduced", (4, 12))])
def test_adapter_arrays_equal_the_reference_adapter(tag, limits):
    """tests/golden/adapter_colmap_small.npz holds what the reference's OWN `_prepare_dataset`
    (benchmark_bundle_adjustment.py:45-106, incl. `Dataset.make_reduced_dataset`) returned for a reference Dataset built
    from the same fixture files (tests/golden/make_adapter_golden.py)."""
    g = np.load(os.path.join(os.path.dirname(GOLD), "adapter_colmap_small.npz"))
    a = mio.read_colmap_model(GOLD).reduced(*limits).to_ba_arrays()
    C = len(a["poses0"])
    np.testing.assert_array_equal(a["point_ids"], g[f"{tag}_point_ids"])
    np.testing.assert_array_equal(a["points0"], g[f"{tag}_p3d"])
    np.testing.assert_allclose(a["poses0"], g[f"{tag}_poses"], atol=1e-14)
    K = g[f"{tag}_K"]
    np.testing.assert_array_equal(a["intrinsics0"], np.stack([K[:, 0, 0], K[:, 1, 1], K[:, 0, 2], K[:, 1, 2], K[:, 0, 1]], axis=1))
    assert a["avg_cam_width"] == float(g[f"{tag}_avg_cam_width"])
    p2d, idx, mask = mio.to_dense_padded(a["cam_idx"], a["pt_idx"], a["obs"], C)
    np.testing.assert_array_equal(p2d, g[f"{tag}_p2d"])
    np.testing.assert_array_equal(idx, g[f"{tag}_idx"])
    np.testing.assert_array_equal(mask, g[f"{tag}_mask"])
    # and the synthetic scenes' own dense adapter is the same rule
    sc = b200ba.synth.make_ba_scene("tiny", seed=3)
    for x, y in zip(sc.to_dense_padded(), mio.to_dense_padded(sc.cam_idx, sc.pt_idx, sc.obs, sc.C)):
        np.testing.assert_array_equal(x, y)


def test_pose_batch_equals_the_reference_pose_adapter():
    """`JaxoptSinglePoseBenchmarkBatched.setup/_prepare_points/_get_mask` (benchmark_pose_optimization.py:32-130), executed
    on a reference Dataset built from the same fixture files, against `SparseModel.to_pose_batch`."""
    g = np.load(os.path.join(os.path.dirname(GOLD), "adapter_colmap_small.npz"))
    b = mio.read_colmap_model(GOLD).to_pose_batch()
    np.testing.assert_allclose(b["poses0"], g["pose_poses"], atol=1e-14)
    K = g["pose_K"]
    np.testing.assert_array_equal(b["intrinsics0"], np.stack([K[:, 0, 0], K[:, 1, 1], K[:, 0, 2], K[:, 1, 2], K[:, 0, 1]], axis=1))
    np.testing.assert_array_equal(b["points"], g["pose_points"])
    np.testing.assert_array_equal(b["observations"], g["pose_obs"])
    np.testing.assert_array_equal(b["mask"], g["pose_mask"])
    assert b["avg_cam_width"] == float(g["pose_avg_cam_width"]) and (b["mask"] == 0).any()
