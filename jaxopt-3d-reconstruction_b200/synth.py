"""Seeded synthetic scenes for the BA and single-pose hot paths.

The reference ships no data (datasets are git-ignored, `.gitignore:162-165`), so every
workload here is generated: points U[-1,1]^3, cameras on a sphere of radius ~10 looking at
the centroid (W2C `[R|t]`, COLMAP right-handed frame, `loader.py:84-86`), 1000x1000 images,
fx=fy=1200 (`colmap/src/optim/bundle_adjustment_test.cc:131-132`), cx=cy=500, skew=0, exact
projection, then the noise model of `Dataset.with_noise` (`src/dataset/dataset.py:43-63`):
every perturbation is `magnitude * (positive-orthant random unit direction)`
(`dataset.py:32-41` uses `rand`, not `randn`).

Pure NumPy, vectorised, deterministic for a given seed.  Observations are rounded ONCE to
float32 and returned as float64 holding those rounded values, so the oracle and the CUDA
kernels consume bit-identical inputs.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

# (C, P, N) of BASELINE.json's configs
SHAPES = {
    "tiny": (3, 8, 20),
    "small": (15, 300, 1500),
    "ref15x400": (15, 400, 2000),   # the reference's own BA benchmark size (benchmark_impl_bundle_adjustment.py:28-32)
    "ladybug49": (49, 7776, 31843),
    "trafalgar257": (257, 65132, 225911),
    "venice1778": (1778, 993923, 5001946),
    "final13682": (13682, 4456117, 28987644),
    # what ONE rank of a point-sharded Final-13682 solve holds at 4 / 8 ranks (all cameras, 1/4 or 1/8 of the points):
    # single-GPU stand-ins for profiling the multi-rank kernels (B200BA_FORCE_SPLIT=1)
    "final13682_r2": (13682, 2228058, 14493822),
    "final13682_r4": (13682, 1114029, 7246911),
    "final13682_r8": (13682, 557015, 3623456),
    "venice1778_r2": (1778, 496962, 2500973),
    "venice1778_r4": (1778, 248481, 1250487),
    "venice1778_r8": (1778, 124240, 625243),
}

WIDTH = 1000.0
FOCAL = 1.2 * 1000.0


def rodrigues_np(w):
    """(...,3) rotation vectors -> (...,3,3); same formula as `utils.py:21-28` (no guard)."""
    w = np.asarray(w, dtype=np.float64)
    theta = np.linalg.norm(w, axis=-1)[..., None, None]
    k = w / theta[..., 0]
    K = np.zeros(w.shape[:-1] + (3, 3))
    K[..., 0, 1] = -k[..., 2]
    K[..., 0, 2] = k[..., 1]
    K[..., 1, 0] = k[..., 2]
    K[..., 1, 2] = -k[..., 0]
    K[..., 2, 0] = -k[..., 1]
    K[..., 2, 1] = k[..., 0]
    eye = np.broadcast_to(np.eye(3), K.shape)
    kk = k[..., :, None] * k[..., None, :]
    return np.cos(theta) * eye + (1.0 - np.cos(theta)) * kk + np.sin(theta) * K


def _pos_dir(rng, n, d):
    r = rng.random((n, d))
    return r / np.linalg.norm(r, axis=1, keepdims=True)


def _look_at_cameras(rng, C, radius=10.0):
    """C world->camera poses on a sphere looking at the origin. Returns (C,3,4)."""
    v = rng.standard_normal((C, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    centre = v * (radius * (1.0 + 0.05 * rng.random((C, 1))))
    z = -centre / np.linalg.norm(centre, axis=1, keepdims=True)  # optical axis -> origin
    up = np.tile(np.array([0.0, 0.0, 1.0]), (C, 1))
    par = np.abs((z * up).sum(1)) > 0.95
    up[par] = np.array([0.0, 1.0, 0.0])
    x = np.cross(up, z)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    y = np.cross(z, x)
    R = np.stack([x, y, z], axis=1)  # rows = camera axes in world coords -> W2C rotation
    # small roll so no rotation is exactly degenerate
    roll = rng.uniform(-0.3, 0.3, C)
    cr, sr = np.cos(roll), np.sin(roll)
    Rz = np.zeros((C, 3, 3))
    Rz[:, 0, 0] = cr
    Rz[:, 0, 1] = -sr
    Rz[:, 1, 0] = sr
    Rz[:, 1, 1] = cr
    Rz[:, 2, 2] = 1.0
    R = Rz @ R
    t = -(R @ centre[..., None])[..., 0]
    return np.concatenate([R, t[..., None]], axis=2)


def apply_pose_noise(poses, move, rotvec):
    """W2C poses (C,3,4) after `apply_move(move)` then `apply_transform_3d(R(rotvec))` as `Dataset.with_noise` INTENDS
    them (`dataset.py:50-58`): the camera centre is displaced by `move`, then `[R|t] <- Rn [R|t]`.
    Measured against the reference's own `CameraPose` (tests/golden/make_noise_golden.py): `apply_transform_3d`
    (`camera_pose.py:93-95`) is exactly the second step; `apply_move` (`camera_pose.py:101-107`, "TODO: Check validity")
    is a NO-OP on W2C poses — it restores the translation it just changed — so on COLMAP-loaded datasets the
    reference's noise never moves a camera centre.  The synthetic workloads keep the intended 5e-2 displacement (a
    harder start); `pos_noise=0` reproduces the reference's actual behaviour."""
    R = poses[:, :, :3]
    t = poses[:, :, 3]
    t = t - (R @ move[..., None])[..., 0]  # camera centre moved by `move`
    Rn = rodrigues_np(rotvec)
    R = Rn @ R
    t = (Rn @ t[..., None])[..., 0]
    return np.concatenate([R, t[..., None]], axis=2)


def _noisy_poses(rng, poses, rot_noise, pos_noise):
    C = poses.shape[0]
    d = _pos_dir(rng, C, 3) * pos_noise
    return apply_pose_noise(poses, d, _pos_dir(rng, C, 3) * rot_noise)


def project_np(poses, intr, X):
    """poses (...,3,4), intr (...,5)=[fx,fy,cx,cy,skew], X (...,3) -> (...,2)."""
    q = (poses[..., :, :3] @ X[..., None])[..., 0] + poses[..., :, 3]
    u = (intr[..., 0] * q[..., 0] + intr[..., 4] * q[..., 1]) / q[..., 2] + intr[..., 2]
    v = intr[..., 1] * q[..., 1] / q[..., 2] + intr[..., 3]
    return np.stack([u, v], axis=-1)


def _coprime_steps(rng, C, n):
    cands = np.array([g for g in range(1, max(C, 2)) if math.gcd(g, C) == 1] or [1])
    return cands[rng.integers(0, len(cands), n)]


@dataclass
class BAScene:
    """Sparse BA problem: one row per observation (unordered), noisy initial estimates."""
    C: int
    P: int
    N: int
    poses0: np.ndarray       # (C,3,4) noisy W2C
    intrinsics0: np.ndarray  # (C,5) noisy [fx,fy,cx,cy,skew]
    points0: np.ndarray      # (P,3) noisy
    cam_idx: np.ndarray      # (N,) int32
    pt_idx: np.ndarray       # (N,) int32
    obs: np.ndarray          # (N,2) float64 holding float32-rounded pixels
    avg_cam_width: float
    poses_gt: np.ndarray
    intrinsics_gt: np.ndarray
    points_gt: np.ndarray

    def to_dense_padded(self):
        """The reference adapter's layout (`benchmark_bundle_adjustment.py:45-106`):
        `(C,Nmax,2)` obs / `(C,Nmax)` idx / mask; pad = camera's FIRST observation, mask 0."""
        order = np.argsort(self.cam_idx, kind="stable")
        cam_s = self.cam_idx[order]
        counts = np.bincount(cam_s, minlength=self.C)
        nmax = int(counts.max())
        start = np.concatenate([[0], np.cumsum(counts)[:-1]])
        slot = np.arange(self.N) - start[cam_s]
        p2d = np.zeros((self.C, nmax, 2))
        idx = np.zeros((self.C, nmax), dtype=np.int64)
        mask = np.zeros((self.C, nmax))
        first = order[start]
        p2d[:] = self.obs[first][:, None, :]
        idx[:] = self.pt_idx[first][:, None]
        p2d[cam_s, slot] = self.obs[order]
        idx[cam_s, slot] = self.pt_idx[order]
        mask[cam_s, slot] = 1.0
        return p2d, idx, mask


def make_ba_scene(shape="ladybug49", seed=0, point_noise=3e-2, rot_noise=5e-2, pos_noise=5e-2,
                  intr_noise=10.0, px_noise=1.0, C=None, P=None, N=None) -> BAScene:
    if C is None:
        C, P, N = SHAPES[shape]
    assert N >= 2 * P and C >= 2
    rng = np.random.default_rng(seed)
    pts = rng.uniform(-1.0, 1.0, (P, 3))
    poses = _look_at_cameras(rng, C)
    intr = np.tile(np.array([FOCAL, FOCAL, WIDTH / 2, WIDTH / 2, 0.0]), (C, 1))

    # visibility: every point seen by >=2 distinct cameras, degrees sum to N exactly
    deg = 2 + rng.multinomial(N - 2 * P, np.full(P, 1.0 / P))
    over = deg > C
    while over.any():  # only possible for toy shapes
        excess = int((deg[over] - C).sum())
        deg[over] = C
        room = np.flatnonzero(deg < C)
        add = rng.choice(room, excess)
        np.add.at(deg, add, 1)
        over = deg > C
    pt_idx = np.repeat(np.arange(P, dtype=np.int64), deg)
    k = np.arange(N) - np.repeat(np.cumsum(deg) - deg, deg)
    start = rng.integers(0, C, P)
    step = _coprime_steps(rng, C, P)
    cam_idx = (start[pt_idx] + k * step[pt_idx]) % C

    uv = project_np(poses[cam_idx], intr[cam_idx], pts[pt_idx])
    uv = uv + _pos_dir(rng, N, 2) * px_noise
    obs = uv.astype(np.float32).astype(np.float64)

    points0 = pts + _pos_dir(rng, P, 3) * point_noise
    poses0 = _noisy_poses(rng, poses, rot_noise, pos_noise)
    intr0 = intr.copy()
    intr0 += rng.random((C, 5)) * intr_noise  # fx,fy,cx,cy,skew each +U[0,10) (`camera.py:30-47`)

    # shuffle rows so that no consumer may rely on an input ordering
    perm = rng.permutation(N)
    return BAScene(C, P, N, poses0, intr0, points0, cam_idx[perm].astype(np.int32),
                   pt_idx[perm].astype(np.int32), obs[perm], WIDTH, poses, intr, pts)


@dataclass
class PoseBatch:
    B: int
    Npts: int
    poses0: np.ndarray       # (B,3,4) noisy
    intrinsics0: np.ndarray  # (B,5) noisy
    points: np.ndarray       # (B,Npts,3) exact, zero-padded
    obs: np.ndarray          # (B,Npts,2) noisy pixels (float32-rounded), zero-padded
    mask: np.ndarray         # (B,Npts) {0.,1.}
    avg_cam_width: float


def make_pose_batch(B=4096, npts=512, seed=0, rot_noise=5e-3, pos_noise=5e-3, intr_noise=10.0,
                    px_noise=1.0, ragged=False) -> PoseBatch:
    """Independent camera + `npts` points in its frustum per pose.
    `ragged`: valid count U[npts/2, npts], zero padding with mask 0
    (`benchmark_pose_optimization.py:99-105`)."""
    rng = np.random.default_rng(seed)
    poses = _look_at_cameras(rng, B)
    intr = np.tile(np.array([FOCAL, FOCAL, WIDTH / 2, WIDTH / 2, 0.0]), (B, 1))
    pts = rng.uniform(-1.0, 1.0, (B, npts, 3))
    uv = project_np(poses[:, None], intr[:, None], pts)
    uv = uv + (_pos_dir(rng, B * npts, 2) * px_noise).reshape(B, npts, 2)
    obs = uv.astype(np.float32).astype(np.float64)
    mask = np.ones((B, npts))
    if ragged:
        valid = rng.integers(npts // 2, npts + 1, B)
        mask = (np.arange(npts)[None, :] < valid[:, None]).astype(np.float64)
        pts = pts * mask[..., None]
        obs = obs * mask[..., None]
    poses0 = _noisy_poses(rng, poses, rot_noise, pos_noise)
    intr0 = intr + rng.random((B, 5)) * intr_noise
    return PoseBatch(B, npts, poses0, intr0, pts, obs, mask, WIDTH)
