"""ORACLE (test infrastructure, NOT product code) — literal float64 torch restatement of the
reference's two residual functions in their ORIGINAL dense-padded array contract, used only to
validate `ba_oracle.py` / `pose_oracle.py` (sparse rows, analytic Jacobians) by autodiff.
jax is not installed here, so torch.float64 on CPU plays jax.numpy's role.  Its residuals are themselves pinned to the
reference's source executed under a jax.numpy stand-in (tests/golden/refsrc_seed0.npz).

  ba_get_residuals    <- BundleAdjustment.get_residuals, bundle_adjustment.py:52-74
                         (reproject_point :17-21, double vmap :24-29)
  pose_get_residuals  <- PoseOptimization.get_residuals, pose_optimization.py:33-52
  rot_mat_from_vec    <- utils.py:21-28 ; parse_intrinsics utils.py:47-57
  l2_loss/cauchy_loss <- loss.py:7-14
"""
import torch


def rot_mat_from_vec(w):
    theta = torch.linalg.norm(w)
    r = w / theta
    eye = torch.eye(3, dtype=w.dtype)
    r_rT = torch.outer(r, r)
    r_cross = torch.cross(eye, r.expand(3, 3), dim=1)
    return torch.cos(theta) * eye + (1 - torch.cos(theta)) * r_rT + torch.sin(theta) * r_cross


def parse_cam_pose(v):
    return torch.cat([rot_mat_from_vec(v[:3]), v[3:6, None]], dim=1)


def parse_intrinsics(f, cx_cy_skew):
    z = torch.zeros((), dtype=f.dtype)
    o = torch.ones((), dtype=f.dtype)
    return torch.stack([
        torch.stack([f[0], cx_cy_skew[2], cx_cy_skew[0]]),
        torch.stack([z, f[1], cx_cy_skew[1]]),
        torch.stack([z, z, o]),
    ])


def l2_loss(y, x):
    return (y - x) ** 2


def cauchy_loss(y, x, scale=1):
    return torch.log(1 + l2_loss(y, x) ** scale)


def ba_get_residuals(opt_params, points_2d_all, p3d_indices_all, cx_cy_skew, masks_all, cam_num,
                     avg_cam_width_sqr):
    C = cam_num
    cam = opt_params[:6 * C].reshape(-1, 6)
    f = opt_params[6 * C:8 * C].reshape(-1, 2)
    points_3d = opt_params[8 * C:].reshape(-1, 3)
    poses = torch.stack([parse_cam_pose(cam[c]) for c in range(C)])
    intr = torch.stack([parse_intrinsics(f[c], cx_cy_skew[c]) for c in range(C)])
    KE = torch.einsum("bij,bjk->bik", intr, poses)
    X = points_3d[p3d_indices_all]                                  # (C,Nmax,3)
    proj = torch.einsum("cij,cnj->cni", KE[:, :, :3], X) + KE[:, None, :, 3]
    proj = proj[..., :2] / proj[..., 2:3]
    err = (l2_loss(proj, points_2d_all) * masks_all[..., None]).sum(-1)
    return err.flatten() / avg_cam_width_sqr


def pose_get_residuals(params, points, observations, cx_cy_skew, mask, avg_cam_width_sqr,
                       loss_fn=cauchy_loss):
    pose = parse_cam_pose(params)
    intr = parse_intrinsics(params[6:8], cx_cy_skew)
    KE = intr @ pose
    p = torch.einsum("ij,bj->bi", KE[:, :3], points) + KE[:, 3]
    p = p[..., :2] / p[..., 2:3]
    res = loss_fn(observations, p)
    return res.sum(dim=1) * mask / avg_cam_width_sqr
