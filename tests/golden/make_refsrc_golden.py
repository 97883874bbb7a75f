"""Golden vectors from the reference's OWN source files for the residual side of the hot path, executed in the build
container with NumPy standing in for jax.numpy (tests/golden/jax_numpy_shim.py; jax/jaxopt are not installable here):
  * `JaxBundleAdjustment.prepare_params` + `BundleAdjustment.get_residuals`  (bundle_adjustment.py:52-74,93-102)
    on the dense-padded contract of the `small` scene (pad = first observation, mask 0);
  * `JaxPoseOptimizer.prepare_params` + `PoseOptimization.get_residuals` for both loss functions
    (pose_optimization.py:33-52,73-84) on a ragged, masked batch;
  * `pose_mat_to_vec` and `rot_mat_from_vec` (utils.py:21-44);
  * the constructor arguments the two call sites hand to jaxopt.LevenbergMarquardt.
Run from the repo root:  python tests/golden/make_refsrc_golden.py   -> tests/golden/refsrc_seed0.npz"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import jax_numpy_shim  # noqa: E402

jax_numpy_shim.install()
sys.path.insert(1, "/root/reference")
import b200ba  # noqa: E402
from src.reconstruction.bundle_adjustment.bundle_adjustment import JaxBundleAdjustment  # noqa: E402  (the reference)
from src.reconstruction.bundle_adjustment.loss import JaxLossFunction  # noqa: E402
from src.reconstruction.bundle_adjustment.pose_optimization import JaxPoseOptimizer  # noqa: E402
from src.reconstruction.bundle_adjustment.utils import pose_mat_to_vec, rot_mat_from_vec  # noqa: E402

out = {}
# ---- BA
sc = b200ba.synth.make_ba_scene("small", seed=0)
ref = JaxBundleAdjustment(sc.C, sc.avg_cam_width)
x0, k3 = ref.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
p2d, idx, mask = sc.to_dense_padded()
res = ref.ba.get_residuals(x0, p2d, idx, k3, mask)
out.update(ba_x0=np.asarray(x0), ba_k3=np.asarray(k3), ba_residual=np.asarray(res), ba_nmax=p2d.shape[1])
kw = ref.optimizer.kwargs
out.update(ba_lm_tol=kw["tol"], ba_lm_maxiter=kw["maxiter"], ba_lm_solver=str(kw.get("solver", "default")))
# ---- pose (ragged batch, zero-padded, mask 0 on the padding: benchmark_pose_optimization.py:99-105)
pb = b200ba.synth.make_pose_batch(B=6, npts=48, seed=2)
rng = np.random.default_rng(4)
pmask = pb.mask.copy()
pts, obs = pb.points.copy(), pb.obs.copy()
for b in range(pb.B):
    nv = int(rng.integers(24, 49))
    pmask[b, nv:] = 0.0
    pts[b, nv:] = 0.0
    obs[b, nv:] = 0.0
for name, lf in (("cauchy", JaxLossFunction.CAUCHY), ("l2", JaxLossFunction.L2)):
    po = JaxPoseOptimizer(pb.avg_cam_width, loss_fn=lf)
    q0, c3 = po.prepare_params(pb.poses0, pb.intrinsics0)
    with np.errstate(all="ignore"):  # padded rows divide 0/0 before the mask multiplies them (as in the reference)
        r = np.stack([po.po.get_residuals(q0[b], pts[b], obs[b], c3[b], pmask[b]) for b in range(pb.B)])
    out[f"pose_residual_{name}"] = r
    kw = po.optimizer.kwargs
out.update(pose_q0=np.asarray(q0), pose_c3=np.asarray(c3), pose_points=pts, pose_obs=obs, pose_mask=pmask,
           pose_lm_tol=kw["tol"], pose_lm_maxiter=kw["maxiter"], pose_lm_solver=str(kw.get("solver")))
# ---- rotation helpers
W = rng.normal(size=(20, 3)) * np.array([0.01, 1.0, 3.0, 0.3])[rng.integers(0, 4, 20)][:, None]
R = np.stack([rot_mat_from_vec(w) for w in W])
T = np.concatenate([R, rng.normal(size=(20, 3, 1))], axis=2)
out.update(rot_w=W, rot_R=R, log_in=T, log_out=np.stack([pose_mat_to_vec(t) for t in T]))
np.savez_compressed(os.path.join(HERE, "refsrc_seed0.npz"), **out)
print({k: np.shape(v) for k, v in out.items()})
print("nan rows in pose residual:", {k: int(np.isnan(v).sum()) for k, v in out.items() if k.startswith("pose_residual")})
