"""Jacobians of the reference's OWN `get_residuals` functions (bundle_adjustment.py:52-74, pose_optimization.py:33-52):
the source files are executed on float64 torch tensors (tests/golden/jax_numpy_shim.py, install("torch")) and
differentiated by torch.autograd — the role jax's jvp/vjp play inside jaxopt.  Pins the oracle's analytic Jacobian.
Run from the repo root:  python tests/golden/make_refsrc_jac_golden.py  -> tests/golden/refsrc_jac_seed1.npz"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import jax_numpy_shim  # noqa: E402

jax_numpy_shim.install("torch")
sys.path.insert(1, "/root/reference")
import b200ba  # noqa: E402
from oracle import ba_oracle as bo  # noqa: E402
from src.reconstruction.bundle_adjustment.bundle_adjustment import BundleAdjustment  # noqa: E402  (the reference)
from src.reconstruction.bundle_adjustment.loss import JaxLossFunction  # noqa: E402
from src.reconstruction.bundle_adjustment.pose_optimization import PoseOptimization  # noqa: E402

T = lambda a: torch.tensor(np.asarray(a))  # noqa: E731
out = {}
sc = b200ba.synth.make_ba_scene("tiny", seed=1)
x0, k3 = bo.pack_params(sc.poses0, sc.intrinsics0, sc.points0)  # packing itself is pinned by refsrc_seed0.npz
p2d, idx, mask = sc.to_dense_padded()
ba = BundleAdjustment(sc.C, sc.avg_cam_width ** 2)
f = lambda p: ba.get_residuals(p, T(p2d), T(idx), T(k3), T(mask))  # noqa: E731
out.update(ba_x0=x0, ba_residual=f(T(x0)).numpy(), ba_jacobian=torch.autograd.functional.jacobian(f, T(x0)).numpy())

pb = b200ba.synth.make_pose_batch(B=3, npts=16, seed=4, ragged=True)
q0 = np.concatenate([bo.pose_mat_to_vec(pb.poses0), pb.intrinsics0[:, :2]], axis=1)
c3 = pb.intrinsics0[:, 2:]
for name, lf in (("cauchy", JaxLossFunction.CAUCHY), ("l2", JaxLossFunction.L2)):
    po = PoseOptimization(pb.avg_cam_width ** 2, lf)
    J, R = [], []
    for b in range(pb.B):
        g = lambda p: po.get_residuals(p, T(pb.points[b]), T(pb.obs[b]), T(c3[b]), T(pb.mask[b]))  # noqa: E731
        R.append(g(T(q0[b])).numpy())
        J.append(torch.autograd.functional.jacobian(g, T(q0[b])).numpy())
    out[f"pose_residual_{name}"], out[f"pose_jacobian_{name}"] = np.stack(R), np.stack(J)
out.update(pose_q0=q0)
np.savez_compressed(os.path.join(HERE, "refsrc_jac_seed1.npz"), **out)
print({k: np.shape(v) for k, v in out.items()}, "nan:", sum(int(np.isnan(v).sum()) for v in out.values()))
