"""Generates the golden fixtures in this directory from the oracle (the reference itself cannot be
imported: jax / jaxopt are absent).  Run from the repo root:
    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import b200ba  # noqa: E402
from oracle import ba_oracle as bo  # noqa: E402
from oracle import pose_oracle as po  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

for shape, seed in [("small", 0), ("ladybug49", 0), ("trafalgar257", 0)]:
    sc = b200ba.synth.make_ba_scene(shape, seed=seed)
    x0, ccs = bo.pack_params(sc.poses0, sc.intrinsics0, sc.points0)
    res = bo.lm_solve(x0, sc.C, sc.P, sc.cam_idx, sc.pt_idx, sc.obs, ccs, sc.avg_cam_width)
    rms = bo.rms_reprojection_px(res.params, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, ccs)
    out = dict(x0=x0 if shape == "small" else x0[:64], iter_num=res.iter_num, loss=np.array(res.trace.loss),
               grad_l2=np.array(res.trace.grad_l2), mu=np.array(res.trace.mu), accepted=np.array(res.trace.accepted),
               rms_px=rms, params_head=res.params[:64])
    np.savez_compressed(os.path.join(HERE, f"ba_{shape}_seed{seed}.npz"), **out)
    print(shape, res.iter_num, rms)

pb = b200ba.synth.make_pose_batch(B=16, npts=512, seed=3, rot_noise=5e-3, pos_noise=5e-3)
x0 = np.concatenate([bo.pose_mat_to_vec(pb.poses0), pb.intrinsics0[:, :2]], axis=1)
res = po.pose_lm_batched(x0, pb.points, pb.obs, pb.intrinsics0[:, 2:], pb.mask, pb.avg_cam_width)
np.savez_compressed(os.path.join(HERE, "pose_b16_seed3.npz"), iter_num=res.iter_num, loss=res.loss, params=res.params,
                    grad_l2=res.grad_l2)
print("pose", res.iter_num)
