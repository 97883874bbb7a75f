import sys, time; sys.path.insert(0,'.')
import numpy as np, torch, b200ba
from oracle import ba_oracle as bo
sc = b200ba.synth.make_ba_scene('small', seed=0)
ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, verbose=True)
x0,k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
p, st = ba.optimize(x0, prob, None, k3, None)
print(st.iter_num, st.loss_trajectory, st.pcg_iters)
ref = bo.lm_solve(x0, sc.C, sc.P, sc.cam_idx, sc.pt_idx, sc.obs, k3, sc.avg_cam_width)
print(ref.iter_num, ref.trace.loss)
