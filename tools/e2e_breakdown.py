"""Where does one end-to-end solve (host buffers in, host parameters out) spend its time?  Stage timers with a stream
synchronisation after every stage (so the sum is an upper bound of the pipelined total).  GPU required."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import b200ba

sc = b200ba.synth.make_ba_scene(sys.argv[1] if len(sys.argv) > 1 else "venice1778", seed=0)
ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
h = [torch.from_numpy(a).pin_memory() for a in (sc.cam_idx, sc.pt_idx, sc.obs, x0, k3)]
dev = ba.device
sync = torch.cuda.synchronize


def stage(acc, name, t0):
    sync()
    t1 = time.perf_counter()
    acc[name] = acc.get(name, 0.0) + 1e3 * (t1 - t0)
    return t1


for rep in range(4):
    acc = {}
    sync(); t = time.perf_counter(); t_all = t
    d = [b200ba.api._dev(a, dt, dev) for a, dt in zip(h[:3], (torch.int32, torch.int32, torch.float64))]
    t = stage(acc, "h2d_rows", t)
    prob = b200ba.BAProblem(sc.C, sc.P, d[0], d[1], d[2], dev, True)
    prob.pt_lo, prob.pt_hi, prob.P_global = 0, sc.P, sc.P
    t = stage(acc, "plan_create+set_obs", t)
    xd = h[3].to(dev, non_blocking=True)
    Cn = sc.C
    cam8 = torch.cat([xd[:6 * Cn].view(Cn, 6), xd[6 * Cn:8 * Cn].view(Cn, 2)], dim=1).contiguous()
    pts = xd[8 * Cn:].view(-1, 3).contiguous().clone()
    k3d = h[4].to(dev, non_blocking=True)
    t = stage(acc, "h2d_params", t)
    res, traj = ba.solve_device(prob, cam8, k3d, pts)
    t = stage(acc, "lm_solve", t)
    out = torch.cat([cam8[:, :6].reshape(-1), cam8[:, 6:].reshape(-1), pts.reshape(-1)]).cpu().numpy()
    t = stage(acc, "d2h_params", t)
    prob.close()
    t = stage(acc, "plan_destroy", t)
    tot = 1e3 * (time.perf_counter() - t_all)
    print(f"rep {rep}: total {tot:.1f} ms, {res.iter_num} its | " + ", ".join(f"{k} {v:.2f}" for k, v in acc.items()), flush=True)
