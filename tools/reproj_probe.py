"""Times the reprojection-error evaluator (k_reproj_err) with CUDA events on the launching stream and
prints one JSON line with its HBM-roofline fraction.  GPU required.  usage: python tools/reproj_probe.py [shape] [reps]
Algorithmic bytes per launch: N·(16 obs + 4 pt + 4 posA + 4 order + 8 err) + 24·P + 88·C + 8·C."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import b200ba

shape = sys.argv[1] if len(sys.argv) > 1 else "venice1778"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
sc = b200ba.synth.make_ba_scene(shape, seed=0)
ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
xd = torch.from_numpy(x0).cuda()
cam8 = torch.cat([xd[:6 * sc.C].view(sc.C, 6), xd[6 * sc.C:8 * sc.C].view(sc.C, 2)], 1).contiguous()
pts = xd[8 * sc.C:].view(-1, 3).contiguous()
k3d = torch.from_numpy(k3).cuda()
K = b200ba.BAKernels(prob, sc.avg_cam_width)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > L2 (126 MB)
peak = 6538.6
try:
    peak = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
alg = sc.N * (16 + 4 + 4 + 4 + 8) + 24 * sc.P + 96 * sc.C
out = {"shape": shape, "N": sc.N, "alg_bytes": alg, "peak_gbs": peak}
for code, name in ((0, "trivial"), (1, "cauchy"), (2, "ceres_cauchy")):
    for _ in range(3):
        K.reprojection_errors(cam8, k3d, pts, code)
    ms = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        err, sums = K.reprojection_errors(cam8, k3d, pts, code)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    us = 1e3 * float(np.median(ms))  # includes the two output allocations of the facade (stream-ordered, no sync)
    out[name] = {"us": us, "gbs": alg / us / 1e3, "frac": alg / us / 1e3 / peak, "obs_per_s": sc.N / us * 1e6,
                 "rms_px": float(torch.sqrt(err.mean())) if code == 0 else None}
print(json.dumps(out))
