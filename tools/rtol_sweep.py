"""Trajectory deviation from the exact-oracle goldens as a function of the PCG tolerance (GPU required):
    python tools/rtol_sweep.py            -> one JSON line per (shape, rtol): LM iterations, PCG iterations, max rel. loss deviation
Decides the library's default pcg_rtol: the loosest tolerance that keeps every golden shape an order of magnitude inside the
1e-5 trajectory bar of north_star."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import b200ba

G = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
for shape in ("small", "ladybug49", "trafalgar257", "venice1778"):
    gold = np.load(os.path.join(G, f"ba_{shape}_seed0.npz"))
    sc = b200ba.synth.make_ba_scene(shape, seed=0)
    for rtol in (1e-10, 1e-9, 1e-8, 1e-7, 1e-6):
        ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, pcg_rtol=rtol)
        x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
        prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
        params, st = ba.optimize(x0, prob, None, k3, None)
        prob.close()
        n = min(len(st.loss_trajectory), len(gold["loss"]))
        dev = float(np.max(np.abs(st.loss_trajectory[:n] - gold["loss"][:n]) / gold["loss"][:n]))
        print(json.dumps({"shape": shape, "rtol": rtol, "iters": int(st.iter_num), "golden_iters": int(gold["iter_num"]),
                          "pcg_iters": int(st.pcg_iters.sum()), "max_rel_loss_dev": dev}), flush=True)
