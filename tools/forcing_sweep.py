"""Trajectory deviation from the exact-oracle goldens under the inexact-Newton forcing schedule of b200ba_lm_solve
(rtol_k = min(max, pcg_rtol (|g_k| / tol)^p), csrc/b200ba.cu forcing_rtol), PCG path on every shape (GPU required):
    python tools/forcing_sweep.py "0:1e-6" "0.5:1e-6" ...   -> one JSON line per (schedule, shape)"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import json, os, sys
sys.path.insert(0, %r)
import numpy as np
import b200ba
G = os.path.join(%r, "tests", "golden")
for shape in ("small", "ladybug49", "trafalgar257", "venice1778"):
    gold = np.load(os.path.join(G, f"ba_{shape}_seed0.npz"))
    sc = b200ba.synth.make_ba_scene(shape, seed=0)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
    params, st = ba.optimize(x0, prob, None, k3, None)
    prob.close()
    n = min(len(st.loss_trajectory), len(gold["loss"]))
    dev = float(np.max(np.abs(st.loss_trajectory[:n] - gold["loss"][:n]) / gold["loss"][:n]))
    print(json.dumps({"forcing": os.environ.get("B200BA_FORCING"), "shape": shape, "iters": int(st.iter_num),
                      "golden_iters": int(gold["iter_num"]), "pcg_iters": int(st.pcg_iters.sum()),
                      "pcg_per_it": [int(v) for v in st.pcg_iters], "max_rel_loss_dev": dev}), flush=True)
''' % (ROOT, ROOT)
for f in sys.argv[1:]:
    env = dict(os.environ, B200BA_FORCING=f, B200BA_DENSE="0")
    subprocess.run([sys.executable, "-c", CHILD], env=env)
