"""Frozen exact-oracle trajectory for the benched config, Venice-1778 (1 778 cams / 993 923 pts / 5 001 946 obs):
    python tests/golden/make_venice_golden.py            # ~10 min on 8 host cores, ~6 GB
The LM state machine of oracle/ba_oracle.py with the inner system solved EXACTLY: reduced camera system formed densely
in C (oracle/ba_oracle_c.c::ba_c_schur_dense, 14 224^2 fp64) + LAPACK Cholesky (`lm_solve_dense_c`).  The reference
itself cannot run (jax / jaxopt absent) and could not hold this shape anyway (dense C x Nmax padding).
Consumers: tests/test_gpu_ba.py::test_venice_config_trajectory_golden and bench.py's `extra.parity` (every N)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import b200ba  # noqa: E402
from oracle import ba_oracle as bo  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
shape, seed = (sys.argv[1] if len(sys.argv) > 1 else "venice1778"), 0
sc = b200ba.synth.make_ba_scene(shape, seed=seed)
x0, ccs = bo.pack_params(sc.poses0, sc.intrinsics0, sc.points0)
t0 = time.time()
res = bo.lm_solve_dense_c(x0, sc.C, sc.P, sc.cam_idx, sc.pt_idx, sc.obs, ccs, sc.avg_cam_width, verbose=True)
rms0 = bo.rms_reprojection_px(x0, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, ccs)
rms = bo.rms_reprojection_px(res.params, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, ccs)
np.savez_compressed(os.path.join(HERE, f"ba_{shape}_seed{seed}.npz"), x0=x0[:64], iter_num=res.iter_num,
                    loss=np.array(res.trace.loss), grad_l2=np.array(res.trace.grad_l2), mu=np.array(res.trace.mu),
                    accepted=np.array(res.trace.accepted), rho=np.array(res.trace.rho), rms_px=rms, rms_px_initial=rms0,
                    params_head=res.params[:64], params_cam=res.params[:8 * sc.C])
print(shape, res.iter_num, rms0, rms, f"{time.time() - t0:.0f} s")
