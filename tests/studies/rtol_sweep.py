"""Parity vs PCG tolerance / Jacobian storage: max relative deviation of the per-iteration cost from
the exact-solve oracle, and PCG iterations spent.  GPU required."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import b200ba
from oracle import ba_oracle as bo

for shape in sys.argv[1:] or ["small", "ladybug49"]:
    sc = b200ba.synth.make_ba_scene(shape, seed=0)
    x0, k3 = bo.pack_params(sc.poses0, sc.intrinsics0, sc.points0)
    t = time.time()
    ref = bo.lm_solve(x0, sc.C, sc.P, sc.cam_idx, sc.pt_idx, sc.obs, k3, sc.avg_cam_width)
    rl = np.array(ref.trace.loss)
    print(f"{shape}: oracle {ref.iter_num} its in {time.time()-t:.1f}s", flush=True)
    for f32 in (True,):
        for rtol in (1e-7, 1e-8, 1e-9, 1e-10, 1e-11):
            ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, jac_f32=f32, pcg_rtol=rtol)
            prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
            p, st = ba.optimize(x0, prob, None, k3, None)
            n = min(len(rl), len(st.loss_trajectory))
            dev = np.abs(st.loss_trajectory[:n] - rl[:n]) / rl[:n]
            rms = bo.rms_reprojection_px(p, sc.C, sc.cam_idx, sc.pt_idx, sc.obs, k3)
            print(f"  f32={int(f32)} rtol={rtol:.0e} its={st.iter_num} pcg={list(st.pcg_iters)} maxdev={dev.max():.2e} "
                  f"dev={['%.1e' % d for d in dev]} rms={rms:.8f}", flush=True)
