"""Default (fp32-stored rows, rtol 1e-10) vs fp64-row (rtol 1e-12) GPU trajectories at a large shape."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import b200ba
shape = sys.argv[1] if len(sys.argv) > 1 else "venice1778"
sc = b200ba.synth.make_ba_scene(shape, seed=0)
res = []
for f32, rtol in ((True, 1e-10), (False, 1e-12)):
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, jac_f32=f32, pcg_rtol=rtol)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    p, st = ba.optimize(x0, ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P), None, k3, None)
    res.append(st)
    print(f"f32={f32} its={st.iter_num} pcg={list(map(int, st.pcg_iters))} loss={st.loss:.12e} |g|={st.error:.3e}", flush=True)
a, b = res
n = min(len(a.loss_trajectory), len(b.loss_trajectory))
dev = np.abs(a.loss_trajectory[:n] - b.loss_trajectory[:n]) / b.loss_trajectory[:n]
print(shape, "max rel loss dev f32 vs f64:", dev.max(), ["%.1e" % d for d in dev])
