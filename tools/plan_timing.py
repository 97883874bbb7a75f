"""Where does plan construction / one end-to-end solve spend its host time?  GPU required."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200BA_VERBOSE"] = "1"
import numpy as np, torch
import b200ba
sc = b200ba.synth.make_ba_scene(sys.argv[1] if len(sys.argv) > 1 else "venice1778", seed=0)
ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
h = [torch.from_numpy(a).pin_memory() for a in (sc.cam_idx, sc.pt_idx, sc.obs)]
for rep in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    pr = ba.build_problem_sparse(h[0], h[1], h[2], sc.P)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    params, st = ba.optimize(x0, pr, None, k3, None)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    pr.close()
    torch.cuda.synchronize(); t3 = time.perf_counter()
    print(f"rep {rep}: build {1e3*(t1-t0):.1f} ms, solve {1e3*(t2-t1):.1f} ms ({st.iter_num} its), close {1e3*(t3-t2):.1f} ms", flush=True)
