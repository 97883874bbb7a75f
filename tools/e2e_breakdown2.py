"""End-to-end stages through the public API (pinned host buffers), synchronised after every stage: build (H2D rows + plan),
optimize (H2D parameters + LM solve + D2H), close; and the device-only solve for comparison.  GPU required."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import b200ba
sc = b200ba.synth.make_ba_scene(sys.argv[1] if len(sys.argv) > 1 else "venice1778", seed=0)
ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
h = [torch.from_numpy(a).pin_memory() for a in (sc.cam_idx, sc.pt_idx, sc.obs, x0, k3)]
sync = torch.cuda.synchronize
for rep in range(5):
    sync(); t0 = time.perf_counter()
    pr = ba.build_problem_sparse(h[0], h[1], h[2], sc.P)
    t1h = time.perf_counter(); sync(); t1 = time.perf_counter()
    params, st = ba.optimize(h[3], pr, None, h[4], None)
    t2h = time.perf_counter(); sync(); t2 = time.perf_counter()
    pr.close(); sync(); t3 = time.perf_counter()
    print(f"rep {rep}: total {1e3*(t3-t0):.2f} ms | build {1e3*(t1-t0):.2f} (host returns after {1e3*(t1h-t0):.2f}) | optimize {1e3*(t2-t1):.2f} "
          f"({st.iter_num} its) | close {1e3*(t3-t2):.2f}", flush=True)
# device-only solve on a resident plan
pr = ba.build_problem_sparse(h[0], h[1], h[2], sc.P)
xd = h[3].to(ba.device); Cn = sc.C
k3d = h[4].to(ba.device)
for rep in range(3):
    cam8 = torch.cat([xd[:6 * Cn].view(Cn, 6), xd[6 * Cn:8 * Cn].view(Cn, 2)], 1).contiguous()
    pts = xd[8 * Cn:].view(-1, 3).contiguous().clone()
    sync(); t0 = time.perf_counter()
    res, traj = ba.solve_device(pr, cam8, k3d, pts)
    sync(); t1 = time.perf_counter()
    print(f"device solve {1e3*(t1-t0):.2f} ms ({res.iter_num} its)")
# pieces of optimize
for rep in range(3):
    sync(); t0 = time.perf_counter()
    xd2 = h[3].to(ba.device, non_blocking=True); k3d2 = h[4].to(ba.device, non_blocking=True); sync(); t1 = time.perf_counter()
    host = torch.empty(xd2.shape, dtype=xd2.dtype, pin_memory=True); t2 = time.perf_counter()
    host.copy_(xd2, non_blocking=True); sync(); t3 = time.perf_counter()
    print(f"h2d params {1e3*(t1-t0):.2f} | pinned alloc {1e3*(t2-t1):.2f} | d2h {1e3*(t3-t2):.2f}")
