"""us per LM iteration of the dense path on the small scenes as a function of the cooperative grid sizes of the tile
Cholesky / triangular sweeps (B200BA_DN_GRID / B200BA_DN_GRID_V).  GPU required."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import json, os, sys, time
sys.path.insert(0, %r)
import torch, b200ba
for shape in ("ref15x400", "ladybug49"):
    sc = b200ba.synth.make_ba_scene(shape, seed=0)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
    Cn = sc.C
    xd = torch.from_numpy(x0).cuda(); k3d = torch.from_numpy(k3).cuda()
    def once():
        cam8 = torch.cat([xd[:6 * Cn].view(Cn, 6), xd[6 * Cn:8 * Cn].view(Cn, 2)], 1).contiguous()
        pts = xd[8 * Cn:].view(-1, 3).contiguous().clone()
        return ba.solve_device(prob, cam8, k3d, pts)
    for _ in range(5): res, _t = once()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    its = 0; e0.record()
    for _ in range(40):
        res, _t = once(); its += res.iter_num
    e1.record(); torch.cuda.synchronize()
    print(json.dumps({"env": {k: os.environ.get(k) for k in ("B200BA_DN_GRID", "B200BA_DN_GRID_V")}, "shape": shape,
                      "us_per_lm_iteration": 1e3 * e0.elapsed_time(e1) / its, "iters": res.iter_num, "loss": res.loss}), flush=True)
''' % ROOT
for g, gv in (("", ""), ("1", "1"), ("1", ""), ("2", "1"), ("4", "2")):
    env = dict(os.environ)
    if g: env["B200BA_DN_GRID"] = g
    if gv: env["B200BA_DN_GRID_V"] = gv
    subprocess.run([sys.executable, "-c", CHILD], env=env)
