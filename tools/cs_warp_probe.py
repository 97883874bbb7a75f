"""Timing probe for the multi-rank kernel sequence on ONE GPU (B200BA_FORCE_SPLIT=1): per shape and task size,
average time of k_cs_warp (both passes), k_ps and k_cs_vec over 30 PCG iterations.
usage: python tools/cs_warp_probe.py "shape[,shape..]" "rows[,rows..]"   "unroll[,unroll..]"   (0 = the plan's own choice)"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200BA_FORCE_SPLIT"] = "1"
import torch  # noqa: E402

import b200ba  # noqa: E402

NAMES = ["resjac", "pt_blocks", "cam_blocks", "pt_prepare", "cam_schur", "cam_finish", "mv_pt", "mv_cam", "pcg_vec",
         "backsub", "cost", "other"]
shapes = sys.argv[1].split(",")
rows = sys.argv[2].split(",") if len(sys.argv) > 2 else ["0"]
unrolls = sys.argv[3].split(",") if len(sys.argv) > 3 else ["0"]
for shape in shapes:
    sc = b200ba.synth.make_ba_scene(shape, seed=0)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, jac_f32=True)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    xd = torch.from_numpy(x0).cuda()
    cam8 = torch.cat([xd[:6 * sc.C].view(sc.C, 6), xd[6 * sc.C:8 * sc.C].view(sc.C, 2)], 1).contiguous()
    pts = xd[8 * sc.C:].view(-1, 3).contiguous()
    k3d = torch.from_numpy(k3).cuda()
    for r, u in [(r, u) for u in unrolls for r in rows]:
        if u == "0":
            os.environ.pop("B200BA_CW_U", None)
        else:
            os.environ["B200BA_CW_U"] = u
        if r == "0":
            os.environ.pop("B200BA_TASK_ROWS", None)
        else:
            os.environ["B200BA_TASK_ROWS"] = r
        prob = ba.build_problem_sparse(sc.cam_idx, sc.pt_idx, sc.obs, sc.P)
        K = b200ba.BAKernels(prob, sc.avg_cam_width)
        K.linearize(cam8, k3d, pts)
        K.schur_prepare(1e-4)
        K.schur_solve(1e-30, 10)
        prob.lib.b200ba_profile_enable(prob.plan, 1)
        K.schur_prepare(1e-4)
        K.schur_solve(1e-30, 30)
        ms, cnt = (C.c_double * 16)(), (C.c_int64 * 16)()
        prob.lib.b200ba_profile_get(prob.plan, ms, cnt)
        prob.lib.b200ba_profile_enable(prob.plan, 0)
        out = {n: round(1e3 * ms[i] / cnt[i], 1) for i, n in enumerate(NAMES) if cnt[i] and n in ("mv_pt", "mv_cam", "pcg_vec")}
        print(f"{shape} U={u} task_rows={r} lib={os.path.basename(os.environ.get('B200BA_LIB', 'default'))} {out} launches={cnt[7]}", flush=True)
        prob.close()
