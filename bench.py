#!/usr/bin/env python
"""bench.py — BA LM iterations/s (+ Jacobian obs/s, pose solves/s) on B200.

A "step" is ONE Levenberg–Marquardt iteration of the bundle-adjustment hot path on the
synthetic Venice-1778 shape (1,778 cams / 993,923 pts / 5,001,946 obs): Schur preparation,
block-Jacobi PCG to rtol on the reduced camera system, back-substitution, trial cost, gain-ratio
decision and (on accept) re-linearisation (residual + Jacobian + Schur blocks).  Full LM solves
are run from the same noisy start (restarting when one converges) until exactly K iterations
have been timed.

  python bench.py --gpus N --steps K --warmup W             # our CUDA path
  python bench.py --impl reference --steps K --warmup W      # CPU restatement of the jaxopt path

One JSON line on stdout (rank 0).  See DESIGN.md §Measurement for the byte models.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("OMP_WAIT_POLICY", "PASSIVE")  # spinning OpenMP workers oversubscribe shared host cores
os.environ["NCCL_DEBUG"] = "WARN"  # NCCL prints its version banner on stdout at VERSION/INFO; stdout is one JSON line

METRIC = "ba_lm_iterations_per_s"
UNIT = "LM iterations/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=14)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="venice1778")
    ap.add_argument("--jac-f64", action="store_true", help="store Jacobian rows in fp64 (default: fp32 rows, fp64 accumulation)")
    ap.add_argument("--pcg-rtol", type=float, default=1e-10)
    ap.add_argument("--no-pose", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=20.0)
    ap.add_argument("--seed", type=int, default=0)
    return ap.parse_args()


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------- byte models
def bytes_models(C, P, N, jsz):
    """Algorithmic bytes per launch of each kernel class (what the kernel MUST move given the
    stored-Jacobian formulation actually built; jsz = bytes per stored Jacobian value)."""
    return {
        # camera-run order: read obs 16 + point index 4 + A position 4; write Jc 8*jsz + r 8 (B order) and one 32 B
        # record {b, r} scattered to A order; points once (gathered inside the point block), cameras once
        "resjac": N * (16 + 4 + 4 + 8 * jsz + 8 + 32) + 24 * P + 88 * C,
        # point kernel: Jp 3*jsz + posB 4 + sigma 8 -> omega 8 ; per point Vinv 48 + ptr 4.  (No per-row point index: the
        # row -> point map is rebuilt in shared memory, PS_OWN in ba_pcg.cuh; earlier bench lines counted 4 B/row for it.)
        "mv_pt": N * (3 * jsz + 4 + 8 + 8) + 52 * P,
        # camera kernel: Jc 8*jsz + omega 8 -> sigma 8 ; per camera PCG record 576 r+w, U 512, Minv 512, run table
        "mv_cam": N * (8 * jsz + 8 + 8) + (2 * 576 + 1024 + 160) * C,
        "cost": N * (16 + 4) + 24 * P + 88 * C,
    }


# ---------------------------------------------------------------------------- CPU baseline (oracle)
def cpu_lm_iteration_sample(sc, x0, k3, cg_cap):
    """One LM iteration of the reference's formulation on ALL host cores (oracle/ba_oracle_c.c, OpenMP):
    residual + Jacobian, jaxopt-style unpreconditioned CG on the full normal equations TRUNCATED at
    `cg_cap` iterations (the tol=1e-5 solve needs hundreds), trial residual, gain ratio, and (on
    accept) re-linearisation.  Because the CG is cut short the resulting iterations/s is an UPPER
    bound on the CPU path's speed."""
    from oracle import c_binding as cb
    t0 = time.perf_counter()
    res = cb.lm_solve(x0, sc.C, sc.P, sc.cam_idx, sc.pt_idx, sc.obs, k3, sc.avg_cam_width, maxiter=1, cg_cap=cg_cap)
    return time.perf_counter() - t0, int(res["cg_iters"][0]), float(res["loss"][-1])


def cpu_threads():
    from oracle import c_binding as cb
    return cb.num_threads()


def run_reference(args):
    """--impl reference: the reference's CPU path. jax/jaxopt are not installable here (no wheels,
    no network), so this times the oracle port (oracle/ba_oracle_c.c: float64 C + OpenMP restatement of
    the reference's residual and jaxopt's LM with full-system CG) on all host cores."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    import b200ba
    sc = b200ba.synth.make_ba_scene(args.workload, seed=args.seed)
    from oracle import ba_oracle as bo
    x0, k3 = bo.pack_params(sc.poses0, sc.intrinsics0, sc.points0)
    cg_cap = 25
    for _ in range(min(args.warmup, 1)):
        cpu_lm_iteration_sample(sc, x0, k3, 2)
    ts = []
    for _ in range(args.steps):
        t, inner, rho = cpu_lm_iteration_sample(sc, x0, k3, cg_cap)
        ts.append(t)
    total = float(np.sum(ts))
    val = args.steps / total
    line = {"metric": METRIC, "value": val, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": args.workload, "C": sc.C, "P": sc.P, "N": sc.N},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cpu_threads(), "kind": "port",
                             "sample": f"{args.steps} LM iterations at {args.workload}: residual+Jacobian, "
                                       f"unpreconditioned CG truncated at {cg_cap} iterations (jaxopt tol=1e-5 "
                                       "needs hundreds), trial residual; upper bound on CPU it/s"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "host_cores": os.cpu_count()}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------- our arm
PC_NAMES = ["resjac", "pt_blocks", "cam_blocks", "pt_prepare", "cam_schur", "cam_finish", "mv_pt", "mv_cam",
            "pcg_vec", "backsub", "cost", "other"]


def run_b200(args):
    import ctypes as C

    import torch
    import torch.distributed as dist

    import b200ba

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sc = b200ba.synth.make_ba_scene(args.workload, seed=args.seed)
    ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, device=dev, jac_f32=not args.jac_f64,
                                     pcg_rtol=args.pcg_rtol)
    x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
    # pinned host copies (the e2e leg copies from these every solve)
    h_cam = torch.from_numpy(sc.cam_idx).pin_memory()
    h_pt = torch.from_numpy(sc.pt_idx).pin_memory()
    h_obs = torch.from_numpy(sc.obs).pin_memory()
    h_x0 = torch.from_numpy(x0).pin_memory()
    h_k3 = torch.from_numpy(k3).pin_memory()

    prob = ba.build_problem_sparse(h_cam, h_pt, h_obs, sc.P)
    Cn = sc.C
    xd = h_x0.to(dev)
    cam8_0 = torch.cat([xd[:6 * Cn].view(Cn, 6), xd[6 * Cn:8 * Cn].view(Cn, 2)], 1).contiguous()
    pts_0 = xd[8 * Cn:].view(-1, 3)[prob.pt_lo:prob.pt_hi].contiguous()
    k3d = h_k3.to(dev)

    def run_iters(n):
        """exactly n LM iterations: full solves from the noisy start, restarted on convergence"""
        done, info = 0, []
        while done < n:
            cam8, pts = cam8_0.clone(), pts_0.clone()
            res, traj = ba.solve_device(prob, cam8, k3d, pts, maxiter=min(100, n - done))
            done += res.iter_num
            info.append((res.iter_num, res.loss, res.grad_l2, res.n_pcg_iters_total, res.n_jac_evals, res.status))
        return info

    run_iters(max(args.warmup, 3))
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = b200ba.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    info = run_iters(args.steps)
    ev1.record()
    barrier()
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    launches = b200ba.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    value = args.steps / (ms * 1e-3)

    # ---- same region again with per-kernel CUDA-event timing (roofline numbers)
    lib = prob.lib
    lib.b200ba_profile_enable(prob.plan, 1)
    run_iters(args.steps)
    ms_cls = (C.c_double * 16)()
    cnt_cls = (C.c_int64 * 16)()
    lib.b200ba_profile_get(prob.plan, ms_cls, cnt_cls)
    lib.b200ba_profile_enable(prob.plan, 0)
    jsz = 8 if args.jac_f64 else 4
    bm = bytes_models(sc.C, prob.P, prob.N, jsz)
    peak, peak_src = load_peaks()
    kern = {}
    for i, nm in enumerate(PC_NAMES):
        if cnt_cls[i]:
            avg_us = 1e3 * ms_cls[i] / cnt_cls[i]
            kern[nm] = {"launches": int(cnt_cls[i]), "total_ms": round(ms_cls[i], 4), "avg_us": round(avg_us, 3)}
            if nm in bm:
                gbs = bm[nm] / (avg_us * 1e-6) / 1e9
                kern[nm].update({"alg_bytes": int(bm[nm]), "gbs": round(gbs, 1), "frac": round(gbs / peak, 4)})
    dom = max((k for k in kern if k in bm), key=lambda k: kern[k]["total_ms"])
    # dram__bytes_read+write per launch of the dominant kernel from the committed ncu --set full capture of this
    # workload (profiles/ncu_r1_v8_summary.md); other workloads / rank counts: not captured
    ncu_traffic = {"mv_pt": 208.1e6, "mv_cam": 232.9e6} if (args.workload == "venice1778" and world == 1 and not args.jac_f64) else {}
    roofline = {"bound": "hbm", "kernel": dom, "achieved": kern[dom]["gbs"], "peak": peak, "unit": "GB/s",
                "frac": kern[dom]["frac"], "traffic": ncu_traffic.get(dom), "peak_source": peak_src,
                "share_of_step": round(kern[dom]["total_ms"] / sum(v["total_ms"] for v in kern.values()), 4)}
    lin_ms = sum(kern[k]["total_ms"] for k in ("resjac", "pt_blocks", "cam_blocks") if k in kern)
    n_lin = kern["resjac"]["launches"] if "resjac" in kern else 0
    jac_obs_s = (sc.N * n_lin) / (lin_ms * 1e-3) if lin_ms else None

    # ---- end to end through the public API with HOST buffers (H2D of rows+params, D2H of params)
    e2e = None
    if True:
        # one untimed warm-up solve through the same API: the first plan built next to the resident one grows
        # the stream-ordered memory pool (a one-off of tens of ms that is not part of the steady state)
        pr = ba.build_problem_sparse(h_cam, h_pt, h_obs, sc.P)
        ba.optimize(h_x0, pr, None, h_k3, None)
        pr.close()
        barrier()
        t0 = time.perf_counter()
        its, h2d, d2h, nsolves = 0, 0, 0, 0
        while its < args.steps:
            pr = ba.build_problem_sparse(h_cam, h_pt, h_obs, sc.P)
            params, st = ba.optimize(h_x0, pr, None, h_k3, None)
            its += int(st.iter_num)
            nsolves += 1
            h2d += h_cam.numel() * 4 + h_pt.numel() * 4 + h_obs.numel() * 8 + h_x0.numel() * 8 + h_k3.numel() * 8
            d2h += params.nbytes
            pr.close()
        barrier()
        t_e2e = time.perf_counter() - t0
        t_e2e = max_over_ranks(t_e2e * 1e3) * 1e-3
        e2e = {"value": its / t_e2e, "unit": UNIT, "h2d_bytes_per_step": int(h2d / its),
               "d2h_bytes_per_step": int(d2h / its), "solves": nsolves, "lm_iterations": its,
               "includes": "after one untimed warm-up solve: H2D of index/observation rows and parameters from pinned memory, plan build "
                           "(GPU sort + workspace allocation), LM solve, D2H of optimised parameters"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64" if args.jac_f64 else "f64 (f32-stored Jacobian rows, f64 accumulation)",
            "data": "synthetic",
            "config": {"workload": args.workload, "C": sc.C, "P": sc.P, "N": sc.N, "parallelism": f"point-shard x{world}",
                       "pcg_rtol": args.pcg_rtol, "l2": "inputs exceed L2 (Jacobian rows %.2f GB per matvec)" %
                       ((bm["mv_pt"] + bm["mv_cam"]) / 1e9), "jacobian_rows": "fp64" if args.jac_f64 else "fp32"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
            "extra": {"solves": [{"iters": i[0], "loss": i[1], "grad_l2": i[2], "pcg_iters": i[3], "jac_evals": i[4],
                                  "status": i[5]} for i in info],
                      "jacobian_obs_per_s": jac_obs_s, "kernels": kern, "host_cores": os.cpu_count()}}

    if world == 1:
        # RMS reprojection error (pixels) before / after the last end-to-end solve, from the host copy of its result with the
        # package's NumPy projection helper (outside every timed region; never allowed to break the line)
        try:
            def _rms_px(xv):
                poses, intr, pts3 = b200ba.unpack_params(np.asarray(xv), k3, sc.C)
                uv = b200ba.synth.project_np(poses[sc.cam_idx], intr[sc.cam_idx], pts3[sc.pt_idx])
                return float(np.sqrt(((uv - sc.obs) ** 2).sum(axis=1).mean()))
            line["extra"]["rms_reprojection_px"] = {"initial": _rms_px(x0), "final": _rms_px(params),
                                                    "of": "last end-to-end solve (full LM solve from the noisy start)"}
        except Exception as exc:  # noqa: BLE001
            line["extra"]["rms_reprojection_px"] = {"error": repr(exc)}

    if not args.no_pose and world == 1:
        line["extra"]["pose"] = bench_pose(b200ba, dev, args)
    if not args.no_cpu_baseline and world == 1:
        t, inner, rho = cpu_lm_iteration_sample(sc, x0, k3, 25)
        line["cpu_baseline"] = {"value": 1.0 / t, "unit": UNIT, "cores": cpu_threads(), "kind": "port",
                                "sample": f"1 LM iteration at {args.workload} (residual+Jacobian, unpreconditioned CG "
                                          f"truncated at {inner} its, trial residual); upper bound on CPU it/s"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def bench_pose(b200ba, dev, args):
    import torch
    pb = b200ba.synth.make_pose_batch(B=4096, npts=512, seed=args.seed, rot_noise=5e-3, pos_noise=5e-3)
    opt = b200ba.B200PoseOptimizer(pb.avg_cam_width, device=dev)
    q0, c3 = opt.prepare_params(pb.poses0, pb.intrinsics0)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    q0d, c3d, ptd, obd, mkd = d(q0), d(c3), d(pb.points), d(pb.obs), d(pb.mask)
    for _ in range(3):
        opt.optimize_device(q0d.clone(), ptd, obd, c3d, mkd)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    e0.record()
    for _ in range(reps):
        q = q0d.clone()
        _, it, loss, err, mu, st = opt.optimize_device(q, ptd, obd, c3d, mkd)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    its = it.cpu().numpy()
    t0 = time.perf_counter()
    for _ in range(3):
        opt.optimize(q0, pb.points, pb.obs, c3, pb.mask)
    torch.cuda.synchronize()
    t_e2e = (time.perf_counter() - t0) / 3
    return {"workload": "4096 poses x 512 correspondences, Cauchy, pose noise 5e-3", "solves_per_s": 4096 / (ms * 1e-3),
            "pose_iterations_per_s": float(its.sum()) / (ms * 1e-3), "ms_per_batch": ms,
            "iters_mean": float(its.mean()), "iters_max": int(its.max()),
            "iters_histogram": [int(v) for v in np.bincount(its.astype(np.int64).clip(min=0))],  # [k] = poses that took k LM iterations
            "e2e_solves_per_s": 4096 / t_e2e, "e2e_h2d_bytes": int(q0.nbytes + c3.nbytes + pb.points.nbytes + pb.obs.nbytes + pb.mask.nbytes)}


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
