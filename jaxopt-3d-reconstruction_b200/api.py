"""Host-side mirror of the reference's optimiser façade (the drop-in boundary).

  B200BundleAdjustment  <->  JaxBundleAdjustment   (bundle_adjustment.py:77-129)
  B200PoseOptimizer     <->  JaxPoseOptimizer      (pose_optimization.py:55-99)

Same method names, argument meaning, parameter packing and error behaviour (the solver never
raises on NaN; it propagates into the outputs and the caller checks, `helpers.py:24-27,82-85`).
Arrays come in as NumPy (or torch) host arrays, are copied to the GPU, solved by the CUDA library
through its C-ABI, and copied back.  torch is used for device memory, streams and
torch.distributed only.  No CPU path: without a GPU / without libb200ba.so these raise.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional

import numpy as np
import torch

from . import _lib
from ._lib import LMOpts, LMResult, check

L2, CAUCHY = 0, 1


class LossFunction:
    """Stand-in for JaxLossFunction (loss.py:17-19)."""
    L2 = L2
    CAUCHY = CAUCHY
    TRIVIAL = L2  # dataset-side name of the same choice (src/dataset/loss_functions.py:15-17)
    CERES_CAUCHY = 2   # evaluator only: rho(|r|^2) of COLMAP's 2-vector residual (bundle_adjustment.cc:53-68)
    CERES_SOFT_L1 = 3


def pose_mat_to_vec(pose):
    """[R|t] (...,3,4) -> [w, t] (...,6); host NumPy like the reference (utils.py:31-44)."""
    pose = np.asarray(pose, dtype=np.float64)
    R = pose[..., :3, :3]
    n = np.zeros(R.shape[:-2] + (3,))
    angle = np.arccos(np.clip((R[..., 0, 0] + R[..., 1, 1] + R[..., 2, 2] - 1) / 2, -1, 1))
    n[..., 0] = R[..., 2, 1] - R[..., 1, 2]
    n[..., 1] = R[..., 0, 2] - R[..., 2, 0]
    n[..., 2] = R[..., 1, 0] - R[..., 0, 1]
    norms = np.linalg.norm(n, axis=-1)
    nz = norms != 0
    n[nz] *= (angle[nz] / norms[nz])[..., None]
    return np.concatenate([n, pose[..., :3, 3]], axis=-1)


def rot_vec_to_mat(w):
    """Rodrigues vector (...,3) -> R (...,3,3); host NumPy.  The reference's output side builds the same rotation with
    `Rotation.from_rotvec` (helpers.py:29,93); unlike `rot_mat_from_vec` (utils.py:21-28) it is finite at w = 0."""
    w = np.asarray(w, dtype=np.float64)
    th = np.linalg.norm(w, axis=-1)[..., None, None]
    safe = np.where(th > 0, th, 1.0)
    k = w[..., None, :] / safe
    K = np.zeros(w.shape[:-1] + (3, 3))
    K[..., 0, 1], K[..., 0, 2], K[..., 1, 2] = -w[..., 2], w[..., 1], -w[..., 0]
    K = (K - np.swapaxes(K, -1, -2)) / safe
    eye = np.broadcast_to(np.eye(3), K.shape)
    return np.cos(th) * eye + (1 - np.cos(th)) * (np.swapaxes(k, -1, -2) * k) + np.sin(th) * K


def unpack_params(params, cx_cy_skew, cam_num):
    """The inverse of `prepare_params`, i.e. the array part of `_parse_output_params_bundle`
    (src/benchmark/jaxopt_benchmark/helpers.py:59-115): packed `[C*6 | C*2 | P*3]` + cx_cy_skew (C,3) ->
    poses (C,3,4) W2C `[R|t]`, intrinsics (C,5) = [fx, fy, cx, cy, skew], points (P,3).  Raises on NaN camera
    parameters like the reference's parser (helpers.py:82-85) — the solver itself never raises."""
    x = np.asarray(params, dtype=np.float64)
    Cn = int(cam_num)
    cam6, f = x[:6 * Cn].reshape(Cn, 6), x[6 * Cn:8 * Cn].reshape(Cn, 2)
    if np.isnan(cam6).any() or np.isnan(f).any():
        raise ValueError("nan detected in optimised camera parameters")
    poses = np.concatenate([rot_vec_to_mat(cam6[:, :3]), cam6[:, 3:, None]], axis=2)
    return poses, np.concatenate([f, np.asarray(cx_cy_skew, dtype=np.float64).reshape(Cn, 3)], axis=1), x[8 * Cn:].reshape(-1, 3)


def _require_cuda():
    if not torch.cuda.is_available():
        raise _lib.B200BAError("b200ba needs a CUDA device (sm_100a); there is no CPU fallback")
    return _lib.load()


def _dev(a, dtype, device):
    """host/any array -> contiguous device tensor.  A pinned torch CPU tensor is copied asynchronously at full PCIe
    speed; a NumPy array (pageable) goes through the driver's staged copy."""
    if isinstance(a, torch.Tensor):
        return a.to(device=device, dtype=dtype, non_blocking=True).contiguous()
    t = torch.from_numpy(np.ascontiguousarray(a))
    if t.dtype != dtype:
        t = t.to(dtype)
    return t.to(device, non_blocking=True)


def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() else C.c_void_p(0)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


@dataclass
class LMState:
    """The fields of jaxopt's LevenbergMarquardtState that the reference's callers read
    (`benchmark_bundle_adjustment.py:183,196`; `benchmark_pose_optimization.py:190-191,237-239`)
    plus solver diagnostics."""
    iter_num: object
    loss: object
    error: object               # ||J^T r||_2
    damping_factor: object
    increase_factor: object = None
    status: object = None
    gradient: object = None
    loss_trajectory: object = None
    grad_trajectory: object = None
    mu_trajectory: object = None
    pcg_iters: object = None
    accepted: object = None
    n_jac_evals: int = 0
    n_res_evals: int = 0
    aux: dict = field(default_factory=dict)


class BAProblem:
    """Problem structure resident on the GPU: plan (sorted rows, CSR, workspaces) + observations."""

    def __init__(self, C_, P, cam_idx, pt_idx, obs, device, jac_f32=True, residual_model=0, obs_ready=None, robust=None):
        lib = _require_cuda()
        self.lib = lib
        self.C, self.P, self.N = int(C_), int(P), int(cam_idx.numel())
        self.device = device
        self.cam_idx = cam_idx
        self.pt_idx = pt_idx
        plan = C.c_void_p()
        with torch.cuda.device(device):
            check(lib.b200ba_plan_create(self.C, self.P, self.N, _ptr(cam_idx), _ptr(pt_idx),
                                         int(bool(jac_f32)), _stream(), C.byref(plan)), "plan_create")
            self.plan = plan
            if residual_model:
                check(lib.b200ba_plan_set_residual_model(plan, int(residual_model)), "set_residual_model")
            if robust is not None:  # (loss, scale): rows in (x, y) pairs with the full observation on both
                check(lib.b200ba_plan_set_robust_loss(plan, {"cauchy": 1, "soft_l1": 2}[robust[0]], float(robust[1])), "set_robust_loss")
            if obs_ready is not None:  # the observations were uploaded on a side stream while the plan was being built
                torch.cuda.current_stream().wait_event(obs_ready)
            check(lib.b200ba_plan_set_observations(plan, _ptr(obs), _stream()), "set_observations")
            torch.cuda.current_stream().synchronize()

    def close(self):
        if getattr(self, "plan", None):
            self.lib.b200ba_plan_destroy(self.plan)
            self.plan = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def shard_points(pt_idx, P: int, world: int):
    """Contiguous point ranges balanced by observation count.
    Returns bounds (world+1,) int64 such that rank r owns points [bounds[r], bounds[r+1]).  `pt_idx` may be a NumPy
    array (host) or a torch tensor (then the histogram / scan / search run on its device and only world+1 integers
    come back)."""
    if isinstance(pt_idx, torch.Tensor):
        counts = torch.bincount(pt_idx.to(torch.int64), minlength=P)
        cum = torch.cat([counts.new_zeros(1), torch.cumsum(counts, 0)])
        total = int(cum[-1])
        targets = torch.arange(world + 1, device=cum.device, dtype=torch.float64) * (total / world)
        bounds = torch.searchsorted(cum.to(torch.float64), targets, right=False).cpu().numpy().astype(np.int64)
    else:
        counts = np.bincount(np.asarray(pt_idx, dtype=np.int64), minlength=P)
        cum = np.concatenate([[0], np.cumsum(counts)])
        total = cum[-1]
        bounds = np.searchsorted(cum, np.arange(world + 1) * (total / world), side="left")
    bounds[0], bounds[-1] = 0, P
    return np.maximum.accumulate(np.minimum(bounds, P)).astype(np.int64)


def shard_rows(cam_idx, pt_idx, obs, P: int, world: int, rank: int):
    """One rank's share of a point-sharded problem, on the HOST (data placement, done once per problem): the rows whose
    point lies in this rank's range, point indices rebased to the range.  Returns
    (cam_idx, pt_idx_local, obs, (lo, hi), bounds) — feed them to `build_problem_sparse(..., pt_range=(lo, hi))`."""
    pt = np.asarray(pt_idx)
    bounds = shard_points(pt, P, world)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    keep = (pt >= lo) & (pt < hi)
    return (np.ascontiguousarray(np.asarray(cam_idx)[keep]), np.ascontiguousarray(pt[keep] - lo).astype(np.int32),
            np.ascontiguousarray(np.asarray(obs)[keep]), (lo, hi), bounds)


class B200BundleAdjustment:
    """Drop-in for JaxBundleAdjustment(cam_num, avg_cam_width)."""

    def __init__(self, cam_num, avg_cam_width, device=None, tol=1e-6, maxiter=100,
                 damping_parameter=1e-6, pcg_rtol=None, pcg_maxiter=1000, jac_f32=None,
                 verbose=False, process_group=None, distributed=None, residual_model="scalar", robust_loss=None,
                 loss_scale=1.0):
        """residual_model: 'scalar' (the reference) or 'vector2' (the COLMAP comparator's 2-vector cost).  robust_loss (vector2
        only): None = TrivialLoss, 'cauchy' / 'soft_l1' = ceres::CauchyLoss / SoftLOneLoss(loss_scale [pixels]) as
        `BundleAdjustmentOptions.CreateLossFunction` selects them (colmap/src/optim/bundle_adjustment.cc:53-68)."""
        if residual_model not in ("scalar", "vector2"):
            raise ValueError("residual_model must be 'scalar' (the reference) or 'vector2' (the COLMAP comparator's cost)")
        if robust_loss not in (None, "trivial", "cauchy", "soft_l1"):
            raise ValueError("robust_loss must be None / 'trivial', 'cauchy' or 'soft_l1'")
        if robust_loss in ("cauchy", "soft_l1") and residual_model != "vector2":
            raise ValueError("robust losses belong to residual_model='vector2' (the reference's own cost has no robustifier)")
        self.residual_model = residual_model
        self.robust_loss = None if robust_loss == "trivial" else robust_loss
        self.loss_scale = float(loss_scale)
        if jac_f32 is None:
            # stored Jacobian rows: fp32 (fp64 accumulation) for the reference's small-residual scalar model; the 2-vector
            # model is a LARGE-residual problem whose last iterations stall on fp32 storage noise (measured: 21 LM
            # iterations instead of 10 on the `small` scene, same final cost) -> fp64 rows
            jac_f32 = residual_model == "scalar"
        if pcg_rtol is None:
            pcg_rtol = _lib.default_pcg_rtol()
        self.cam_num = int(cam_num)
        self.avg_cam_width_sqr = float(avg_cam_width) ** 2
        self.device = torch.device(device if device is not None else "cuda")
        self.tol, self.maxiter, self.damping_parameter = tol, maxiter, damping_parameter
        self.pcg_rtol, self.pcg_maxiter = pcg_rtol, pcg_maxiter
        self.jac_f32, self.verbose = jac_f32, verbose
        self.fixed_pcg_iters = 0
        self.group = process_group
        # Point-sharded multi-GPU mode is OPT-IN (distributed=True or an explicit process_group): a process group that
        # happens to be initialised for unrelated work must not turn every solve into a collective.
        import torch.distributed as dist
        on = bool(distributed) or process_group is not None
        if on and not dist.is_initialized():
            raise _lib.B200BAError("distributed=True needs an initialised torch.distributed process group")
        self.world = dist.get_world_size(process_group) if on else 1
        self.rank = dist.get_rank(process_group) if self.world > 1 else 0

    # ------------------------------------------------------------------ reference API
    def prepare_params(self, poses0, intrinsics0, points0):
        """bundle_adjustment.py:93-102: opt_params = [C*6 | C*2 | P*3], cx_cy_skew (C,3)."""
        intrinsics0 = np.asarray(intrinsics0, dtype=np.float64)
        cam = pose_mat_to_vec(np.asarray(poses0, dtype=np.float64)).reshape(-1)
        opt = np.concatenate([cam, intrinsics0[..., :2].reshape(-1),
                              np.asarray(points0, dtype=np.float64).reshape(-1)])
        return opt, intrinsics0[..., 2:].copy()

    def compile(self, points_num=0, indices_num=0):
        """bundle_adjustment.py:122-129 warms the XLA cache; here: load the library and create
        the CUDA context (nothing is shape-specialised)."""
        _require_cuda()
        torch.zeros(1, device=self.device)
        torch.cuda.synchronize(self.device)

    def optimize(self, opt_params, points_2d_all, p3d_indices_all=None, cx_cy_skew=None, masks_all=None):
        """bundle_adjustment.py:104-120.  `points_2d_all` may also be a BAProblem built by
        build_problem()/build_problem_sparse() (then indices/masks are ignored)."""
        if isinstance(points_2d_all, BAProblem):
            return self._solve(opt_params, cx_cy_skew, points_2d_all)
        # like the reference, the arrays are re-read on every call (no cache keyed on object identity: a caller that
        # rewrites its buffers in place must get the new problem); build a BAProblem once for repeated solves
        n_pts = (np.asarray(opt_params).shape[0] - 8 * self.cam_num) // 3
        prob = self.build_problem(points_2d_all, p3d_indices_all, masks_all, n_pts)
        try:
            return self._solve(opt_params, cx_cy_skew, prob)
        finally:
            prob.close()

    # ------------------------------------------------------------------ problem builders
    def build_problem(self, points_2d_all, p3d_indices_all, masks_all, points_num):
        """Dense padded (C,Nmax) arrays -> sparse rows on the device (mask==1 only; padded slots
        contribute exactly zero in the reference, `benchmark_bundle_adjustment.py:81-84`)."""
        _require_cuda()
        dev = self.device
        mask = _dev(masks_all, torch.float64, dev)
        idx = _dev(p3d_indices_all, torch.int64, dev)
        p2d = _dev(points_2d_all, torch.float64, dev)
        sel = mask != 0
        cam_idx = sel.nonzero()[:, 0].to(torch.int32)
        pt_idx = idx[sel].to(torch.int32)
        obs = p2d[sel].contiguous()
        return self.build_problem_sparse(cam_idx, pt_idx, obs, points_num)

    def build_problem_sparse(self, cam_idx, pt_idx, obs, points_num, pt_range=None):
        """One row per observation (any order).  Multi-rank (distributed=True): either every rank passes the FULL
        problem and the rows are sharded by point range on the device (only world+1 integers travel to the host), or
        every rank passes its OWN rows from `shard_rows` together with `pt_range=(lo, hi)` (point indices already local)
        and uploads nothing else."""
        _require_cuda()
        dev = self.device
        cam_idx = _dev(cam_idx, torch.int32, dev)
        pt_idx = _dev(pt_idx, torch.int32, dev)
        obs_ready = None
        if (isinstance(obs, torch.Tensor) and obs.device.type == "cpu" and obs.is_pinned() and obs.dtype == torch.float64
                and (pt_range is not None or self.world == 1) and getattr(self, "residual_model", "scalar") != "vector2"):
            # pinned host observations (two thirds of the upload): copy them on a side stream while the plan is built from
            # the index arrays (sorts, run tables); the plan's stream waits for the copy before it permutes them
            if getattr(self, "_obs_stream", None) is None:
                self._obs_stream = torch.cuda.Stream(device=dev)
            main = torch.cuda.current_stream(dev)
            with torch.cuda.stream(self._obs_stream):
                obs = obs.contiguous().to(dev, non_blocking=True)
                obs_ready = torch.cuda.Event()
                obs_ready.record(self._obs_stream)
            obs.record_stream(main)
        else:
            obs = _dev(obs, torch.float64, dev)
        P = int(points_num)
        lo, hi = 0, P
        if pt_range is not None:
            lo, hi = int(pt_range[0]), int(pt_range[1])
        elif self.world > 1:
            bounds = shard_points(pt_idx, P, self.world)
            lo, hi = int(bounds[self.rank]), int(bounds[self.rank + 1])
            keep = (pt_idx >= lo) & (pt_idx < hi)
            cam_idx, pt_idx, obs = cam_idx[keep].contiguous(), (pt_idx[keep] - lo).contiguous(), obs[keep].contiguous()
        model, robust = 0, None
        if getattr(self, "residual_model", "scalar") == "vector2":
            # 2-vector model: the same pipeline over 2N rows, (x, NaN) / (NaN, y) mark the component
            model = 1
            if self.robust_loss is None:
                o2 = torch.full((2 * obs.shape[0], 2), float("nan"), dtype=torch.float64, device=dev)
                o2[0::2, 0], o2[1::2, 1] = obs[:, 0], obs[:, 1]
            else:  # robust models: adjacent (x, y) row pairs, both with the full observation
                o2 = obs.repeat_interleave(2, dim=0).contiguous()
                robust = (self.robust_loss, self.loss_scale)
            cam_idx, pt_idx, obs = cam_idx.repeat_interleave(2).contiguous(), pt_idx.repeat_interleave(2).contiguous(), o2
        prob = BAProblem(self.cam_num, hi - lo, cam_idx, pt_idx, obs, dev, self.jac_f32, residual_model=model, obs_ready=obs_ready,
                         robust=robust)
        prob.pt_lo, prob.pt_hi, prob.P_global = lo, hi, P
        if self.world > 1:
            self._init_comm(prob)
        return prob

    _comm_ready = False  # one NCCL communicator per process, shared by every plan

    def _init_comm(self, prob):
        import torch.distributed as dist
        if B200BundleAdjustment._comm_ready:
            with torch.cuda.device(self.device):
                check(prob.lib.b200ba_plan_set_comm(prob.plan, self.rank, self.world, None), "set_comm")
            return
        uid = torch.zeros(128, dtype=torch.uint8)
        if self.rank == 0:
            buf = C.create_string_buffer(128)
            check(prob.lib.b200ba_nccl_unique_id(buf), "nccl_unique_id")
            uid = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
        backend = dist.get_backend(self.group)
        t = uid.to(self.device) if backend == "nccl" else uid
        dist.broadcast(t, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0, group=self.group)
        raw = bytes(t.cpu().numpy().tobytes())
        with torch.cuda.device(self.device):
            check(prob.lib.b200ba_plan_set_comm(prob.plan, self.rank, self.world, raw), "set_comm")
        B200BundleAdjustment._comm_ready = True

    # ------------------------------------------------------------------ solve
    def solve_device(self, prob: BAProblem, cam8, k3, pts, maxiter=None, tol=None):
        """LM solve on device-resident tensors: cam8 (C,8) and pts (P_local,3) are updated IN PLACE.
        No host<->device copies of problem data. Returns (LMResult struct, trajectory buffers)."""
        lib = prob.lib
        opts = LMOpts()
        lib.b200ba_lm_default_opts(C.byref(opts))
        opts.maxiter = int(self.maxiter if maxiter is None else maxiter)
        opts.tol = float(self.tol if tol is None else tol)
        opts.damping_parameter = float(self.damping_parameter)
        opts.pcg_rtol, opts.pcg_maxiter = float(self.pcg_rtol), int(self.pcg_maxiter)
        opts.fixed_pcg_iters, opts.verbose = int(self.fixed_pcg_iters), int(bool(self.verbose))
        m = max(opts.maxiter, 1)
        loss_t = (C.c_double * (m + 1))()
        grad_t = (C.c_double * (m + 1))()
        mu_t = (C.c_double * (m + 1))()
        pcg_t = (C.c_int * m)()
        acc_t = (C.c_int * m)()
        res = LMResult()
        res.loss_trajectory_host = C.cast(loss_t, C.POINTER(C.c_double))
        res.grad_trajectory_host = C.cast(grad_t, C.POINTER(C.c_double))
        res.mu_trajectory_host = C.cast(mu_t, C.POINTER(C.c_double))
        res.pcg_iters_host = C.cast(pcg_t, C.POINTER(C.c_int))
        res.accepted_host = C.cast(acc_t, C.POINTER(C.c_int))
        with torch.cuda.device(self.device):
            check(lib.b200ba_lm_solve(prob.plan, _ptr(cam8), _ptr(k3), _ptr(pts),
                                      1.0 / self.avg_cam_width_sqr, C.byref(opts), C.byref(res), _stream()), "lm_solve")
        return res, (loss_t, grad_t, mu_t, pcg_t, acc_t)

    def compute_reprojection_errors_alt(self, opt_params, cx_cy_skew, prob: BAProblem, loss_function=LossFunction.TRIVIAL,
                                        loss_scale=1.0):
        """`Dataset.compute_reprojection_errors_alt(loss_function)` (src/dataset/dataset.py:116-123) on packed solver
        parameters: returns (per-row errors in the caller's row order, per-camera sums), both NumPy, pixels^2.
        One kernel launch for all cameras (`b200ba_reprojection_errors`); multi-rank: this rank's rows only.
        `LossFunction.CERES_CAUCHY / CERES_SOFT_L1` (+ loss_scale): the per-observation rho(|r|^2) of the COLMAP comparator's
        cost (half the sum is what Ceres reports)."""
        Cn, dev = self.cam_num, self.device
        xd = _dev(opt_params, torch.float64, dev)
        cam8 = torch.cat([xd[:6 * Cn].view(Cn, 6), xd[6 * Cn:8 * Cn].view(Cn, 2)], dim=1).contiguous()
        pts = xd[8 * Cn:].view(-1, 3)[prob.pt_lo:prob.pt_hi].contiguous()
        k3 = _dev(cx_cy_skew, torch.float64, dev)
        err, sums = BAKernels(prob, self.avg_cam_width_sqr ** 0.5).reprojection_errors(cam8, k3, pts, loss_function, loss_scale)
        return err.cpu().numpy(), sums.cpu().numpy()

    def _solve(self, opt_params, cx_cy_skew, prob: BAProblem):
        lib = prob.lib
        Cn, dev = self.cam_num, self.device
        x = np.asarray(opt_params, dtype=np.float64) if not isinstance(opt_params, torch.Tensor) else opt_params
        xd = _dev(x, torch.float64, dev)
        cam8 = torch.cat([xd[:6 * Cn].view(Cn, 6), xd[6 * Cn:8 * Cn].view(Cn, 2)], dim=1).contiguous()
        pts_all = xd[8 * Cn:].view(-1, 3)
        pts = pts_all[prob.pt_lo:prob.pt_hi].contiguous().clone()
        k3 = _dev(cx_cy_skew, torch.float64, dev)
        res, traj = self.solve_device(prob, cam8, k3, pts)
        loss_t, grad_t, mu_t, pcg_t, acc_t = traj
        if self.world > 1:
            # every rank returns the FULL parameter vector (the reference's contract): one all-reduce of a zero-filled
            # (P,3) buffer that holds this rank's point range (ranges are disjoint; works for unequal shard sizes and on
            # any backend)
            import torch.distributed as dist
            pts_full = torch.zeros(prob.P_global, 3, dtype=torch.float64, device=dev)
            pts_full[prob.pt_lo:prob.pt_hi] = pts
            if dist.get_backend(self.group) == "nccl":
                dist.all_reduce(pts_full, group=self.group)
            else:
                h = pts_full.cpu()
                dist.all_reduce(h, group=self.group)
                pts_full = h.to(dev)
        else:
            pts_full = pts
        # result read-back through a pinned buffer (a pageable D2H of Venice's 24 MB runs at a fraction of the link rate)
        flat = torch.cat([cam8[:, :6].reshape(-1), cam8[:, 6:].reshape(-1), pts_full.reshape(-1)])
        host = torch.empty(flat.shape, dtype=flat.dtype, pin_memory=True)
        host.copy_(flat, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        out = host.numpy()  # keeps the pinned tensor alive
        n = res.iter_num
        state = LMState(iter_num=n, loss=res.loss, error=res.grad_l2, damping_factor=res.damping_factor,
                        increase_factor=res.increase_factor, status=res.status,
                        loss_trajectory=np.array(loss_t[:n + 1]), grad_trajectory=np.array(grad_t[:n + 1]),
                        mu_trajectory=np.array(mu_t[:n + 1]), pcg_iters=np.array(pcg_t[:n]),
                        accepted=np.array(acc_t[:n]), n_jac_evals=res.n_jac_evals, n_res_evals=res.n_res_evals)
        return out, state


class B200PoseOptimizer:
    """Drop-in for JaxPoseOptimizer(avg_cam_width, loss_fn=JaxLossFunction.CAUCHY)."""

    def __init__(self, avg_cam_width, loss_fn=CAUCHY, device=None, tol=1e-7, maxiter=100,
                 damping_parameter=1e-6):
        self.avg_cam_width_sqr = float(avg_cam_width) ** 2
        self.loss_fn = int(loss_fn)
        self.device = torch.device(device if device is not None else "cuda")
        self.tol, self.maxiter, self.damping_parameter = tol, maxiter, damping_parameter

    def prepare_params(self, poses0, intrinsics0):
        """pose_optimization.py:73-84: (B,8) = [w, t, fx, fy], cx_cy_skew (B,3)."""
        intrinsics0 = np.asarray(intrinsics0, dtype=np.float64)
        cam = pose_mat_to_vec(np.asarray(poses0, dtype=np.float64))
        return np.concatenate([cam, intrinsics0[..., :2]], axis=-1), intrinsics0[..., 2:].copy()

    def compile(self, points_num=0, batch_size=8):
        _require_cuda()
        torch.zeros(1, device=self.device)
        torch.cuda.synchronize(self.device)

    def optimize_device(self, params, points, observations, cx_cy_skew, mask):
        """Device tensors in (params is updated IN PLACE), device tensors out. No host copies."""
        lib = _require_cuda()
        B, N = int(points.shape[0]), int(points.shape[1])
        dev = params.device
        it = torch.empty(B, dtype=torch.int32, device=dev)
        st = torch.empty(B, dtype=torch.int32, device=dev)
        loss = torch.empty(B, dtype=torch.float64, device=dev)
        err = torch.empty(B, dtype=torch.float64, device=dev)
        mu = torch.empty(B, dtype=torch.float64, device=dev)
        with torch.cuda.device(dev):
            check(lib.b200ba_pose_lm_batched(B, N, _ptr(params), _ptr(cx_cy_skew), _ptr(points), _ptr(observations),
                                             _ptr(mask), 1.0 / self.avg_cam_width_sqr, self.loss_fn, int(self.maxiter),
                                             float(self.tol), float(self.damping_parameter), _ptr(it), _ptr(loss),
                                             _ptr(err), _ptr(mu), _ptr(st), _stream()), "pose_lm_batched")
        return params, it, loss, err, mu, st

    def optimize(self, opt_params, points, observations, cx_cy_skew, mask):
        """pose_optimization.py:86-89. Host arrays in, (params (B,8), state) out; inputs untouched.
        The batch is cut into chunks: chunk k+1 travels host->device on a copy stream while chunk k is being solved
        (pinned torch CPU tensors copy at full PCIe speed; NumPy arrays take the driver's pageable path); ONE
        device->host copy returns parameters and per-pose state."""
        _require_cuda()
        dev = self.device
        host = [a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
                for a in (opt_params, points, observations, cx_cy_skew, mask)]
        host = [t if t.dtype == torch.float64 else t.to(torch.float64) for t in host]
        B = int(host[0].shape[0])
        nchunk = max(1, min(8, B // 256))
        with torch.cuda.device(dev):
            main = torch.cuda.current_stream()
            if getattr(self, "_copy_stream", None) is None:
                self._copy_stream = torch.cuda.Stream(device=dev)
            cs = self._copy_stream
            d = [torch.empty(t.shape, dtype=torch.float64, device=dev) for t in host]
            out = torch.empty(B, 13, dtype=torch.float64, device=dev)
            cs.wait_stream(main)
            bounds = [B * k // nchunk for k in range(nchunk + 1)]
            events = []
            with torch.cuda.stream(cs):
                for k in range(nchunk):
                    lo, hi = bounds[k], bounds[k + 1]
                    for dst, src in zip(d, host):
                        dst[lo:hi].copy_(src[lo:hi], non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(cs)
                    events.append(ev)
            for k in range(nchunk):
                lo, hi = bounds[k], bounds[k + 1]
                main.wait_event(events[k])
                p, it, loss, err, mu, st = self.optimize_device(d[0][lo:hi], d[1][lo:hi], d[2][lo:hi], d[3][lo:hi], d[4][lo:hi])
                o = out[lo:hi]
                o[:, :8] = p
                o[:, 8], o[:, 9], o[:, 10], o[:, 11], o[:, 12] = loss, err, mu, it.to(torch.float64), st.to(torch.float64)
            out = out.cpu().numpy()  # one D2H copy
        state = LMState(iter_num=out[:, 11].astype(np.int64), loss=out[:, 8].copy(), error=out[:, 9].copy(),
                        damping_factor=out[:, 10].copy(), status=out[:, 12].astype(np.int64))
        return out[:, :8].copy(), state


# ---------------------------------------------------------------------- building blocks (tests, bench)
class BAKernels:
    """Thin per-kernel access used by the parity tests and per-kernel benchmarks."""

    def __init__(self, prob: BAProblem, avg_cam_width):
        self.prob, self.lib = prob, prob.lib
        self.inv_w2 = 1.0 / float(avg_cam_width) ** 2
        self.dev = prob.device

    def residual_jacobian(self, cam8, k3, pts):
        N, d = self.prob.N, self.dev
        r = torch.empty(N, dtype=torch.float64, device=d)
        Jc = torch.empty(N, 8, dtype=torch.float64, device=d)
        Jp = torch.empty(N, 3, dtype=torch.float64, device=d)
        cost = C.c_double()
        with torch.cuda.device(d):
            check(self.lib.b200ba_residual_jacobian(self.prob.plan, _ptr(cam8), _ptr(k3), _ptr(pts), self.inv_w2,
                                                    _ptr(r), _ptr(Jc), _ptr(Jp), C.byref(cost), _stream()), "residual_jacobian")
        return r, Jc, Jp, cost.value

    def linearize(self, cam8, k3, pts, want_blocks=True, want_scalars=True):
        Cn, P, d = self.prob.C, self.prob.P, self.dev
        U = torch.empty(Cn, 8, 8, dtype=torch.float64, device=d) if want_blocks else None
        gc = torch.empty(Cn, 8, dtype=torch.float64, device=d) if want_blocks else None
        V = torch.empty(P, 6, dtype=torch.float64, device=d) if want_blocks else None
        gp = torch.empty(P, 3, dtype=torch.float64, device=d) if want_blocks else None
        sc = (C.c_double * 3)()
        with torch.cuda.device(d):
            check(self.lib.b200ba_linearize(self.prob.plan, _ptr(cam8), _ptr(k3), _ptr(pts), self.inv_w2, _ptr(U),
                                            _ptr(gc), _ptr(V), _ptr(gp),
                                            C.cast(sc, C.POINTER(C.c_double)) if want_scalars else None, _stream()), "linearize")
        return U, gc, V, gp, (sc[0], sc[1], sc[2])

    def schur_prepare(self, mu, want_rhs=True):
        rhs = torch.empty(self.prob.C, 8, dtype=torch.float64, device=self.dev) if want_rhs else None
        with torch.cuda.device(self.dev):
            check(self.lib.b200ba_schur_prepare(self.prob.plan, float(mu), _ptr(rhs), _stream()), "schur_prepare")
        return rhs

    def schur_matvec(self, x, y=None):
        y = torch.empty_like(x) if y is None else y
        with torch.cuda.device(self.dev):
            check(self.lib.b200ba_schur_matvec(self.prob.plan, _ptr(x), _ptr(y), _stream()), "schur_matvec")
        return y

    def schur_solve(self, rtol=1e-12, maxiter=1000):
        vc = torch.empty(self.prob.C, 8, dtype=torch.float64, device=self.dev)
        vp = torch.empty(self.prob.P, 3, dtype=torch.float64, device=self.dev)
        it = C.c_int()
        with torch.cuda.device(self.dev):
            check(self.lib.b200ba_schur_solve(self.prob.plan, float(rtol), int(maxiter), _ptr(vc), _ptr(vp), C.byref(it), _stream()), "schur_solve")
        return vc, vp, it.value

    def reprojection_errors(self, cam8, k3, pts, loss=LossFunction.TRIVIAL, loss_scale=1.0):
        """Per-row errors (caller's row order, pixels^2) and ordered per-camera sums — the reference's
        `Dataset.compute_reprojection_errors_alt(loss_function)` (dataset.py:116-123) for all cameras in one launch."""
        err = torch.empty(self.prob.N, dtype=torch.float64, device=self.dev)
        sums = torch.zeros(self.prob.C, dtype=torch.float64, device=self.dev)
        with torch.cuda.device(self.dev):
            check(self.lib.b200ba_reprojection_errors(self.prob.plan, _ptr(cam8), _ptr(k3), _ptr(pts), int(loss),
                                                      float(loss_scale), _ptr(err), _ptr(sums), _stream()), "reprojection_errors")
        return err, sums

    def cost(self, cam8, k3, pts):
        c = C.c_double()
        with torch.cuda.device(self.dev):
            check(self.lib.b200ba_cost(self.prob.plan, _ptr(cam8), _ptr(k3), _ptr(pts), self.inv_w2, C.byref(c), _stream()), "cost")
        return c.value


def with_noise(poses, intrinsics, points, observations, seed=0, point3d_noise=3e-2, camera_rotation_noise=5e-2,
               camera_translation_noise=5e-2, camera_intrinsics_noise=10.0, point2d_noise=1.0, move_cameras=False,
               explicit=None, device=None):
    """`Dataset.with_noise` (src/dataset/dataset.py:43-63, same keyword names and defaults) on the GPU: returns noisy COPIES
    (poses (C,3,4) W2C, intrinsics (C,5), points (P,3), observations (N,2)) as NumPy arrays; inputs untouched.
    `explicit`: dict with any of pts_vec / move_vec / rot_vec / intr_add / obs_vec to replace the generated perturbations."""
    lib = _require_cuda()
    dev = torch.device(device if device is not None else "cuda")
    ex = explicit or {}
    d = [_dev(np.asarray(a, dtype=np.float64), torch.float64, dev).clone() for a in (poses, intrinsics, points, observations)]
    e = [(_dev(np.asarray(ex[k], dtype=np.float64), torch.float64, dev) if k in ex else None)
         for k in ("pts_vec", "move_vec", "rot_vec", "intr_add", "obs_vec")]
    Cn, P, N = int(d[0].shape[0]), int(d[2].shape[0]), int(d[3].shape[0])
    with torch.cuda.device(dev):
        check(lib.b200ba_with_noise(Cn, P, N, _ptr(d[0]), _ptr(d[1]), _ptr(d[2]), _ptr(d[3]), int(seed), float(point3d_noise),
                                    float(camera_rotation_noise), float(camera_translation_noise), float(camera_intrinsics_noise),
                                    float(point2d_noise), int(bool(move_cameras)), *[_ptr(t) for t in e], _stream()), "with_noise")
    return tuple(t.cpu().numpy() for t in d)


def launch_count() -> int:
    return int(_lib.load().b200ba_launch_count())
