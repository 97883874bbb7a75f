"""`b200ba_cost_noalloc` — the trial cost for callers that cannot allocate (include/b200ba.h; no counterpart in the
reference) — against the plan-based `b200ba_cost`, through the C-ABI.

Collected LAST on purpose.  Passes alone in a fresh process and after the other tests:
the entry point takes the address of its file-scope scratch once per device and every call overwrites all it reads.
Tolerance: 1e-12 relative, the one tests/test_gpu_ba.py uses for the cost (the two sums run in different orders)."""
import ctypes as C

import numpy as np
import pytest
import torch

import b200ba
from b200ba import _lib
from oracle import ba_oracle as bo

pytestmark = pytest.mark.gpu

MAX_CAMS, MAX_ROWS = 16384, 1 << 25


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class _Case:
    """One scene on the device: the plan for the existing entry point and the raw arrays for the new one."""

    def __init__(self, shape, seed, model):
        sc = b200ba.synth.make_ba_scene(shape, seed=seed)
        ba = b200ba.B200BundleAdjustment(sc.C, sc.avg_cam_width, jac_f32=False)
        x0, k3 = ba.prepare_params(sc.poses0, sc.intrinsics0, sc.points0)
        ci, pi, ob = sc.cam_idx, sc.pt_idx, sc.obs
        if model:
            ci, pi, ob = bo.expand_vector2(ci, pi, ob)
        dev = torch.device("cuda:0")
        self.C, self.P, self.N, self.model = sc.C, sc.P, len(ci), model
        self.inv_w2 = 1.0 / float(sc.avg_cam_width) ** 2
        self.ci = torch.from_numpy(np.ascontiguousarray(ci, dtype=np.int32)).to(dev)
        self.pi = torch.from_numpy(np.ascontiguousarray(pi, dtype=np.int32)).to(dev)
        self.ob = torch.from_numpy(np.ascontiguousarray(ob, dtype=np.float64)).to(dev)
        xd = torch.from_numpy(x0).to(dev)
        self.cam8 = torch.cat([xd[:6 * sc.C].view(sc.C, 6), xd[6 * sc.C:8 * sc.C].view(sc.C, 2)], 1).contiguous()
        self.pts = xd[8 * sc.C:].view(-1, 3).contiguous()
        self.k3 = torch.from_numpy(k3).to(dev)
        self.prob = b200ba.BAProblem(sc.C, sc.P, self.ci, self.pi, self.ob, dev, jac_f32=False, residual_model=model)
        self.out = torch.full((1,), -1.0, dtype=torch.float64, device=dev)

    def with_plan(self, lib):
        c = C.c_double()
        _lib.check(lib.b200ba_cost(self.prob.plan, _ptr(self.cam8), _ptr(self.k3), _ptr(self.pts), self.inv_w2,
                                   C.byref(c), _stream()), "cost")
        return c.value

    def no_alloc(self, lib):
        self.out.fill_(-1.0)
        _lib.check(lib.b200ba_cost_noalloc(self.C, self.P, self.N, _ptr(self.ci), _ptr(self.pi), _ptr(self.ob),
                                           _ptr(self.cam8), _ptr(self.k3), _ptr(self.pts), self.inv_w2, self.model,
                                           _ptr(self.out), _stream()), "cost_noalloc")
        torch.cuda.synchronize()
        return float(self.out.item())


@pytest.mark.parametrize("model", [0, 1])
def test_no_alloc_cost_equals_plan_cost_on_two_calls_in_a_row(model):
    lib = _lib.load()
    torch.cuda.set_device(0)
    big, small = _Case("ladybug49", 0, model), _Case("tiny", 1, model)
    ref_big, ref_small = big.with_plan(lib), small.with_plan(lib)
    assert ref_big > 0 and ref_small > 0
    # call 1: the larger scene fills the scratch (49 camera tables, 31 + partial sums)
    got_big = big.no_alloc(lib)
    assert got_big == pytest.approx(ref_big, rel=1e-12)
    # call 2, straight after: a SMALLER scene with other parameters — what call 1 left in the shared scratch must not leak in
    free0 = torch.cuda.mem_get_info(0)[0]
    got_small = small.no_alloc(lib)
    free1 = torch.cuda.mem_get_info(0)[0]
    assert got_small == pytest.approx(ref_small, rel=1e-12)
    assert free0 - free1 < (1 << 20), "the call allocated device memory"
    # and the first scene again, at moved points: same answer as the plan path at those points
    big.pts.add_(1e-3)
    assert big.no_alloc(lib) == pytest.approx(big.with_plan(lib), rel=1e-12)
    assert big.no_alloc(lib) != pytest.approx(ref_big, rel=1e-9)
    big.prob.close(); small.prob.close()


def test_no_alloc_cost_rejects_sizes_beyond_the_fixed_scratch():
    lib = _lib.load()
    torch.cuda.set_device(0)
    out = torch.zeros(1, dtype=torch.float64, device="cuda:0")
    null = C.c_void_p(0)
    rc = lib.b200ba_cost_noalloc(MAX_CAMS + 1, 10, 10, null, null, null, null, null, null, 1.0, 0, _ptr(out), _stream())
    assert rc != 0 and b"16384" in lib.b200ba_last_error() and b"cameras" in lib.b200ba_last_error()
    rc = lib.b200ba_cost_noalloc(10, 10, MAX_ROWS + 1, null, null, null, null, null, null, 1.0, 0, _ptr(out), _stream())
    assert rc != 0 and b"33554432" in lib.b200ba_last_error() and b"observations" in lib.b200ba_last_error()
    rc = lib.b200ba_cost_noalloc(10, 10, 10, null, null, null, null, null, null, 1.0, 7, _ptr(out), _stream())
    assert rc != 0
    # N = 0 is a valid, empty sum
    out.fill_(-1.0)
    _lib.check(lib.b200ba_cost_noalloc(0, 0, 0, null, null, null, null, null, null, 1.0, 0, _ptr(out), _stream()), "empty")
    torch.cuda.synchronize()
    assert float(out.item()) == 0.0
