"""A stand-in `jax` / `jaxopt` made of NumPy, used ONLY by tests/golden/make_refsrc_golden.py to execute the
reference's own source files for the residual side of the hot path (jax and jaxopt are absent from this image and
from /opt/wheelhouse).  `jax.numpy` -> numpy (float64) — or, with install("torch"), the few names the
residual code uses on float64 torch tensors so torch.autograd can differentiate it —, `jax.jit` -> identity,
`jax.vmap` -> a Python loop over the mapped axes, pytree registration -> identity.  `jaxopt.LevenbergMarquardt` only RECORDS its constructor
arguments (so the call sites' tol / maxiter / solver are captured); it cannot run — the LM state machine stays
"parity unpinned" (DESIGN.md §7).  Test infrastructure, never imported by the product."""
import sys
import types

import numpy as np


def _vmap(f, in_axes=0):
    def mapped(*args):
        axes = tuple(in_axes) if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        n = next(np.shape(a)[ax] for a, ax in zip(args, axes) if ax is not None)
        outs = [f(*[np.take(a, i, axis=ax) if ax is not None else a for a, ax in zip(args, axes)]) for i in range(n)]
        return np.stack(outs)
    return mapped


class _Config:
    def update(self, *a, **k):
        pass


class RecordedLM:
    """Records what the reference passes to jaxopt.LevenbergMarquardt (bundle_adjustment.py:83-91,
    pose_optimization.py:62-71)."""
    def __init__(self, residual_fun, **kw):
        self.residual_fun, self.kwargs = residual_fun, dict(kw)

    def run(self, *a, **k):
        raise RuntimeError("the shim cannot run jaxopt's LM")


def _torch_numpy():
    """The handful of jax.numpy names the reference's residual code uses, on float64 torch tensors, so that
    torch.autograd can differentiate the reference's own `get_residuals` (what jaxopt does with jvp/vjp)."""
    import torch
    m = types.ModuleType("jax.numpy")

    def array(x, dtype=None):
        if isinstance(x, torch.Tensor):
            return x
        if isinstance(x, (list, tuple)):
            return torch.stack([array(v) for v in x])
        return torch.as_tensor(x, dtype=torch.float64)
    m.array = array
    m.eye = lambda n, dtype=None: torch.eye(n, dtype=torch.float64)
    m.outer, m.cos, m.sin, m.log, m.einsum = torch.outer, torch.cos, torch.sin, torch.log, torch.einsum
    m.cross = lambda a, b: torch.linalg.cross(a, b.expand_as(a))
    m.concatenate = lambda xs, axis=0: torch.cat([array(x) for x in xs], dim=axis)
    m.linalg = types.SimpleNamespace(norm=torch.linalg.norm)
    m.newaxis = None
    return m


def _vmap_torch(f, in_axes=0):
    import torch

    def mapped(*args):
        axes = tuple(in_axes) if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        n = next(a.shape[ax] for a, ax in zip(args, axes) if ax is not None)
        return torch.stack([f(*[a.select(ax, i) if ax is not None else a for a, ax in zip(args, axes)]) for i in range(n)])
    return mapped


def install(backend="numpy"):
    if backend == "torch":
        jnp = _torch_numpy()
    else:
        jnp = types.ModuleType("jax.numpy")
        for name in dir(np):
            if not name.startswith("_"):
                setattr(jnp, name, getattr(np, name))
        jnp.newaxis = None
    jax = types.ModuleType("jax")
    jax.numpy = jnp
    jax.jit = lambda f=None, **k: f
    jax.vmap = _vmap_torch if backend == "torch" else _vmap
    jax.device_put = lambda x: x
    jax.config = _Config()
    tu = types.ModuleType("jax.tree_util")
    tu.register_pytree_node_class = lambda cls: cls
    jax.tree_util = tu
    jaxopt = types.ModuleType("jaxopt")
    jaxopt.LevenbergMarquardt = RecordedLM
    sys.modules.update({"jax": jax, "jax.numpy": jnp, "jax.tree_util": tu, "jaxopt": jaxopt})
