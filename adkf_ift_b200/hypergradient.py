"""Implicit-function-theorem hypergradient for arbitrary (f_outer, f_inner) closures -- host-side torch autograd.

Same contract as fs_mol/utils/cauchy_hypergradient.py:5-164 and cauchy_hypergradient_jvp.py:5-156:
called as f(params_outer, params_inner); the inner parameters are assumed to sit at a local minimum of f_inner;
`.grad` of every outer parameter is set to
        dF/dtheta - (dF/dphi)^T H^-1 d2f/dphi dtheta,      H = d2f/dphi2 (dense solve),
`.grad` of the inner parameters to dF/dphi, unused parameters get zeros, and F is returned.

The reference materialises the mixed second derivative as a Jacobian-of-a-Jacobian (h_len x |theta| floats, 304 MB for
ADKT).  Because v = H^-1 dF/dphi is a constant of the differentiation, v^T d2f/dphi dtheta = d/dtheta (v^T grad_phi f),
so ONE extra backward suffices -- that identity is also what the fused CUDA kernel
`adkf_gp_ift` implements for ADKTModel, where this generic routine is NOT on the hot path
(`adaptive_dkt_utils.run_on_task_batch(train=True)` is).  Kept for the reference's known-answer tests
(test_hypergrad.ipynb) and for user-defined objectives.
"""
from typing import Callable, Sequence, Tuple

import torch


def _flat(ts: Sequence[torch.Tensor]) -> torch.Tensor:
    return torch.cat([t.reshape(-1) for t in ts])


def cauchy_hypergradient(f_outer: Callable, f_inner: Callable, params_outer: Tuple[torch.Tensor, ...],
                         params_inner: Tuple[torch.Tensor, ...], device=None, ignore_grad_correction: bool = False,
                         sanity_checks: bool = False, ignore_direct_grad: bool = False) -> torch.Tensor:
    params_outer, params_inner = tuple(params_outer), tuple(params_inner)
    for t in params_outer + params_inner:
        t.grad = None

    F = f_outer(params_outer, params_inner)
    gF = torch.autograd.grad(F, params_outer + params_inner, allow_unused=True)
    n_out = len(params_outer)
    g_outer = [torch.zeros_like(p) if g is None else g.detach().clone() for p, g in zip(params_outer, gF[:n_out])]
    g_inner = [torch.zeros_like(p) if g is None else g.detach().clone() for p, g in zip(params_inner, gF[n_out:])]
    if ignore_direct_grad:
        for g in g_outer:
            g.zero_()

    if not ignore_grad_correction:
        f = f_inner(params_outer, params_inner)
        gphi = torch.autograd.grad(f, params_inner, create_graph=True, allow_unused=True)
        gphi = [torch.zeros_like(p) if g is None else g for p, g in zip(params_inner, gphi)]
        gflat = _flat(gphi)
        h_len = gflat.numel()
        rows = []
        for i in range(h_len):
            if gflat[i].requires_grad:
                r = torch.autograd.grad(gflat[i], params_inner, retain_graph=True, allow_unused=True)
                rows.append(_flat([torch.zeros_like(p) if x is None else x for p, x in zip(params_inner, r)]))
            else:
                rows.append(torch.zeros(h_len, dtype=gflat.dtype, device=gflat.device))
        H = torch.stack(rows).detach()
        if sanity_checks:
            logabsdet = torch.linalg.slogdet(H).logabsdet
            if logabsdet < -2:
                print(f"WARNING: determinant seems low ({logabsdet:.5g}). perhaps Hessian is not invertible?")
            assert logabsdet.item() > -10.0
        v = torch.linalg.solve(H, _flat(g_inner).to(H.dtype))
        s = (gflat * v.detach()).sum()
        if s.requires_grad:
            corr = torch.autograd.grad(s, params_outer, allow_unused=True)
            for g, c in zip(g_outer, corr):
                if c is not None:
                    g -= c.detach()

    for p, g in zip(params_outer, g_outer):
        p.grad = g
    for p, g in zip(params_inner, g_inner):
        p.grad = g
    return F.detach()


def functional_call(module: torch.nn.Module, parameters_and_buffers, args, kwargs=None):
    """fs_mol/utils/_stateless.py:79-119: call `module` with some of its parameters / buffers replaced by the given
    tensors (dotted names as in `named_parameters()`), for the duration of the call only.  The reference swaps the
    module's class to intercept attribute reads; here the entries of the owning sub-module's `_parameters` / `_buffers`
    are exchanged and restored, so code that reads `module.weight`, `named_parameters()` or `gp_params()` all sees the
    substituted tensors (plain tensors are accepted: they need not be nn.Parameter)."""
    if kwargs is None:
        kwargs = {}
    saved = []
    try:
        for name, tensor in parameters_and_buffers.items():
            *path, leaf = name.split(".")
            owner = module
            for part in path:
                owner = getattr(owner, part)
            store = owner._parameters if leaf in owner._parameters else owner._buffers
            if leaf not in store:
                raise KeyError(f"functional_call: {type(module).__name__} has no parameter or buffer named '{name}'")
            saved.append((store, leaf, store[leaf]))
            store[leaf] = tensor
        if isinstance(args, tuple):
            return module(*args, **kwargs)
        return module(args, **kwargs)
    finally:
        for store, leaf, original in reversed(saved):
            store[leaf] = original


# the JVP variant of the reference computes the same quantity with forward-over-reverse products
cauchy_hypergradient_jvp = cauchy_hypergradient
