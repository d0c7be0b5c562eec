"""Dispatcher registration of the C-ABI entry points a caller composes most often: `torch.ops.adkf.*` (op schemas on top
of the `extern "C"` boundary).  Registered with torch.library for the CUDA key only -- a CPU tensor reaches no kernel and raises
(no fallback) -- with fake (meta) implementations, so the ops trace under torch.export / torch.compile and capture in CUDA graphs
(they allocate through torch's caching allocator only and never synchronise).

    import adkf_ift_b200.torch_ops          # registers the library once
    y = torch.ops.adkf.linear(x, weight, bias, 1)                     # act: 0 none | 1 relu | 2 leaky_relu(0.01); differentiable
    d = torch.ops.adkf.gemm_tf32x3(a, False, b, True, None, 0)        # D = opA(a) @ opB(b)^T on the tcgen05 fp32-accurate GEMM
    dz, db = torch.ops.adkf.act_bwd_colsum(dy, y, 1, True)

The reference has no native interface (it is pure PyTorch); these are the ops its nn.Linear / autograd calls map to
(fs_mol/modules/gnn.py:79-95, mlp.py:6-22), see include/adkf_b200.h for the file:line of each entry point underneath."""
from typing import Tuple

import torch

from . import ops

_ACT_NAMES = {0: None, 1: "relu", 2: "leaky_relu"}
_lib = torch.library.Library("adkf", "DEF")
_lib.define("gemm_tf32x3(Tensor a, bool trans_a, Tensor b, bool trans_b, Tensor? bias, int act) -> Tensor")
_lib.define("gemm_tf32x3_masked(Tensor a, Tensor b, Tensor mask, int mask_act) -> Tensor")
_lib.define("linear(Tensor x, Tensor weight, Tensor? bias, int act) -> Tensor")
_lib.define("act_bwd_colsum(Tensor dy, Tensor? y, int act, bool want_db) -> (Tensor, Tensor)")
_lib.define("layernorm128(Tensor x, Tensor gamma, Tensor beta, float eps) -> Tensor")
_lib.define("pna_fused_fwd(Tensor h, Tensor W, Tensor b, Tensor rowptr, Tensor col_src, Tensor col_tgt, Tensor col_type, "
            "Tensor tile_node0) -> (Tensor, Tensor, Tensor)")


def _gemm(a, trans_a, b, trans_b, bias, act):
    return ops.gemm_tf32x3(a, trans_a, b, trans_b, bias, _ACT_NAMES[int(act)])


def _gemm_fake(a, trans_a, b, trans_b, bias, act):
    M = a.shape[1] if trans_a else a.shape[0]
    N = b.shape[1] if trans_b else b.shape[0]
    return a.new_empty(M, N)


def _gemm_masked(a, b, mask, mask_act):
    return ops.gemm_tf32x3_masked(a, b, mask, int(mask_act))


def _linear(x, weight, bias, act):
    # (called below autograd: register_autograd supplies the backward; weight_b -> pre-split weight images where they pay off)
    return ops.gemm_tf32x3(x, False, weight, False, bias, _ACT_NAMES[int(act)], weight_b=True)


def _linear_setup(ctx, inputs, output):
    x, weight, bias, act = inputs
    ctx.act = int(act)
    ctx.has_bias = bias is not None
    ctx.save_for_backward(x, weight, output if ctx.act else None)


def _linear_backward(ctx, dy):
    x, weight, y = ctx.saved_tensors
    need_x, need_w, need_b = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.has_bias and ctx.needs_input_grad[2]
    dy = dy.contiguous()
    dz, db = torch.ops.adkf.act_bwd_colsum(dy, y, ctx.act, bool(need_b)) if (ctx.act or need_b) else (dy, None)
    if not ctx.act:
        dz = dy
    dx = torch.ops.adkf.gemm_tf32x3(dz, False, weight, True, None, 0) if need_x else None
    dw = torch.ops.adkf.gemm_tf32x3(dz, True, x, True, None, 0) if need_w else None
    return dx, dw, (db if need_b else None), None


def _act_bwd_colsum(dy, y, act, want_db) -> Tuple[torch.Tensor, torch.Tensor]:
    """(dZ, db): dZ = dy * act'(y); with act 0 dZ IS dy, and an op may not return an alias of its input: an empty tensor stands for
    it.  db is empty when it is not wanted."""
    dz, db = ops.act_bwd_colsum(dy, y, int(act), bool(want_db))
    return (dz if int(act) else dy.new_empty(0)), (db if db is not None else dy.new_empty(0))


def _act_bwd_colsum_fake(dy, y, act, want_db):
    return dy.new_empty(dy.shape if act else (0,)), dy.new_empty(dy.shape[1] if want_db else 0)


def _layernorm128(x, gamma, beta, eps):
    return ops.layernorm128(x, gamma, beta, float(eps))


def _pna_fused_fwd(h, W, b, rowptr, col_src, col_tgt, col_type, tile_node0):
    from ._cabi import check, dptr, load, stream_ptr
    lib = load()
    V, E = int(h.shape[0]), int(col_src.shape[0])
    out = torch.empty(V, 1024, dtype=torch.float32, device=h.device)
    cnt = torch.empty(V, 4, 64, dtype=torch.uint8, device=h.device)
    arg = torch.empty(V, 4, 64, dtype=torch.uint8, device=h.device)
    check(lib.adkf_pna_fused_fwd(dptr(h.contiguous(), torch.float32), dptr(W.contiguous(), torch.float32), dptr(b.contiguous(), torch.float32),
                                 dptr(rowptr), dptr(col_src), dptr(col_tgt), dptr(col_type), dptr(tile_node0), V, E, dptr(out), dptr(cnt),
                                 dptr(arg), stream_ptr()), "adkf_pna_fused_fwd")
    return out, cnt, arg


def _pna_fused_fwd_fake(h, W, b, rowptr, col_src, col_tgt, col_type, tile_node0):
    V = h.shape[0]
    return h.new_empty(V, 1024), h.new_empty(V, 4, 64, dtype=torch.uint8), h.new_empty(V, 4, 64, dtype=torch.uint8)


_lib.impl("gemm_tf32x3", _gemm, "CUDA")
_lib.impl("gemm_tf32x3_masked", _gemm_masked, "CUDA")
_lib.impl("linear", _linear, "CUDA")
_lib.impl("act_bwd_colsum", _act_bwd_colsum, "CUDA")
_lib.impl("layernorm128", _layernorm128, "CompositeImplicitAutograd")     # (ops.layernorm128 is its own autograd.Function)
_lib.impl("pna_fused_fwd", _pna_fused_fwd, "CUDA")
torch.library.register_fake("adkf::gemm_tf32x3", _gemm_fake, lib=_lib)
torch.library.register_fake("adkf::gemm_tf32x3_masked", lambda a, b, mask, mask_act: a.new_empty(a.shape[0], b.shape[1]), lib=_lib)
torch.library.register_fake("adkf::linear", lambda x, weight, bias, act: x.new_empty(x.shape[0], weight.shape[0]), lib=_lib)
torch.library.register_fake("adkf::act_bwd_colsum", _act_bwd_colsum_fake, lib=_lib)
torch.library.register_fake("adkf::pna_fused_fwd", _pna_fused_fwd_fake, lib=_lib)
torch.library.register_autograd("adkf::linear", _linear_backward, setup_context=_linear_setup, lib=_lib)


OPS = ("gemm_tf32x3", "gemm_tf32x3_masked", "linear", "act_bwd_colsum", "layernorm128", "pna_fused_fwd")


def registered_ops():
    return sorted(n for n in OPS if hasattr(torch.ops.adkf, n))
