"""Host-side operator layer over the C-ABI: device graph index, autograd Functions for the fused graph
kernels, and the batched GP entry points.  Every function here launches hand-written sm_100a kernels from
libadkf_b200.so; nothing falls back to torch ops or to the CPU."""
import ctypes
import os
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch

from . import _cabi
from ._cabi import check, dptr, stream_ptr

KERNEL_IDS = {"matern": 0, "rbf": 1, "RBF": 1, "linear": 2, "poli1": 3, "poli2": 4, "cossim": 5, "tanimoto": 6}
STAT_SQDIST, STAT_DOT, STAT_TANIMOTO = 0, 1, 2
KERNEL_STATS = {"matern": 0, "rbf": 0, "RBF": 0, "linear": 1, "poli1": 1, "poli2": 1, "cossim": 1, "tanimoto": 2}
FREEZE_NOISE, FREEZE_OUTPUTSCALE, FREEZE_THIRD = 0x100, 0x200, 0x400


class nvtx_range:
    """NVTX range around one phase of the step (embed / gram / fit / ift / backward / all-reduce / clip+adam), so that an
    nsys / ncu timeline reads like the call stack of SURVEY section 3; ADKF_NVTX=0 turns the markers off."""
    enabled = os.environ.get("ADKF_NVTX", "1") != "0"

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        if self.enabled and torch.cuda.is_available():
            torch.cuda.nvtx.range_push(self.name)
        return self

    def __exit__(self, *exc):
        if self.enabled and torch.cuda.is_available():
            torch.cuda.nvtx.range_pop()
        return False


def kernel_arg(kernel, freeze: int = 0) -> int:
    """`kernel` argument of adkf_gp_fit / adkf_gp_ift: family id | ADKF_KERNEL_FREEZE_* bits."""
    return (KERNEL_IDS[kernel] if isinstance(kernel, str) else int(kernel)) | int(freeze)
AUX = 32
AUX_F, AUX_G, AUX_H, AUX_L0, AUX_LS_LOC, AUX_ITERS, AUX_FOUTER, AUX_DF_DPHI, AUX_V, AUX_JITTER, AUX_NFEV = (
    0, 1, 4, 13, 14, 15, 16, 17, 20, 23, 24)
ST_JITTER, ST_NOT_PD, ST_NOT_CONVERGED, ST_HESS_DAMPED, ST_BADWS = 1, 2, 4, 8, 16
MP_PARTIALS = 64


# --------------------------------------------------------------------------------------------------
# graph index
# --------------------------------------------------------------------------------------------------
@dataclass
class GraphIndex:
    """CSR over target nodes of the bidirectional edge set (device, int32) + per-graph node ranges."""
    num_nodes: int
    num_edges: int  # directed (= 2 x bonds)
    rowptr: torch.Tensor
    col_src: torch.Tensor
    col_type: torch.Tensor
    col_tgt: torch.Tensor
    col_eid: torch.Tensor
    rev_pos: torch.Tensor
    degree: torch.Tensor
    graph_ptr: Optional[torch.Tensor] = None  # [G+1] int32
    num_graphs: int = 0
    type_sorted: Optional["TypeSortedEdges"] = None
    max_degree: Optional[int] = None          # largest in-degree (host value: from the batcher, else one device read at build time)
    tile_node0: Optional[torch.Tensor] = None  # fused message passing: first node owned by each 112-row tile of the CSR

    @property
    def fused_ok(self) -> bool:
        """The fused message-passing kernel needs every node's edges inside one 128-row tile: in-degree <= 16."""
        return (self.tile_node0 is not None and self.max_degree is not None
                and self.max_degree <= int(_cabi.load().adkf_pna_fused_max_degree()))


def _histogram(idx: torch.Tensor, bins: int) -> torch.Tensor:
    """Counts of idx in [0, bins) without the device->host synchronisation torch.bincount does to size its output (the
    step must stay asynchronous: a sync here drains the GPU once per step)."""
    idx = idx.to(torch.int64)
    return torch.zeros(bins, dtype=torch.int64, device=idx.device).scatter_add_(0, idx, torch.ones_like(idx))


def build_graph_index(adjacency_lists: Sequence[torch.Tensor], num_nodes: int,
                      node_to_graph: Optional[torch.Tensor] = None, num_graphs: int = 0,
                      max_degree: Optional[int] = None) -> GraphIndex:
    """adjacency_lists: 3 x [E_t, 2] int64 CUDA tensors exactly as the reference batcher emits them.  max_degree: the largest
    in-degree when the host already knows it (the collation computes it); None reads it back from the device once."""
    lib = _cabi.load()
    assert len(adjacency_lists) == 3, "FS-Mol has 3 edge types"
    dev = adjacency_lists[0].device
    adjs = [a.contiguous() for a in adjacency_lists]
    es = [int(a.shape[0]) for a in adjs]
    etot = 2 * sum(es)
    i32 = dict(dtype=torch.int32, device=dev)
    rowptr = torch.empty(num_nodes + 1, **i32)
    degree = torch.empty(max(num_nodes, 1), **i32)
    n = max(etot, 1)
    col_src, col_tgt, col_eid, rev_pos, scratch = (torch.empty(n, **i32) for _ in range(5))
    col_type = torch.empty(n, dtype=torch.uint8, device=dev)
    rc = lib.adkf_graph_csr_build(dptr(adjs[0], torch.int64), es[0], dptr(adjs[1], torch.int64), es[1],
                                  dptr(adjs[2], torch.int64), es[2], num_nodes, dptr(rowptr), dptr(col_src),
                                  dptr(col_type), dptr(col_tgt), dptr(col_eid), dptr(rev_pos), dptr(degree),
                                  dptr(scratch), stream_ptr())
    check(rc, "adkf_graph_csr_build")
    graph_ptr = None
    if node_to_graph is not None:
        # nodes of a graph are contiguous (fsmol_batcher.py:200-202): counts -> exclusive scan
        counts = _histogram(node_to_graph, num_graphs)
        graph_ptr = torch.zeros(num_graphs + 1, **i32)
        graph_ptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
    gi = GraphIndex(num_nodes, etot, rowptr, col_src[:etot], col_type[:etot], col_tgt[:etot], col_eid[:etot],
                    rev_pos[:etot], degree[:num_nodes], graph_ptr, num_graphs)
    if etot > 0:
        gi.type_sorted = _type_sorted(gi, [2 * e for e in es])
    if num_nodes > 0:
        # tiles of the fused message-passing kernel: tile j = CSR rows [112 j, 112 j + 128) owns the nodes whose first row is in
        # [112 j, 112 j + 112)
        ntiles = int(lib.adkf_pna_fused_num_tiles(etot))
        stride = (128 - int(lib.adkf_pna_fused_max_degree()))
        bounds = torch.arange(ntiles + 1, dtype=torch.int32, device=dev) * stride
        gi.tile_node0 = torch.searchsorted(rowptr[:num_nodes], bounds, right=False).to(torch.int32)
        gi.max_degree = int(max_degree) if max_degree is not None else (int(degree[:num_nodes].max().item()) if etot > 0 else 0)
    return gi


# --------------------------------------------------------------------------------------------------
# PNA message passing
# --------------------------------------------------------------------------------------------------
class _PNAMessagePassing(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, W, b, gi: GraphIndex):
        lib = _cabi.load()
        h = h.contiguous()
        W = W.contiguous()
        b = b.contiguous()
        V = gi.num_nodes
        out = torch.empty(V, 4 * 768, dtype=torch.float32, device=h.device)
        aux_cnt = torch.empty(V, 4, 64, dtype=torch.uint8, device=h.device)
        aux_arg = torch.empty(V, 4, 64, dtype=torch.uint8, device=h.device)
        rc = lib.adkf_mp_pna_fwd(dptr(h, torch.float32), dptr(W, torch.float32), dptr(b, torch.float32),
                                 dptr(gi.rowptr), dptr(gi.col_src), dptr(gi.col_type), V, dptr(out),
                                 dptr(aux_cnt), dptr(aux_arg), stream_ptr())
        check(rc, "adkf_mp_pna_fwd")
        ctx.gi = gi
        ctx.save_for_backward(h, W, b, out, aux_cnt, aux_arg)
        return out

    @staticmethod
    def backward(ctx, d_out):
        lib = _cabi.load()
        h, W, b, out, aux_cnt, aux_arg = ctx.saved_tensors
        gi = ctx.gi
        V, E = gi.num_nodes, gi.num_edges
        dev = h.device
        d_out = d_out.contiguous()
        f32 = dict(dtype=torch.float32, device=dev)
        d_pre = torch.empty(max(E, 1), 4, 192, **f32)
        d_xs = torch.empty(max(E, 1), 128, **f32)
        d_xt = torch.empty(V, 128, **f32)
        d_h = torch.empty(V, 128, **f32)
        d_W = torch.empty_like(W)
        d_b = torch.empty_like(b)
        partials = torch.empty(MP_PARTIALS, 12, 192 * 64 + 192, **f32)
        rc = lib.adkf_mp_pna_bwd(dptr(h), dptr(W), dptr(b), dptr(gi.rowptr), dptr(gi.col_src), dptr(gi.col_tgt),
                                 dptr(gi.col_type), dptr(gi.rev_pos), V, E, dptr(out), dptr(aux_cnt), dptr(aux_arg),
                                 dptr(d_out, torch.float32), dptr(d_pre), dptr(d_xs), dptr(d_xt), dptr(d_h),
                                 dptr(d_W), dptr(d_b), dptr(partials), MP_PARTIALS, stream_ptr())
        check(rc, "adkf_mp_pna_bwd")
        return d_h, d_W, d_b, None


@dataclass
class TypeSortedEdges:
    """Edges re-ordered by (type, target node, reference message index): the layout of the GEMM-based PNA path."""
    src: torch.Tensor        # [E] int32 source node of sorted edge r
    tgt: torch.Tensor        # [E] int32
    rev: torch.Tensor        # [E] int32 sorted position of the reversed copy of edge r
    tptr: torch.Tensor       # [3V+1] int32
    type_off: List[int]      # host: first sorted position of every type (+ E at the end)
    csr2ts: Optional[torch.Tensor] = None   # [E] int32 sorted position of CSR row p


def _type_sorted(gi: GraphIndex, edges_per_type: Sequence[int]) -> TypeSortedEdges:
    """Derived from the CSR (target-major, type-major inside a row) by one stable sort on (type, target): integer work done
    once per batch.  edges_per_type = directed edge counts per type (2 x bonds), known on the host from the tensor shapes."""
    V, E = gi.num_nodes, gi.num_edges
    key = gi.col_type.to(torch.int64) * V + gi.col_tgt.to(torch.int64)
    perm = torch.argsort(key, stable=True)
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(E, device=perm.device)
    tptr = torch.zeros(3 * V + 1, dtype=torch.int32, device=perm.device)
    tptr[1:] = torch.cumsum(_histogram(key, 3 * V), 0).to(torch.int32)
    off = [0]
    for e in edges_per_type:
        off.append(off[-1] + int(e))
    assert off[-1] == E
    return TypeSortedEdges(gi.col_src[perm].contiguous(), gi.col_tgt[perm].contiguous(),
                           inv[gi.rev_pos.to(torch.int64)[perm]].to(torch.int32).contiguous(), tptr, off,
                           inv.to(torch.int32).contiguous())


MP_PNA, MP_MULTIAGGR, MP_PLAIN = 0, 1, 2          # ADKF_MP_* of include/adkf_b200.h
_MP_OUT_WIDTH = {MP_PNA: 4 * 768, MP_MULTIAGGR: 4 * 256, MP_PLAIN: 4 * 64}


class _PNAMessagePassingTC(torch.autograd.Function):
    """Tower-style relational message passing on the tensor cores: gather -> batched 3xTF32 GEMM (message MLP, all 4
    towers per launch) -> aggregation kernel (pna_gemm.cu).  mode: MP_PNA / MP_MULTIAGGR (messages 192 wide) or MP_PLAIN
    (messages 64 wide)."""

    @staticmethod
    def forward(ctx, h, W, b, gi: GraphIndex, mode: int = MP_PNA):
        lib = _cabi.load()
        ts: TypeSortedEdges = gi.type_sorted
        h, W, b = h.contiguous(), W.contiguous(), b.contiguous()
        V, E = gi.num_nodes, gi.num_edges
        msg = W.shape[2]                                  # 192 (PNA / MultiAggr) or 64 (Plain)
        if msg != (64 if mode == MP_PLAIN else 192):
            raise ValueError(f"message width {msg} does not match message-passing mode {mode}")
        f32 = dict(dtype=torch.float32, device=h.device)
        XG = torch.empty(max(E, 1), 256, **f32)
        pre = torch.empty(max(E, 1), 4 * msg, **f32)
        check(lib.adkf_pna_gather_inputs(dptr(h, torch.float32), dptr(ts.src), dptr(ts.tgt), E, dptr(XG), stream_ptr()),
              "adkf_pna_gather_inputs")
        W2, b1 = W.view(12 * msg, 64), b.view(-1)
        for t in range(3):
            e0, e1 = ts.type_off[t], ts.type_off[t + 1]
            if e1 > e0:   # pre[e0:e1, msg k : msg k + msg] = XG[e0:e1, 64k:64k+64] @ W[k][t]^T + b[k][t]
                gemm_tf32x3_batched(XG[e0:e1], False, 0, 64, W2[t * msg:], False, 3 * msg, 0, e1 - e0, msg, 64, 4, pre[e0:e1], 4 * msg,
                                    msg, bias=b1[t * msg:], bias_stride=3 * msg)
        out = torch.empty(V, _MP_OUT_WIDTH[mode], **f32)
        if mode == MP_PLAIN:
            aux_cnt = aux_arg = torch.empty(0, dtype=torch.uint8, device=h.device)
        else:
            aux_cnt = torch.empty(V, 4, 64, dtype=torch.uint8, device=h.device)
            aux_arg = torch.empty(V, 4, 64, dtype=torch.uint8, device=h.device)
        null = ctypes.c_void_p(0)
        check(lib.adkf_mp_aggregate_fwd(dptr(pre), dptr(ts.tptr), V, int(mode), dptr(out), dptr(aux_cnt) if mode != MP_PLAIN else null,
                                        dptr(aux_arg) if mode != MP_PLAIN else null, stream_ptr()), "adkf_mp_aggregate_fwd")
        ctx.gi, ctx.mode = gi, mode
        ctx.save_for_backward(W, XG, pre, out, aux_cnt, aux_arg)
        return out

    @staticmethod
    def backward(ctx, d_out):
        lib = _cabi.load()
        W, XG, pre, out, aux_cnt, aux_arg = ctx.saved_tensors
        gi, mode = ctx.gi, ctx.mode
        ts: TypeSortedEdges = gi.type_sorted
        V, E = gi.num_nodes, gi.num_edges
        msg = W.shape[2]
        f32 = dict(dtype=torch.float32, device=W.device)
        d_out = d_out.contiguous()
        d_pre = torch.empty(max(E, 1), 4 * msg, **f32)
        null = ctypes.c_void_p(0)
        plain = mode == MP_PLAIN
        check(lib.adkf_mp_dpre(dptr(pre), dptr(ts.tptr), V, int(mode), null if plain else dptr(out), null if plain else dptr(aux_cnt),
                               null if plain else dptr(aux_arg), dptr(d_out, torch.float32), dptr(d_pre), stream_ptr()),
              "adkf_mp_dpre")
        dXG = torch.empty(max(E, 1), 256, **f32)
        d_W = torch.zeros_like(W)
        d_b = torch.zeros(4, 3, msg, **f32)
        W2 = W.view(12 * msg, 64)
        for t in range(3):
            e0, e1 = ts.type_off[t], ts.type_off[t + 1]
            if e1 <= e0:
                continue
            n = e1 - e0
            # dXG[:, 64k:64k+64] = d_pre[:, msg k : msg k + msg] @ W[k][t]
            gemm_tf32x3_batched(d_pre[e0:e1], False, 0, msg, W2[t * msg:], True, 0, 3 * msg, n, 64, msg, 4, dXG[e0:e1], 256, 64)
            # dW[k][t] = d_pre[:, msg k : msg k + msg]^T @ XG[:, 64k:64k+64]   (split-K over the edges)
            gemm_tf32x3_batched(d_pre[e0:e1], True, msg, 0, XG[e0:e1], True, 64, 0, msg, 64, n, 4,
                                d_W.view(-1)[t * msg * 64:], 64, 3 * msg * 64, allow_splitk=True)
            d_b[:, t] = act_bwd_colsum(d_pre[e0:e1], None, 0, True)[1].view(4, msg)
        d_h = torch.empty(V, 128, **f32)
        check(lib.adkf_pna_dh_gather(dptr(dXG), dptr(ts.tptr), dptr(ts.rev), V, dptr(d_h), stream_ptr()), "adkf_pna_dh_gather")
        return d_h, d_W, d_b, None, None


def _fused_args(h, W, b, gi: GraphIndex):
    return (dptr(h, torch.float32), dptr(W, torch.float32), dptr(b, torch.float32), dptr(gi.rowptr), dptr(gi.col_src),
            dptr(gi.col_tgt), dptr(gi.col_type), dptr(gi.tile_node0))


class _PNAMessagePassingFused(torch.autograd.Function):
    """MultiAggr-layout message passing (unscaled sum | mean | pseudo-std | max, [V,1024]) with ONE kernel in the forward pass
    (pna_fused.cu: gather -> tcgen05 message MLP -> segment aggregation; neither the gathered inputs nor the pre-activations
    reach HBM) and a recomputing d-pre kernel in the backward pass; the weight / input gradient GEMMs and the d_h gather are
    those of _PNAMessagePassingTC.  Forward output bit-identical to _PNAMessagePassingTC(mode=MP_MULTIAGGR)."""

    @staticmethod
    def forward(ctx, h, W, b, gi: GraphIndex):
        lib = _cabi.load()
        h, W, b = h.contiguous(), W.contiguous(), b.contiguous()
        if W.shape[2] != 192:
            raise ValueError(f"fused message passing needs 192-wide messages, got {W.shape[2]}")
        V, E = gi.num_nodes, gi.num_edges
        out = torch.empty(V, 1024, dtype=torch.float32, device=h.device)
        aux_cnt = torch.empty(V, 4, 64, dtype=torch.uint8, device=h.device)
        aux_arg = torch.empty(V, 4, 64, dtype=torch.uint8, device=h.device)
        check(lib.adkf_pna_fused_fwd(*_fused_args(h, W, b, gi), V, E, dptr(out), dptr(aux_cnt), dptr(aux_arg), stream_ptr()),
              "adkf_pna_fused_fwd")
        ctx.gi = gi
        ctx.save_for_backward(h, W, b, out, aux_cnt, aux_arg)
        return out

    @staticmethod
    def backward(ctx, d_out):
        lib = _cabi.load()
        h, W, b, out, aux_cnt, aux_arg = ctx.saved_tensors
        gi = ctx.gi
        ts: TypeSortedEdges = gi.type_sorted
        V, E = gi.num_nodes, gi.num_edges
        msg = 192
        f32 = dict(dtype=torch.float32, device=W.device)
        d_W = torch.zeros_like(W)
        d_b = torch.zeros(4, 3, msg, **f32)
        d_h = torch.zeros(V, 128, **f32) if E == 0 else torch.empty(V, 128, **f32)
        if E == 0:
            return d_h, d_W, d_b, None
        d_out = d_out.contiguous()
        d_pre = torch.empty(E, 4 * msg, **f32)
        check(lib.adkf_pna_fused_dpre(*_fused_args(h, W, b, gi), dptr(ts.csr2ts), V, E, dptr(out), dptr(aux_cnt), dptr(aux_arg),
                                      dptr(d_out, torch.float32), dptr(d_pre), stream_ptr()), "adkf_pna_fused_dpre")
        XG = torch.empty(E, 256, **f32)                       # inputs of the weight-gradient GEMM, re-gathered (L2-resident h)
        check(lib.adkf_pna_gather_inputs(dptr(h), dptr(ts.src), dptr(ts.tgt), E, dptr(XG), stream_ptr()), "adkf_pna_gather_inputs")
        dXG = torch.empty(E, 256, **f32)
        W2 = W.view(12 * msg, 64)
        for t in range(3):
            e0, e1 = ts.type_off[t], ts.type_off[t + 1]
            if e1 <= e0:
                continue
            n = e1 - e0
            gemm_tf32x3_batched(d_pre[e0:e1], False, 0, msg, W2[t * msg:], True, 0, 3 * msg, n, 64, msg, 4, dXG[e0:e1], 256, 64)
            gemm_tf32x3_batched(d_pre[e0:e1], True, msg, 0, XG[e0:e1], True, 64, 0, msg, 64, n, 4,
                                d_W.view(-1)[t * msg * 64:], 64, 3 * msg * 64, allow_splitk=True)
            d_b[:, t] = act_bwd_colsum(d_pre[e0:e1], None, 0, True)[1].view(4, msg)
        check(lib.adkf_pna_dh_gather(dptr(dXG), dptr(ts.tptr), dptr(ts.rev), V, dptr(d_h), stream_ptr()), "adkf_pna_dh_gather")
        return d_h, d_W, d_b, None


# Which path runs (measured on the C2 graph batch, tools/pna_fused_probe.py, profiles/r2z_pna_fused_probe.txt): the fused forward
# takes 0.94 ms; the chain took 0.99 ms when the fused kernel was written and takes 0.78 ms since its GEMM got the TMA-store epilogue.
# A backward pass that has to RECOMPUTE the pre-activations costs 1.93 ms against 1.29 ms for the chain that saved them.  What the
# fused kernel still buys is MEMORY: the chain materialises XG [2E,256] + pre [2E,768] (4 KB per message: 1.2 GB at C2, 19 GB for
# the 256-task meta-test pass) and moves 3.0 GB through HBM per block where the fused kernel moves 0.8 GB.  So it serves the passes
# without gradient whose transient buffers would exceed _FUSED_MIN_BYTES; everything else runs the chain.
# ADKF_PNA_FUSED=0: never fused; =2: fused whenever the graph allows it, training included (A/B runs).
_FUSED_MODE = int(os.environ.get("ADKF_PNA_FUSED", "1"))
_FUSED_MIN_BYTES = 32 << 30


def multiaggr_message_passing(h, W, b, gi: GraphIndex, fused: Optional[bool] = None):
    """Unscaled [V,1024] aggregate of the 4 towers: the three-kernel chain, or the fused kernel (in-degree <= 16) for no-gradient
    passes whose XG / pre buffers would be larger than _FUSED_MIN_BYTES (see above; `fused` overrides the choice)."""
    if fused is None:
        needs_grad = torch.is_grad_enabled() and (h.requires_grad or W.requires_grad or b.requires_grad)
        big = gi.num_edges * 4096 > _FUSED_MIN_BYTES
        fused = _FUSED_MODE == 2 or (_FUSED_MODE == 1 and not needs_grad and big)
    if fused and gi.fused_ok and gi.num_edges > 0:
        return _PNAMessagePassingFused.apply(h, W, b, gi)
    return _PNAMessagePassingTC.apply(h, W, b, gi, MP_MULTIAGGR)


def mp_message_passing(h, W, b, gi: GraphIndex, mode: int):
    """h [V,128], W [4,3,msg,64], b [4,3,msg] -> aggregated messages of all 4 towers for GNN types PNA / MultiAggr / Plain."""
    if getattr(gi, "type_sorted", None) is None:
        raise RuntimeError("mp_message_passing needs a GraphIndex built by build_graph_index (type-sorted edges)")
    if mode == MP_MULTIAGGR:
        return multiaggr_message_passing(h, W, b, gi)
    return _PNAMessagePassingTC.apply(h, W, b, gi, mode)


class _PNACombine(torch.autograd.Function):
    """out = P[:, :128] + amp * P[:, 128:256] + att * P[:, 256:] + bias (pna_combine.cu)."""

    @staticmethod
    def forward(ctx, P, bias, amp, att):
        lib = _cabi.load()
        P = P.contiguous()
        V = P.shape[0]
        out = torch.empty(V, 128, dtype=torch.float32, device=P.device)
        check(lib.adkf_pna_combine_fwd(dptr(P, torch.float32), dptr(amp), dptr(att),
                                       dptr(bias.contiguous(), torch.float32) if bias is not None else ctypes.c_void_p(0), V,
                                       dptr(out), stream_ptr()), "adkf_pna_combine_fwd")
        ctx.save_for_backward(amp, att)
        ctx.has_bias = bias is not None
        return out

    @staticmethod
    def backward(ctx, g):
        lib = _cabi.load()
        amp, att = ctx.saved_tensors
        g = g.contiguous()
        V = g.shape[0]
        dP = torch.empty(V, 384, dtype=torch.float32, device=g.device)
        want_db = ctx.has_bias and ctx.needs_input_grad[1]
        db = torch.empty(128, dtype=torch.float32, device=g.device) if want_db else None
        nws = int(lib.adkf_pna_combine_bwd_ws_floats(V)) if want_db else 0
        ws = torch.empty(nws, dtype=torch.float32, device=g.device) if want_db else None
        null = ctypes.c_void_p(0)
        check(lib.adkf_pna_combine_bwd(dptr(g, torch.float32), dptr(amp), dptr(att), V, dptr(dP), dptr(db) if want_db else null,
                                       dptr(ws) if want_db else null, nws, stream_ptr()), "adkf_pna_combine_bwd")
        return dP, db, None, None


def pna_degree_scalers(gi: GraphIndex):
    """(amp [V], att [V]) of gnn.py:242-263, computed once per graph batch and cached on the index."""
    cached = getattr(gi, "_pna_scalers", None)
    if cached is None:
        lib = _cabi.load()
        ts: TypeSortedEdges = gi.type_sorted
        V = gi.num_nodes
        sc = torch.empty(2, max(V, 1), dtype=torch.float32, device=ts.tptr.device)
        check(lib.adkf_pna_degree_scalers(dptr(ts.tptr), V, dptr(sc[0]), dptr(sc[1]), stream_ptr()), "adkf_pna_degree_scalers")
        cached = (sc[0], sc[1])
        gi._pna_scalers = cached
    return cached


def pna_message_passing_projected(h, W, b, W_out, b_out, gi: GraphIndex):
    """msg_out_projection(PNA aggregate) [V,128] without materialising the [V,3072] scaled aggregate: unscaled aggregation
    ([V,1024], the MultiAggr kernels) -> ONE GEMM with the three scaler blocks of W_out stacked ([384,1024]) -> combine."""
    if getattr(gi, "type_sorted", None) is None:
        raise RuntimeError("pna_message_passing_projected needs a GraphIndex built by build_graph_index (type-sorted edges)")
    agg = multiaggr_message_passing(h, W, b, gi)
    # W_out [128, 4 towers x (1 | amp | att) x 256] -> [3 x 128, 4 x 256]: row block s holds the columns of scaler s
    Wcat = W_out.view(128, 4, 3, 256).permute(2, 0, 1, 3).reshape(384, 1024)
    P = linear_tc(agg, Wcat, None)
    amp, att = pna_degree_scalers(gi)
    return _PNACombine.apply(P, b_out, amp, att)


class _MHAGather(torch.autograd.Function):
    """XG [E,256] = [x_src | x_tgt] over the type-sorted edges (mha_mp.cu)."""

    @staticmethod
    def forward(ctx, x, gi: GraphIndex):
        lib = _cabi.load()
        ts: TypeSortedEdges = gi.type_sorted
        E = gi.num_edges
        XG = torch.empty(max(E, 1), 256, dtype=torch.float32, device=x.device)
        check(lib.adkf_mha_gather_inputs(dptr(x.contiguous(), torch.float32), dptr(ts.src), dptr(ts.tgt), E, dptr(XG), stream_ptr()),
              "adkf_mha_gather_inputs")
        ctx.gi = gi
        return XG

    @staticmethod
    def backward(ctx, dXG):
        lib = _cabi.load()
        gi = ctx.gi
        ts: TypeSortedEdges = gi.type_sorted
        d_x = torch.empty(gi.num_nodes, 128, dtype=torch.float32, device=dXG.device)
        check(lib.adkf_mha_dx_gather(dptr(dXG.contiguous(), torch.float32), dptr(ts.tptr), dptr(ts.rev), gi.num_nodes, dptr(d_x),
                                     stream_ptr()), "adkf_mha_dx_gather")
        return d_x, None


class _MHAAttention(torch.autograd.Function):
    """out [V,256] from message pre-activations pre [E,256] and the node projections QK [V,1536] (mha_mp.cu)."""

    @staticmethod
    def forward(ctx, pre, QK, gi: GraphIndex):
        lib = _cabi.load()
        ts: TypeSortedEdges = gi.type_sorted
        V, E = gi.num_nodes, gi.num_edges
        pre, QK = pre.contiguous(), QK.contiguous()
        out = torch.empty(V, 256, dtype=torch.float32, device=QK.device)
        probs = torch.empty(max(E, 1), 4, dtype=torch.float32, device=QK.device)
        check(lib.adkf_mha_attention_fwd(dptr(pre, torch.float32), dptr(QK, torch.float32), dptr(ts.src), dptr(ts.tptr), V, dptr(out),
                                         dptr(probs), stream_ptr()), "adkf_mha_attention_fwd")
        ctx.gi = gi
        ctx.save_for_backward(pre, QK, probs)
        return out

    @staticmethod
    def backward(ctx, d_out):
        lib = _cabi.load()
        pre, QK, probs = ctx.saved_tensors
        gi = ctx.gi
        ts: TypeSortedEdges = gi.type_sorted
        V, E = gi.num_nodes, gi.num_edges
        d_pre = torch.empty_like(pre)
        ds = torch.empty(max(E, 1), 4, dtype=torch.float32, device=QK.device)
        dQK = torch.empty_like(QK)
        check(lib.adkf_mha_attention_bwd(dptr(pre), dptr(QK), dptr(ts.src), dptr(ts.rev), dptr(ts.tptr), V, dptr(probs),
                                         dptr(d_out.contiguous(), torch.float32), dptr(d_pre), dptr(ds), dptr(dQK), stream_ptr()),
              "adkf_mha_attention_bwd")
        return d_pre, dQK, None


def mha_message_passing(x, msg_weights, msg_biases, Wqk, gi: GraphIndex):
    """RelationalMultiHeadAttentionMP.forward (gnn.py:316-387): x [V,128]; msg_weights[t] [256,256], msg_biases[t] [256];
    Wqk [1536,128] = the three target-query then the three neighbour-key projection matrices stacked -> [V,256]."""
    if getattr(gi, "type_sorted", None) is None:
        raise RuntimeError("mha_message_passing needs a GraphIndex built by build_graph_index (type-sorted edges)")
    ts: TypeSortedEdges = gi.type_sorted
    XG = _MHAGather.apply(x, gi)
    parts = []
    for t in range(3):
        e0, e1 = ts.type_off[t], ts.type_off[t + 1]
        if e1 > e0:
            parts.append(linear_tc(XG[e0:e1], msg_weights[t], msg_biases[t]))
    pre = torch.cat(parts, 0) if parts else XG[:0]
    QK = linear_tc(x, Wqk)
    return _MHAAttention.apply(pre, QK, gi)


def pna_message_passing(h, W, b, gi: GraphIndex, impl: str = "tcgen05"):
    """h [V,128], W [4,3,192,64], b [4,3,192] -> aggregated messages [V, 3072] (all 4 towers).
    impl="tcgen05": gather + batched tensor-core GEMM + segment kernels (pna_gemm.cu; needs gi.type_sorted);
    impl="simt": the fused fp32 CUDA-core kernels of graph.cu."""
    if impl == "tcgen05" and getattr(gi, "type_sorted", None) is not None:
        return _PNAMessagePassingTC.apply(h, W, b, gi, MP_PNA)
    return _PNAMessagePassing.apply(h, W, b, gi)


# --------------------------------------------------------------------------------------------------
# readout
# --------------------------------------------------------------------------------------------------
class _ReadoutWeighted(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scores, values, graph_ptr, num_graphs, heads, head_dim, mode):
        lib = _cabi.load()
        scores = scores.contiguous()
        values = values.contiguous()
        weights = torch.empty_like(scores)
        out = torch.empty(num_graphs, heads * head_dim, dtype=torch.float32, device=scores.device)
        rc = lib.adkf_readout_weighted_fwd(dptr(scores, torch.float32), dptr(values, torch.float32),
                                           dptr(graph_ptr, torch.int32), num_graphs, heads, head_dim, mode,
                                           dptr(weights), dptr(out), stream_ptr())
        check(rc, "adkf_readout_weighted_fwd")
        ctx.save_for_backward(weights, values, graph_ptr)
        ctx.meta = (num_graphs, heads, head_dim, mode)
        return out

    @staticmethod
    def backward(ctx, d_out):
        lib = _cabi.load()
        weights, values, graph_ptr = ctx.saved_tensors
        G, heads, head_dim, mode = ctx.meta
        d_out = d_out.contiguous()
        d_scores = torch.zeros_like(weights)
        d_values = torch.zeros_like(values)
        rc = lib.adkf_readout_weighted_bwd(dptr(weights), dptr(values), dptr(graph_ptr), G, heads, head_dim, mode,
                                           dptr(d_out, torch.float32), dptr(d_scores), dptr(d_values), stream_ptr())
        check(rc, "adkf_readout_weighted_bwd")
        return d_scores, d_values, None, None, None, None, None


def readout_weighted(scores, values, graph_ptr, num_graphs, heads, head_dim, mode):
    """mode 0: per-graph softmax weights ("weighted_mean"); mode 1: sigmoid weights ("weighted_sum")."""
    return _ReadoutWeighted.apply(scores, values, graph_ptr, num_graphs, heads, head_dim, mode)


class _ReadoutMax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, graph_ptr, num_graphs):
        lib = _cabi.load()
        x = x.contiguous()
        dim = x.shape[1]
        out = torch.empty(num_graphs, dim, dtype=torch.float32, device=x.device)
        argmax = torch.empty(num_graphs, dim, dtype=torch.int32, device=x.device)
        rc = lib.adkf_readout_max_fwd(dptr(x, torch.float32), dptr(graph_ptr, torch.int32), num_graphs, dim,
                                      dptr(out), dptr(argmax), stream_ptr())
        check(rc, "adkf_readout_max_fwd")
        ctx.save_for_backward(argmax, graph_ptr)
        ctx.meta = (num_graphs, dim, x.shape[0])
        return out

    @staticmethod
    def backward(ctx, d_out):
        lib = _cabi.load()
        argmax, graph_ptr = ctx.saved_tensors
        G, dim, V = ctx.meta
        d_out = d_out.contiguous()
        d_x = torch.zeros(V, dim, dtype=torch.float32, device=d_out.device)
        rc = lib.adkf_readout_max_bwd(dptr(d_out, torch.float32), dptr(argmax), dptr(graph_ptr), G, dim, dptr(d_x),
                                      stream_ptr())
        check(rc, "adkf_readout_max_bwd")
        return d_x, None, None


def readout_max(x, graph_ptr, num_graphs):
    return _ReadoutMax.apply(x, graph_ptr, num_graphs)


# --------------------------------------------------------------------------------------------------
# sparse ECFP branch of fc[0]
# --------------------------------------------------------------------------------------------------
ECFP_DENSE_BWD_MIN_GRAPHS = 512


class _EcfpLinear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, base, fp, Wt):
        lib = _cabi.load()
        out = base.contiguous().clone()
        Wt = Wt.contiguous()
        fp = fp.contiguous()
        G, fp_dim = fp.shape
        rc = lib.adkf_ecfp_linear_fwd(dptr(fp, torch.int32), G, fp_dim, dptr(Wt, torch.float32), Wt.shape[1],
                                      dptr(out, torch.float32), stream_ptr())
        check(rc, "adkf_ecfp_linear_fwd")
        ctx.save_for_backward(fp)
        ctx.wshape = tuple(Wt.shape)
        return out

    @staticmethod
    def backward(ctx, d_out):
        lib = _cabi.load()
        (fp,) = ctx.saved_tensors
        d_out = d_out.contiguous()
        d_Wt = None
        if ctx.needs_input_grad[2] and fp.shape[0] >= ECFP_DENSE_BWD_MIN_GRAPHS and fp.shape[1] % 4 == 0 \
                and d_out.shape[1] % 4 == 0 and d_out.data_ptr() % 16 == 0:
            # many graphs: d_Wt = fp^T @ d_out as ONE (T,T) tensor-core GEMM (small integer counts are exact in tf32);
            # the per-bit sparse accumulation below is 6x slower at G = 5,120
            d_Wt = gemm_tf32x3(fp.to(torch.float32), True, d_out, True)
        elif ctx.needs_input_grad[2]:
            d_Wt = torch.zeros(ctx.wshape, dtype=torch.float32, device=d_out.device)
            rc = lib.adkf_ecfp_linear_bwd(dptr(fp), fp.shape[0], fp.shape[1], dptr(d_out, torch.float32),
                                          ctx.wshape[1], dptr(d_Wt), stream_ptr())
            check(rc, "adkf_ecfp_linear_bwd")
        return d_out, None, d_Wt


def ecfp_linear(base, fp, Wt):
    """base [G,out] (dense part + bias) + fp[G,2048] (int32 counts) @ Wt[2048,out] through the sparse kernel."""
    return _EcfpLinear.apply(base, fp, Wt)


# --------------------------------------------------------------------------------------------------
# batched GP tail
# --------------------------------------------------------------------------------------------------
def host_to_device(values, dtype, device) -> torch.Tensor:
    """Small host list -> device tensor through PINNED memory with an asynchronous copy: `torch.tensor(list, device=cuda)`
    copies from pageable memory, which blocks the host until the stream reaches the copy -- one hidden synchronisation per
    call on a step that must stay asynchronous."""
    t = torch.tensor(values, dtype=dtype)
    if torch.device(device).type != "cuda":
        return t.to(device)
    return t.pin_memory().to(device, non_blocking=True)


@dataclass
class TaskLayout:
    """Row layout of a batch of T tasks inside Z / y: task t owns rows [z_off[t], z_off[t+1]) with the
    n_sup[t] support rows first; its N_t x N_t squared-distance matrix starts at d2_off[t]."""
    T: int
    n_sup_host: List[int]
    n_rows_host: List[int]
    z_off: torch.Tensor
    n_sup: torch.Tensor
    d2_off: torch.Tensor
    max_n: int
    total_rows: int
    d2_total: int

    @staticmethod
    def build(n_sup: Sequence[int], n_query: Sequence[int], device) -> "TaskLayout":
        T = len(n_sup)
        rows = [int(a) + int(b) for a, b in zip(n_sup, n_query)]
        z_off = [0]
        d2_off = [0]
        for r in rows:
            z_off.append(z_off[-1] + r)
            d2_off.append(d2_off[-1] + r * r)
        return TaskLayout(T, [int(x) for x in n_sup], rows,
                          host_to_device(z_off, torch.int32, device),
                          host_to_device([int(x) for x in n_sup], torch.int32, device),
                          host_to_device(d2_off, torch.int64, device),
                          max(rows) if rows else 0, z_off[-1], d2_off[-1])

    def ws_floats(self, which: int) -> int:
        lib = _cabi.load()
        zo = [0]
        for r in self.n_rows_host:
            zo.append(zo[-1] + r)
        a = (ctypes.c_int32 * (self.T + 1))(*zo)
        b = (ctypes.c_int32 * max(self.T, 1))(*self.n_sup_host)
        return int(lib.adkf_gp_ws_floats(self.T, ctypes.cast(a, ctypes.c_void_p), ctypes.cast(b, ctypes.c_void_p), which))


@dataclass
class SqDist:
    D2: torch.Tensor
    mu: torch.Tensor
    nrm: torch.Tensor


def gp_sqdist(Z: torch.Tensor, layout: TaskLayout, qq_block: bool, impl: str = "tcgen05", stat: int = STAT_SQDIST) -> SqDist:
    """Per-task pair statistics (stat 0: centred squared distances, 1: dot products, 2: Tanimoto similarities).
    impl="tcgen05": 3xTF32 tensor-core Gram (default, needs d % 4 == 0); impl="simt": the fp32 CUDA-core tile
    kernel (any d)."""
    lib = _cabi.load()
    Z = Z.contiguous()
    d = Z.shape[1]
    f32 = dict(dtype=torch.float32, device=Z.device)
    D2 = torch.zeros(max(layout.d2_total, 1), **f32)
    mu = torch.empty(max(layout.T, 1), d, **f32)
    nrm = torch.empty(max(layout.total_rows, 1), **f32)
    if impl == "tcgen05" and d % 4 == 0 and Z.data_ptr() % 16 == 0:
        rc = lib.adkf_gram_pairstat_tc(int(stat), dptr(Z, torch.float32), d, layout.T, Z.shape[0], dptr(layout.z_off),
                                       dptr(layout.n_sup), dptr(layout.d2_off), layout.max_n, 1 if qq_block else 0,
                                       dptr(mu), dptr(nrm), dptr(D2), stream_ptr())
        check(rc, "adkf_gram_pairstat_tc")
        return SqDist(D2, mu, nrm)
    rc = lib.adkf_gram_pairstat(int(stat), dptr(Z, torch.float32), d, layout.T, dptr(layout.z_off), dptr(layout.n_sup),
                                dptr(layout.d2_off), layout.max_n, 1 if qq_block else 0, dptr(mu), dptr(nrm), dptr(D2),
                                stream_ptr())
    check(rc, "adkf_gram_pairstat")
    return SqDist(D2, mu, nrm)


def gp_sqdist_bwd(Z, layout: TaskLayout, Wm, mu, scale: float, stat: int = STAT_SQDIST) -> torch.Tensor:
    lib = _cabi.load()
    Z = Z.contiguous()
    dZ = torch.zeros_like(Z)
    rc = lib.adkf_gram_pairstat_bwd(int(stat), dptr(Z, torch.float32), Z.shape[1], layout.T, dptr(layout.z_off),
                                    dptr(layout.d2_off), layout.max_n, dptr(Wm, torch.float32), dptr(mu), float(scale),
                                    dptr(dZ), stream_ptr())
    check(rc, "adkf_gram_pairstat_bwd")
    return dZ


def gp_fit(layout: TaskLayout, D2, y, kernel, noise_init: float, noise_prior_loc: float, ls_prior: int = 1,
           init_mode: int = 1, max_iter: int = 50, gtol: float = 2e-6, phi: Optional[torch.Tensor] = None,
           aux: Optional[torch.Tensor] = None):
    """Returns (phi [T,3] raw, aux [T,32], status [T] int32)."""
    lib = _cabi.load()
    dev = D2.device
    if phi is None:
        phi = torch.zeros(layout.T, 3, dtype=torch.float32, device=dev)
    else:
        phi = phi.contiguous().clone()
    if aux is None:
        aux = torch.zeros(layout.T, AUX, dtype=torch.float32, device=dev)
    else:
        aux = aux.contiguous().clone()
    status = torch.zeros(layout.T, dtype=torch.int32, device=dev)
    nws = layout.ws_floats(0)
    ws = torch.empty(nws, dtype=torch.float32, device=dev)
    rc = lib.adkf_gp_fit(layout.T, dptr(layout.z_off), dptr(layout.n_sup), dptr(layout.d2_off),
                         dptr(D2, torch.float32), dptr(y, torch.float32), kernel_arg(kernel), float(noise_init),
                         float(noise_prior_loc), int(ls_prior), int(init_mode), int(max_iter), float(gtol),
                         dptr(phi), dptr(aux), dptr(status), dptr(ws), nws, stream_ptr())
    check(rc, "adkf_gp_fit")
    return phi, aux, status


def gp_predict(layout: TaskLayout, D2, y, kernel, phi):
    """Returns (mean, var) aligned with the Z rows (only the query rows are written) and status."""
    lib = _cabi.load()
    dev = D2.device
    mean = torch.zeros(layout.total_rows, dtype=torch.float32, device=dev)
    var = torch.zeros(layout.total_rows, dtype=torch.float32, device=dev)
    status = torch.zeros(layout.T, dtype=torch.int32, device=dev)
    nws = layout.ws_floats(1)
    ws = torch.empty(nws, dtype=torch.float32, device=dev)
    rc = lib.adkf_gp_predict(layout.T, dptr(layout.z_off), dptr(layout.n_sup), dptr(layout.d2_off),
                             dptr(D2, torch.float32), dptr(y, torch.float32), kernel_arg(kernel) & 0xff,
                             dptr(phi.contiguous(), torch.float32), dptr(mean), dptr(var), dptr(status), dptr(ws), nws,
                             stream_ptr())
    check(rc, "adkf_gp_predict")
    return mean, var, status


def gp_ift(layout: TaskLayout, D2, y, kernel, noise_prior_loc: float, ls_prior: int, phi, aux,
           ignore_grad_correction: bool = False, mode=None):
    """Returns (Wm same layout as D2, aux updated, status)."""
    lib = _cabi.load()
    dev = D2.device
    if mode is None:
        mode = 1 if ignore_grad_correction else 0
    if os.environ.get("ADKF_GP_PROF"):      # development probe: phase timings of task 0 printed by the kernel
        mode = int(mode) | 256
    Wm = torch.zeros_like(D2)
    aux = aux.contiguous().clone()
    status = torch.zeros(layout.T, dtype=torch.int32, device=dev)
    nws = layout.ws_floats(2)
    ws = torch.empty(nws, dtype=torch.float32, device=dev)
    rc = lib.adkf_gp_ift(layout.T, dptr(layout.z_off), dptr(layout.n_sup), dptr(layout.d2_off),
                         dptr(D2, torch.float32), dptr(y, torch.float32), kernel_arg(kernel), float(noise_prior_loc),
                         int(ls_prior), int(mode), dptr(phi.contiguous(), torch.float32),
                         dptr(aux), dptr(Wm), dptr(status), dptr(ws), nws, stream_ptr())
    check(rc, "adkf_gp_ift")
    return Wm, aux, status


# --------------------------------------------------------------------------------------------------
# dense layers on the tcgen05 3xTF32 GEMM (forward, input gradient, weight gradient)
# --------------------------------------------------------------------------------------------------
ACT_IDS = {None: 0, "none": 0, "relu": 1, "leaky_relu": 2}


# pre-split weight images (adkf_gemm_presplit): built once per weight VERSION and kept on the parameter object itself, so they live
# and die with the model; ADKF_GEMM_PRESPLIT=0 switches the path off (A/B runs)
_PRESPLIT = os.environ.get("ADKF_GEMM_PRESPLIT", "1") != "0"
_PRE_MIN_ROWS, _PRE_MIN_K = 32768, 256   # (below ~32k rows a step is host-launch bound: the image kernels cost more than they save)


# Weights that are updated THROUGH AN ALIAS do not bump their own version counter: FlatParams re-seats every parameter's .data on
# a view of one flat buffer and Adam updates that buffer in place (p._version stays put).  Whoever updates weights that way calls
# weights_changed(); the epoch is part of every image's key.
_WEIGHT_EPOCH = [0]


def weights_changed() -> None:
    """Invalidate every cached weight image (call after an optimizer step that wrote parameters through an aliasing buffer)."""
    _WEIGHT_EPOCH[0] += 1


def _presplit_mode() -> bool:
    """Pre-split images follow the default correction scheme only (scaled fp16 forward / bf16 backward on CTA pairs)."""
    if not _PRESPLIT:
        return False
    prev = int(_cabi.load().adkf_gemm_configure(-1, -1))
    return (prev & 0xff) == 2 and (prev >> 8) == 3


def weight_images(w: torch.Tensor, transposed: bool):
    """(hi, co, corr) images of a weight for adkf_gemm_tf32x3_pre: of w [N,K] itself (forward, scaled-fp16 corrections) or of w^T
    (input gradient, bf16 corrections).  Cached on the tensor object, keyed by its version counter and data pointer."""
    corr = 1 if transposed else 2
    cache = getattr(w, "_adkf_images", None)
    if cache is None:
        cache = {}
        try:
            w._adkf_images = cache
        except AttributeError:
            pass
    key = (transposed, w.data_ptr(), w._version, tuple(w.shape), _WEIGHT_EPOCH[0])
    ent = cache.get(transposed)
    if ent is not None and ent[0] == key:
        return ent[1], ent[2], corr
    src = w.detach()
    src = src.t().contiguous() if transposed else (src if src.stride(1) == 1 and src.stride(0) % 2 == 0 else src.contiguous())
    rows, K = src.shape
    if ent is not None and ent[1].shape == (rows, K):
        hi, co = ent[1], ent[2]                      # same stream order as the GEMMs that read them: safe to overwrite
    else:
        hi = torch.empty(rows, K, dtype=torch.float32, device=w.device)
        co = torch.empty(rows, K, dtype=torch.float32, device=w.device)
    check(_cabi.load().adkf_gemm_presplit(ctypes.c_void_p(src.data_ptr()), rows, K, src.stride(0), corr, dptr(hi), dptr(co), stream_ptr()),
          "adkf_gemm_presplit")
    cache[transposed] = (key, hi, co)
    return hi, co, corr


def _pre_eligible(a: torch.Tensor, w: torch.Tensor, transposed: bool) -> bool:
    K = w.shape[0] if transposed else w.shape[1]
    return (a.dim() == 2 and a.shape[0] >= _PRE_MIN_ROWS and K % 32 == 0 and K >= _PRE_MIN_K and w.dim() == 2
            and a.stride(1) == 1 and a.stride(0) % 4 == 0 and a.data_ptr() % 16 == 0 and _presplit_mode())


def gemm_weight_pre(a: torch.Tensor, w: torch.Tensor, transposed: bool, bias: Optional[torch.Tensor] = None, act=None,
                    mask: Optional[torch.Tensor] = None, mask_act: int = 0) -> torch.Tensor:
    """act(a @ w.T + bias) (transposed False: w [N,K]) or a @ w (transposed True: w [K_gemm, N]), optionally times act'(mask), with the
    weight's pre-split images as B operand (adkf_gemm_tf32x3_pre)."""
    lib = _cabi.load()
    hi, co, corr = weight_images(w, transposed)
    M, K = a.shape
    N = hi.shape[0]
    out = torch.empty(M, N, dtype=torch.float32, device=a.device)
    null = ctypes.c_void_p(0)
    rc = lib.adkf_gemm_tf32x3_pre(ctypes.c_void_p(a.data_ptr()), a.stride(0), dptr(hi), dptr(co), corr, M, N, K,
                                  dptr(bias.contiguous(), torch.float32) if bias is not None else null, ACT_IDS[act],
                                  ctypes.c_void_p(mask.data_ptr()) if mask is not None else null, mask.stride(0) if mask is not None else 0,
                                  int(mask_act), dptr(out), N, stream_ptr())
    check(rc, "adkf_gemm_tf32x3_pre")
    return out


def gemm_tf32x3(a: torch.Tensor, trans_a: bool, b: torch.Tensor, trans_b: bool, bias: Optional[torch.Tensor] = None,
                act=None, out: Optional[torch.Tensor] = None, weight_b: bool = False) -> torch.Tensor:
    """D[M,N] = act(opA(a) @ opB(b).T + bias) on the tensor cores with fp32-class accuracy.
    a is [M,K] (trans_a False) or [K,M]; b is [N,K] (trans_b False) or [K,N]; both may be row-strided views
    (last dim contiguous, row stride a multiple of 4 floats, 16-byte aligned)."""
    lib = _cabi.load()
    if weight_b and not trans_a and out is None and a.is_cuda and b.is_cuda and _pre_eligible(a, b, trans_b):
        return gemm_weight_pre(a, b, trans_b, bias, act)      # B is a weight matrix: its pre-split images feed the MMA directly
    for t in (a, b):
        if not t.is_cuda or t.dtype != torch.float32 or t.dim() != 2 or t.stride(1) != 1:
            raise RuntimeError("gemm_tf32x3 needs 2-D fp32 CUDA operands with a contiguous last dimension (no CPU fallback)")
        if t.stride(0) % 4 or t.data_ptr() % 16:
            raise ValueError("gemm_tf32x3 needs row strides that are multiples of 4 floats and 16-byte aligned operands")
    K, M = (a.shape if trans_a else a.shape[::-1])
    Kb, N = (b.shape if trans_b else b.shape[::-1])
    if K != Kb:
        raise ValueError(f"contraction mismatch: {tuple(a.shape)} vs {tuple(b.shape)}")
    if out is None:
        out = torch.empty(M, N, dtype=torch.float32, device=a.device)
    elif (not out.is_cuda or out.dtype != torch.float32 or tuple(out.shape) != (M, N) or out.stride(1) != 1 or out.stride(0) % 4
          or out.data_ptr() % 16):
        raise ValueError("gemm_tf32x3: `out` must be an [M, N] fp32 CUDA tensor (or row-strided view) with 16-byte aligned rows")
    ws, nws = None, 0
    # split-K only for the backward operand forms: a forward (K,K) product must give every output row the same bits whatever the
    # number of rows in the launch (the split plan depends on the shape), so that a task's features -- and with them every
    # ReLU / arg-max decision of its gradient -- do not depend on which other tasks share the batch
    if bias is None and ACT_IDS[act] == 0 and (trans_a or trans_b):
        nws = int(lib.adkf_gemm_splitk_ws_floats(M, N, K))
        if nws:
            ws = torch.empty(nws, dtype=torch.float32, device=a.device)
    rc = lib.adkf_gemm_tf32x3(ctypes.c_void_p(a.data_ptr()), 1 if trans_a else 0, a.stride(0),
                              ctypes.c_void_p(b.data_ptr()), 1 if trans_b else 0, b.stride(0), M, N, K,
                              dptr(bias.contiguous(), torch.float32) if bias is not None else ctypes.c_void_p(0),
                              ACT_IDS[act], ctypes.c_void_p(out.data_ptr()), out.stride(0),
                              dptr(ws) if ws is not None else ctypes.c_void_p(0), nws, stream_ptr())
    check(rc, "adkf_gemm_tf32x3")
    return out


def gemm_tf32x3_masked(a: torch.Tensor, b: torch.Tensor, mask: torch.Tensor, mask_act: int, weight_b: bool = False) -> torch.Tensor:
    """(a [M,K] @ b [K,N]) * act'(mask [M,N]) -- the (K,T) input-gradient form with the activation derivative of the layer
    below applied in the epilogue (mask_act 1 relu, 2 leaky_relu(0.01))."""
    lib = _cabi.load()
    M, K = a.shape
    N = b.shape[1]
    if b.shape[0] != K or tuple(mask.shape) != (M, N):
        raise ValueError(f"shape mismatch: a {tuple(a.shape)} b {tuple(b.shape)} mask {tuple(mask.shape)}")
    if (weight_b and a.is_cuda and mask.is_cuda and mask.stride(1) == 1 and mask.stride(0) % 4 == 0 and mask.data_ptr() % 16 == 0
            and _pre_eligible(a, b, True)):
        return gemm_weight_pre(a, b, True, mask=mask, mask_act=mask_act)
    for t in (a, b, mask):
        if not t.is_cuda or t.dtype != torch.float32 or t.stride(1) != 1 or t.stride(0) % 4 or t.data_ptr() % 16:
            raise RuntimeError("gemm_tf32x3_masked needs 16-byte aligned 2-D fp32 CUDA operands (no CPU fallback)")
    out = torch.empty(M, N, dtype=torch.float32, device=a.device)
    rc = lib.adkf_gemm_tf32x3_masked(ctypes.c_void_p(a.data_ptr()), 0, a.stride(0), ctypes.c_void_p(b.data_ptr()), 1, b.stride(0),
                                     M, N, K, ctypes.c_void_p(mask.data_ptr()), mask.stride(0), int(mask_act), dptr(out), N,
                                     stream_ptr())
    check(rc, "adkf_gemm_tf32x3_masked")
    return out


def linear_tf32x3_forward(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], act=None) -> torch.Tensor:
    """act(x @ weight.T + bias) with fp32-class accuracy on the tensor cores (no autograd; see linear_tc)."""
    return gemm_tf32x3(x.contiguous(), False, weight.contiguous(), False, bias, act)


def _tc_eligible(x, weight) -> bool:
    K, N = weight.shape[1], weight.shape[0]
    # every layer whose K and N are multiples of 4 (16-byte rows for the tensor maps) goes to the tensor-core GEMM, the 12-wide
    # attention-score layer and the 32-wide input projection included: a 128-wide tile wastes MMA columns there but still beats
    # the cuBLAS SIMT kernels they used to fall back to (3 launches, 1.5 ms per step) and keeps the path on our own kernels
    return (x.is_cuda and K % 4 == 0 and N % 4 == 0 and K >= 4 and N >= 4 and x.stride(-1) == 1 and weight.stride(1) == 1
            and x.stride(0) % 4 == 0 and weight.stride(0) % 4 == 0 and x.data_ptr() % 16 == 0 and weight.data_ptr() % 16 == 0)


def act_bwd_colsum(dy: torch.Tensor, y: Optional[torch.Tensor], act: int, want_db: bool):
    """(dZ, db): dZ = dy * act'(y) (act 0: dZ is dy itself), db = dZ.sum(0) when want_db -- one pass over dy
    (adkf_act_bwd_colsum).  dy [M,N] fp32 CUDA, rows 16-byte aligned, N % 4 == 0."""
    lib = _cabi.load()
    if not dy.is_cuda or dy.dtype != torch.float32 or dy.dim() != 2 or dy.stride(1) != 1:
        raise RuntimeError("act_bwd_colsum needs a 2-D fp32 CUDA tensor with a contiguous last dimension (no CPU fallback)")
    M, N = dy.shape
    if act == 0 and not want_db:
        return dy, None
    if M == 0:
        return dy, (torch.zeros(N, dtype=torch.float32, device=dy.device) if want_db else None)
    dz = torch.empty(M, N, dtype=torch.float32, device=dy.device) if act else dy
    db = ws = None
    nws = 0
    if want_db:
        db = torch.empty(N, dtype=torch.float32, device=dy.device)
        nws = int(lib.adkf_act_bwd_colsum_ws_floats(M, N))
        ws = torch.empty(nws, dtype=torch.float32, device=dy.device)
    null = ctypes.c_void_p(0)
    rc = lib.adkf_act_bwd_colsum(ctypes.c_void_p(dy.data_ptr()), dy.stride(0), ctypes.c_void_p(y.data_ptr()) if act else null,
                                 y.stride(0) if act else 0, M, N, int(act), ctypes.c_void_p(dz.data_ptr()) if act else null,
                                 dz.stride(0) if act else 0, dptr(db) if want_db else null, dptr(ws) if want_db else null, nws,
                                 stream_ptr())
    check(rc, "adkf_act_bwd_colsum")
    return dz, db


def rezero_fwd(h: torch.Tensor, x: torch.Tensor, alpha: torch.Tensor) -> torch.Tensor:
    """h + alpha * x, alpha a 1-element device tensor (adkf_rezero_fwd)."""
    lib = _cabi.load()
    if not (h.is_cuda and x.is_cuda and alpha.is_cuda):
        raise RuntimeError("rezero_fwd needs CUDA tensors (no CPU fallback)")
    h, x = h.contiguous(), x.contiguous()
    if h.shape != x.shape or h.numel() % 4 or alpha.numel() != 1:
        raise ValueError(f"rezero_fwd: shapes {tuple(h.shape)} / {tuple(x.shape)} / alpha {tuple(alpha.shape)}")
    out = torch.empty_like(h)
    check(lib.adkf_rezero_fwd(dptr(h, torch.float32), dptr(x, torch.float32), dptr(alpha.contiguous(), torch.float32), h.numel(), dptr(out),
                              stream_ptr()), "adkf_rezero_fwd")
    return out


def rezero_bwd(g: torch.Tensor, x: torch.Tensor, alpha: torch.Tensor, want_dx: bool = True, want_dalpha: bool = True):
    """(dx = alpha * g, dalpha = <g, x> shaped like alpha) in one pass over g (adkf_rezero_bwd); None for what is not wanted."""
    lib = _cabi.load()
    g, x = g.contiguous(), x.contiguous()
    null = ctypes.c_void_p(0)
    dx = torch.empty_like(g) if want_dx else None
    dalpha = torch.empty(alpha.shape, dtype=torch.float32, device=g.device) if want_dalpha else None
    ws = torch.empty(int(lib.adkf_rezero_bwd_ws_doubles()), dtype=torch.float64, device=g.device) if want_dalpha else None
    check(lib.adkf_rezero_bwd(dptr(g, torch.float32), dptr(x, torch.float32), dptr(alpha.contiguous(), torch.float32), g.numel(),
                              dptr(dx) if want_dx else null, dptr(dalpha) if want_dalpha else null,
                              ctypes.c_void_p(ws.data_ptr()) if want_dalpha else null, stream_ptr()), "adkf_rezero_bwd")
    return dx, dalpha


class _LayerNorm128(torch.autograd.Function):
    """nn.LayerNorm(128) on [V,128] node states (gnn.py:504): warp-per-row kernels, deterministic dgamma / dbeta."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        lib = _cabi.load()
        if not x.is_cuda or x.dtype != torch.float32 or x.dim() != 2 or x.shape[1] != 128:
            raise RuntimeError("layernorm128 needs an fp32 CUDA tensor [M,128] (no CPU fallback)")
        x = x.contiguous()
        gamma, beta = gamma.contiguous(), beta.contiguous()
        M = x.shape[0]
        y = torch.empty_like(x)
        stats = torch.empty(2, max(M, 1), dtype=torch.float32, device=x.device)
        rc = lib.adkf_layernorm128_fwd(dptr(x, torch.float32), M, dptr(gamma, torch.float32), dptr(beta, torch.float32),
                                       float(eps), dptr(y), dptr(stats[0]), dptr(stats[1]), stream_ptr())
        check(rc, "adkf_layernorm128_fwd")
        ctx.save_for_backward(x, gamma, stats)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _cabi.load()
        x, gamma, stats = ctx.saved_tensors
        dy = dy.contiguous()
        M = x.shape[0]
        dx = torch.empty_like(x)
        dgb = torch.empty(2, 128, dtype=torch.float32, device=x.device)
        nws = int(lib.adkf_layernorm128_bwd_ws_floats(M))
        ws = torch.empty(nws, dtype=torch.float32, device=x.device)
        rc = lib.adkf_layernorm128_bwd(dptr(dy, torch.float32), dptr(x), dptr(stats[0]), dptr(stats[1]), dptr(gamma), M,
                                       dptr(dx), dptr(dgb[0]), dptr(dgb[1]), dptr(ws), nws, stream_ptr())
        check(rc, "adkf_layernorm128_bwd")
        return dx, dgb[0], dgb[1], None


def layernorm128(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    return _LayerNorm128.apply(x, gamma, beta, eps)


def _padded_tile_area(m: int, n: int) -> int:
    """Output area the GEMM kernels actually compute for an [m, n] product: 256 x 128 tiles on CTA pairs (m > 128), else 128 x 128."""
    bm = 256 if m > 128 else 128
    return -(-m // bm) * bm * (-(-n // 128) * 128)


def weight_grad_gemm(dy: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    """dW [N_out, K_in] = dy^T x (the (T,T) form, contraction over the rows).  The kernels tile the OUTPUT in 256 x 128 (pairs) or
    128 x 128 (M <= 128: one CTA per tile, slower per flop), so the orientation matters: [384, 1024] pads to 512 x 1024, its
    transpose [1024, 384] fits exactly (25% less tensor work), and [128, 1024] runs on pairs as [1024, 128].  The transposed
    product is returned as a strided view (no copy)."""
    m, n = dy.shape[1], x.shape[1]
    direct, swapped = _padded_tile_area(m, n), _padded_tile_area(n, m)
    if swapped < direct or (swapped == direct and m <= 128 < n):
        return gemm_tf32x3(x, True, dy, True).t()
    return gemm_tf32x3(dy, True, x, True)


class _LinearTC(torch.autograd.Function):
    """nn.Linear (+ optional ReLU / LeakyReLU(0.01)) with all three GEMMs on the tcgen05 3xTF32 kernel."""

    @staticmethod
    def forward(ctx, x, weight, bias, act):
        y = gemm_tf32x3(x, False, weight, False, bias, act, weight_b=True)
        ctx.act = ACT_IDS[act]
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x, weight, y if ctx.act else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, y = ctx.saved_tensors
        # one pass over dY: activation derivative + bias gradient (dense_bwd.cu)
        dy, db = act_bwd_colsum(dy.contiguous(), y, ctx.act, ctx.has_bias and ctx.needs_input_grad[2])
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = gemm_tf32x3(dy, False, weight, True, weight_b=True)          # dX[M,K] = dY[M,N] W[N,K]
        if ctx.needs_input_grad[1]:
            dw = weight_grad_gemm(dy, x)                       # dW[N,K] = dY^T[N,M] X[M,K]
        return dx, dw, db, None


class _MLP2TC(torch.autograd.Function):
    """Linear -> ReLU | LeakyReLU(0.01) -> Linear (BOOMLayer gnn.py:79-95; hidden-layer MLPs mlp.py:6-22) as one autograd node,
    so that the backward pass can fuse across the two layers: the input-gradient GEMM of the second layer applies the
    activation derivative in its epilogue and emits dZ1 directly (the [M,hidden] gradient is written once and never
    re-read by an elementwise pass); the bias gradients are read-only column sums."""

    @staticmethod
    def forward(ctx, x, w1, b1, act, w2, b2):
        hid = gemm_tf32x3(x, False, w1, False, b1, act, weight_b=True)
        y = gemm_tf32x3(hid, False, w2, False, b2, None, weight_b=True)
        ctx.act = ACT_IDS[act]
        ctx.has_b = (b1 is not None, b2 is not None)
        ctx.save_for_backward(x, w1, hid, w2)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w1, hid, w2 = ctx.saved_tensors
        dy = dy.contiguous()
        need = ctx.needs_input_grad
        db2 = act_bwd_colsum(dy, None, 0, True)[1] if (ctx.has_b[1] and need[5]) else None
        dw2 = weight_grad_gemm(dy, hid) if need[4] else None                 # dW2[N,H] = dY^T hid
        dz1 = gemm_tf32x3_masked(dy, w2, hid, ctx.act, weight_b=True)                        # dZ1[M,H] = (dY W2) * act'(hid)
        db1 = act_bwd_colsum(dz1, None, 0, True)[1] if (ctx.has_b[0] and need[2]) else None
        dw1 = weight_grad_gemm(dz1, x) if need[1] else None
        dx = gemm_tf32x3(dz1, False, w1, True, weight_b=True) if need[0] else None
        return dx, dw1, db1, None, dw2, db2


def mlp2_tc(x, w1, b1, act, w2, b2):
    """act(x W1^T + b1) W2^T + b2 with the fused backward of _MLP2TC when both layers fit the tensor-core kernel, else two
    linear_tc calls."""
    if (x.is_cuda and x.dim() == 2 and ACT_IDS[act] in (1, 2) and _tc_eligible(x, w1) and w2.shape[1] % 4 == 0 and w2.shape[0] % 4 == 0
            and w2.shape[0] >= 4 and w2.stride(1) == 1 and w2.stride(0) % 4 == 0 and w2.data_ptr() % 16 == 0):
        return _MLP2TC.apply(x, w1, b1, act, w2, b2)
    return linear_tc(linear_tc(x, w1, b1, act), w2, b2)


def linear_tc(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, act=None) -> torch.Tensor:
    """Drop-in for F.linear(+activation) on [rows, K] inputs.  Shapes the tensor-core kernel does not take (K or N not a
    multiple of 4: only the 42 phys-chem descriptors of the non-default `pc-descs` feature set) stay on cuBLAS."""
    if not x.is_cuda:
        raise RuntimeError("adkf_ift_b200 kernels need CUDA tensors (no CPU fallback)")
    if x.dim() == 2 and _tc_eligible(x, weight):
        return _LinearTC.apply(x, weight, bias, act)
    y = torch.nn.functional.linear(x, weight, bias)
    if ACT_IDS[act] == 1:
        y = torch.relu(y)
    elif ACT_IDS[act] == 2:
        y = torch.nn.functional.leaky_relu(y, 0.01)
    return y


# --------------------------------------------------------------------------------------------------
# large-pool GP path: blocked Cholesky joint NLL, expected improvement over a pool
# --------------------------------------------------------------------------------------------------
def gp_joint_nll_blocked(U: torch.Tensor, N: int, n_sup: int, y: torch.Tensor, kernel, phi: torch.Tensor):
    """(out4, status): out4 = [NLL(all N rows), NLL(support rows), F = -log p(y_q | y_s), log det J] for ONE task whose
    pair statistics U [N*N] came from gp_sqdist(..., qq_block=True).  Follows psd_safe_cholesky's jitter ladder
    (0, 1e-6, 1e-5, 1e-4) ON THE DEVICE (adkf_gp_joint_nll_blocked_ladder): no host synchronisation inside the op; status is
    0, ST_JITTER or ST_NOT_PD."""
    lib = _cabi.load()
    dev = U.device
    nws = int(lib.adkf_chol_blocked_ws_floats(N))
    ws = torch.empty(nws, dtype=torch.float32, device=dev)
    out = torch.zeros(4, dtype=torch.float32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    phi = phi.contiguous().view(-1)[:3]
    rc = lib.adkf_gp_joint_nll_blocked_ladder(dptr(U, torch.float32), int(N), int(n_sup), dptr(y.contiguous(), torch.float32),
                                              kernel_arg(kernel) & 0xff, dptr(phi, torch.float32), dptr(ws), nws, dptr(out),
                                              dptr(status), stream_ptr())
    check(rc, "adkf_gp_joint_nll_blocked_ladder")
    return out, status


def ei_pool(mean: torch.Tensor, var: torch.Tensor, noise: float, best_f: float,
            exclude: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Expected improvement (minimisation) of every pool molecule from the NOISY predictive mean / variance."""
    lib = _cabi.load()
    m = mean.numel()
    ei = torch.empty(m, dtype=torch.float32, device=mean.device)
    rc = lib.adkf_ei_pool(dptr(mean.contiguous(), torch.float32), dptr(var.contiguous(), torch.float32), m, float(noise),
                          float(best_f), dptr(exclude.contiguous(), torch.uint8) if exclude is not None else ctypes.c_void_p(0),
                          dptr(ei), stream_ptr())
    check(rc, "adkf_ei_pool")
    return ei


def gemm_tf32x3_batched(a: torch.Tensor, trans_a: bool, a_mn_stride: int, a_k_stride: int, b: torch.Tensor, trans_b: bool,
                        b_mn_stride: int, b_k_stride: int, M: int, N: int, K: int, batch: int, out: torch.Tensor, ldd: int,
                        d_stride: int, bias: Optional[torch.Tensor] = None, bias_stride: int = 0, act=None,
                        allow_splitk: bool = False) -> torch.Tensor:
    """`batch` products D_b[M,N] = act(opA_b @ opB_b^T + bias_b) in one launch, the operands being row / column blocks of the
    2-D arrays `a` and `b` (see adkf_gemm_tf32x3_batched in include/adkf_b200.h).  `out` is written at element offsets
    b*d_stride with leading dimension ldd."""
    lib = _cabi.load()
    for t in (a, b):
        if not t.is_cuda or t.dtype != torch.float32 or t.dim() != 2 or t.stride(1) != 1 or t.stride(0) % 4 or t.data_ptr() % 16:
            raise RuntimeError("gemm_tf32x3_batched needs aligned 2-D fp32 CUDA operands (no CPU fallback)")
    ws, nws = None, 0
    if allow_splitk and bias is None and ACT_IDS[act] == 0 and ldd == N:
        nws = int(lib.adkf_gemm_batched_splitk_ws_floats(M, N, K, batch))
        if nws:
            ws = torch.empty(nws, dtype=torch.float32, device=a.device)
    rc = lib.adkf_gemm_tf32x3_batched(
        ctypes.c_void_p(a.data_ptr()), 1 if trans_a else 0, a.stride(0), a.shape[0], a.shape[1], int(a_mn_stride), int(a_k_stride),
        ctypes.c_void_p(b.data_ptr()), 1 if trans_b else 0, b.stride(0), b.shape[0], b.shape[1], int(b_mn_stride), int(b_k_stride),
        int(M), int(N), int(K), int(batch), dptr(bias, torch.float32) if bias is not None else ctypes.c_void_p(0),
        int(bias_stride), ACT_IDS[act], ctypes.c_void_p(out.data_ptr()), int(ldd), int(d_stride),
        dptr(ws) if ws is not None else ctypes.c_void_p(0), nws, stream_ptr())
    check(rc, "adkf_gemm_tf32x3_batched")
    return out
