"""torch_scatter_lite -- ORACLE / TEST INFRASTRUCTURE ONLY.

Pure-torch restatement of the slice of third-party `torch_scatter` (un-pinned wheel,
environment.yml:33-35) the reference's GNN uses (fs_mol/modules/gnn.py:7, graph_readout.py:8):
scatter_sum / scatter_add / scatter_mean / scatter_max / scatter(reduce=...) / scatter_softmax /
scatter_log_softmax along dim 0.  Published semantics restated: empty segments
give 0 for sum / mean / max; mean divides by clamp(count, 1); scatter_max returns (values, argmax);
softmax variants are max-stabilised per segment.  Double-backward works (plain autograd ops)."""
import torch


def _dim_size(index, dim_size):
    if dim_size is not None:
        return int(dim_size)
    return int(index.max().item()) + 1 if index.numel() > 0 else 0


def _expand_index(index, src):
    if index.dim() == 1 and src.dim() > 1:
        index = index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)
    return index


def scatter_sum(src, index, dim=0, out=None, dim_size=None):
    assert dim == 0
    n = _dim_size(index, dim_size)
    out = torch.zeros((n,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    return out.scatter_add(0, _expand_index(index, src), src)


scatter_add = scatter_sum


def scatter_mean(src, index, dim=0, out=None, dim_size=None):
    n = _dim_size(index, dim_size)
    s = scatter_sum(src, index, dim, None, n)
    cnt = torch.zeros(n, dtype=src.dtype, device=src.device).scatter_add(
        0, index, torch.ones_like(index, dtype=src.dtype)).clamp_(min=1)
    if torch.is_floating_point(src):
        return s / cnt.view(-1, *([1] * (src.dim() - 1)))
    return torch.div(s, cnt.view(-1, *([1] * (src.dim() - 1))), rounding_mode="floor")


def scatter_max(src, index, dim=0, out=None, dim_size=None):
    assert dim == 0
    n = _dim_size(index, dim_size)
    idx = _expand_index(index, src)
    out = torch.zeros((n,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    out = out.scatter_reduce(0, idx, src, reduce="amax", include_self=False)
    # argmax (first index attaining the max) -- returned for API parity, unused by the reference
    arg = None
    return out, arg


def scatter_min(src, index, dim=0, out=None, dim_size=None):
    n = _dim_size(index, dim_size)
    idx = _expand_index(index, src)
    out = torch.zeros((n,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    return out.scatter_reduce(0, idx, src, reduce="amin", include_self=False), None


def scatter(src, index, dim=0, out=None, dim_size=None, reduce="sum"):
    if reduce in ("sum", "add"):
        return scatter_sum(src, index, dim, out, dim_size)
    if reduce == "mean":
        return scatter_mean(src, index, dim, out, dim_size)
    if reduce == "max":
        return scatter_max(src, index, dim, out, dim_size)[0]
    if reduce == "min":
        return scatter_min(src, index, dim, out, dim_size)[0]
    raise ValueError(reduce)


def scatter_log_softmax(src, index, dim=0, eps=1e-12, dim_size=None):
    n = _dim_size(index, dim_size)
    idx = _expand_index(index, src)
    mx = torch.zeros((n,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    mx = mx.scatter_reduce(0, idx, src.detach(), reduce="amax", include_self=False)
    rec = src - mx.gather(0, idx)
    sm = scatter_sum(rec.exp(), index, 0, None, n)
    return rec - (sm + eps).log().gather(0, idx)


def scatter_softmax(src, index, dim=0, dim_size=None):
    n = _dim_size(index, dim_size)
    idx = _expand_index(index, src)
    mx = torch.zeros((n,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    mx = mx.scatter_reduce(0, idx, src.detach(), reduce="amax", include_self=False)
    rec = (src - mx.gather(0, idx)).exp()
    sm = scatter_sum(rec, index, 0, None, n)
    return rec / sm.gather(0, idx)
