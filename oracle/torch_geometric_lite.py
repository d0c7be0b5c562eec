"""oracle/torch_geometric_lite.py -- ORACLE / TEST INFRASTRUCTURE ONLY.

Restates the handful of torch_geometric 1.7 entry points that MoleculeNet/chem_lib/models/encoder.py imports
(MoleculeNet/README.md:36-37 pins torch-geometric 1.7.0; not installable here), so that the reference's GNN_Encoder runs
UNMODIFIED in the authoring container:
  MessagePassing.propagate  flow source_to_target: x_j = x[edge_index[0]], x_i = x[edge_index[1]], messages reduced at
                            edge_index[1] with `aggr` ("add" | "mean" | "max"), then update()
  add_self_loops            appends (i, i) for every node
  global_{add,mean,max}_pool, softmax (segment softmax), glorot / zeros initialisers
GlobalAttention / Set2Set are placeholders (non-default pooling is out of scope).
"""
import inspect
import math
import sys
import types

import torch


class MessagePassing(torch.nn.Module):
    def __init__(self, aggr="add", flow="source_to_target", node_dim=0):
        super().__init__()
        self.aggr = aggr
        self.node_dim = node_dim

    def propagate(self, edge_index, size=None, **kwargs):
        x = kwargs.get("x")
        n = x.size(0) if x is not None else int(edge_index.max()) + 1
        src, tgt = edge_index[0], edge_index[1]
        avail = dict(kwargs)
        avail["edge_index"] = edge_index
        for k, v in list(kwargs.items()):
            if torch.is_tensor(v) and v.dim() >= 1 and v.size(0) == n:
                avail[k + "_j"] = v.index_select(0, src)
                avail[k + "_i"] = v.index_select(0, tgt)
        args = [avail[p] for p in inspect.signature(self.message).parameters]
        msg = self.message(*args)
        out = torch.zeros((n,) + tuple(msg.shape[1:]), dtype=msg.dtype, device=msg.device)
        aggr = getattr(self, "aggr", "add")
        if aggr in ("add", "mean"):
            out.index_add_(0, tgt, msg)
            if aggr == "mean":
                cnt = torch.zeros(n, dtype=msg.dtype, device=msg.device).index_add_(0, tgt, torch.ones_like(tgt, dtype=msg.dtype))
                out = out / cnt.clamp_min(1).view(-1, *([1] * (msg.dim() - 1)))
        else:
            raise NotImplementedError(aggr)
        return self.update(out)

    def message(self, x_j):
        return x_j

    def update(self, aggr_out):
        return aggr_out


def add_self_loops(edge_index, edge_weight=None, fill_value=1, num_nodes=None):
    n = int(num_nodes) if num_nodes is not None else int(edge_index.max()) + 1
    loop = torch.arange(n, dtype=edge_index.dtype, device=edge_index.device)
    return torch.cat([edge_index, torch.stack([loop, loop])], dim=1), edge_weight


def softmax(src, index, num_nodes=None):
    n = int(num_nodes) if num_nodes is not None else int(index.max()) + 1
    mx = torch.full((n,) + tuple(src.shape[1:]), -float("inf"), dtype=src.dtype).scatter_reduce(
        0, index.view(-1, *([1] * (src.dim() - 1))).expand_as(src), src, "amax")
    e = (src - mx.index_select(0, index)).exp()
    den = torch.zeros((n,) + tuple(src.shape[1:]), dtype=src.dtype).index_add_(0, index, e)
    return e / den.index_select(0, index)


def _pool(x, batch, mode):
    G = int(batch.max()) + 1
    out = torch.zeros(G, x.size(1), dtype=x.dtype, device=x.device).index_add_(0, batch, x)
    if mode == "mean":
        cnt = torch.zeros(G, dtype=x.dtype, device=x.device).index_add_(0, batch, torch.ones_like(batch, dtype=x.dtype))
        out = out / cnt.clamp_min(1).unsqueeze(1)
    return out


def global_add_pool(x, batch, size=None):
    return _pool(x, batch, "add")


def global_mean_pool(x, batch, size=None):
    return _pool(x, batch, "mean")


def global_max_pool(x, batch, size=None):
    G = int(batch.max()) + 1
    return torch.full((G, x.size(1)), -float("inf"), dtype=x.dtype).scatter_reduce(0, batch.view(-1, 1).expand_as(x), x, "amax")


class GlobalAttention(torch.nn.Module):
    def __init__(self, gate_nn, nn=None):
        super().__init__()
        raise NotImplementedError("GlobalAttention pooling is not on the ADKF-IFT path")


class Set2Set(torch.nn.Module):
    def __init__(self, *a, **k):
        super().__init__()
        raise NotImplementedError("Set2Set pooling is not on the ADKF-IFT path")


def glorot(tensor):
    if tensor is not None:
        stdv = math.sqrt(6.0 / (tensor.size(-2) + tensor.size(-1)))
        tensor.data.uniform_(-stdv, stdv)


def zeros(tensor):
    if tensor is not None:
        tensor.data.fill_(0)


def install():
    """Register the stand-in as `torch_geometric` (+ .nn, .nn.inits, .utils, .data) unless the real package imports."""
    try:
        import torch_geometric  # noqa: F401
        return "real"
    except Exception:
        pass
    me = sys.modules[__name__]
    tg = types.ModuleType("torch_geometric")
    nn_ = types.ModuleType("torch_geometric.nn")
    inits = types.ModuleType("torch_geometric.nn.inits")
    utils = types.ModuleType("torch_geometric.utils")
    data = types.ModuleType("torch_geometric.data")
    for name in ("MessagePassing", "global_add_pool", "global_mean_pool", "global_max_pool", "GlobalAttention", "Set2Set"):
        setattr(nn_, name, getattr(me, name))
    inits.glorot, inits.zeros = glorot, zeros
    utils.add_self_loops, utils.softmax = add_self_loops, softmax
    data.DataLoader = None
    tg.nn, tg.utils, tg.data, nn_.inits = nn_, utils, data, inits
    sys.modules.update({"torch_geometric": tg, "torch_geometric.nn": nn_, "torch_geometric.nn.inits": inits,
                        "torch_geometric.utils": utils, "torch_geometric.data": data})
    return "restated (oracle/torch_geometric_lite.py)"
