"""MoleculeNet branch of ADKF-IFT (BASELINE config 4) on the B200 kernels.

Surface mirrored (state-dict names included, so `pretrained` GIN checkpoints load):
  GINConv / GNN / GNN_Encoder          MoleculeNet/chem_lib/models/encoder.py:16-58, 202-360
  ADKFModel                            MoleculeNet/chem_lib/models/adkf_model.py:15-164
  ADKF_Meta_Trainer.train_step / test  MoleculeNet/chem_lib/models/adkfift_trainer.py:148-309

The reference encodes every task's support set and query set in separate encoder calls (BatchNorm statistics per call) and
evaluates the extractor ~8 times per task inside cauchy_hypergradient.  Here all graphs of all T tasks go through ONE pass:
the graph kernels of gin.cu take the merged batch, BatchNorm statistics are taken per (task, support|query) segment
(adkf_seg_bn_*), the dense layers run on the tcgen05 GEMM, and the GP tail / IFT hypergradient are the batched kernels of
gp_core.cu.  Only the default configuration (gin, JK=last, mean pooling, batch norm) has kernels; anything else raises.
"""
from dataclasses import dataclass
from types import SimpleNamespace
from typing import List, Optional, Sequence

import torch
import torch.nn as nn

from . import gp as gpk
from . import ops
from ._cabi import check, dptr, load, stream_ptr
from .adaptive_dkt import ExactGPLayer, GaussianLikelihoodParams, MarginalLogLikelihood, _TaskGPLoss

NUM_ATOM_TYPE, NUM_CHIRALITY_TAG, NUM_BOND_TYPE, NUM_BOND_DIRECTION = 120, 3, 6, 3   # encoder.py:9-13
BN_EPS, BN_MOMENTUM = 1e-5, 0.1


# ----------------------------------------------------------------------------------------------------------
# data
# ----------------------------------------------------------------------------------------------------------
@dataclass
class MolBatch:
    """PyG `Batch` duck-type: what the reference's loaders hand to the model."""
    x: torch.Tensor            # [V,2] int64
    edge_index: torch.Tensor   # [2,E] int64
    edge_attr: torch.Tensor    # [E,2] int64
    batch: torch.Tensor        # [V] int64, graph id of every node (sorted)
    y: torch.Tensor            # [G]
    num_graphs: int

    def to(self, device):
        return MolBatch(self.x.to(device), self.edge_index.to(device), self.edge_attr.to(device), self.batch.to(device),
                        self.y.to(device), self.num_graphs)


@dataclass
class MergedMolBatch:
    """All graphs of T tasks, task-major, support graphs before query graphs inside a task."""
    x: torch.Tensor
    edge_index: torch.Tensor
    edge_attr: torch.Tensor
    graph_ptr: torch.Tensor     # [G+1] int32 node offsets
    seg_ptr: torch.Tensor       # [S+1] int32 node offsets of the BatchNorm segments (one per encoder call of the reference)
    labels: torch.Tensor        # [G] float, +-1
    layout: ops.TaskLayout
    num_graphs: int
    num_tasks: int
    index: Optional["GinIndex"] = None


def merge_mol_batches(parts: Sequence[MolBatch], device, n_sup: Sequence[int], n_query: Sequence[int]) -> MergedMolBatch:
    """parts = [task0 support, task0 query, task1 support, ...] (a task without query graphs passes an empty MolBatch or is
    listed with n_query 0 and no query part)."""
    xs, eis, eas, labels, gcounts, seg = [], [], [], [], [], [0]
    v_off = 0
    for p in parts:
        V = int(p.x.shape[0])
        xs.append(p.x)
        eis.append(p.edge_index + v_off)
        eas.append(p.edge_attr)
        labels.append((p.y.float() - 0.5) * 2.0)
        gcounts.append(torch.bincount(p.batch, minlength=p.num_graphs))
        v_off += V
        seg.append(v_off)
    counts = torch.cat(gcounts)
    graph_ptr = torch.zeros(counts.numel() + 1, dtype=torch.int32)
    graph_ptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
    layout = ops.TaskLayout.build(list(n_sup), list(n_query), device)
    return MergedMolBatch(torch.cat(xs).to(device), torch.cat(eis, 1).contiguous().to(device), torch.cat(eas).to(device),
                          graph_ptr.to(device), torch.tensor(seg, dtype=torch.int32).to(device),
                          torch.cat(labels).to(device), layout, int(counts.numel()), len(n_sup))


@dataclass
class GinIndex:
    V: int
    E: int
    rowptr_t: torch.Tensor
    col_src: torch.Tensor
    col_eid_t: torch.Tensor
    rowptr_s: torch.Tensor
    col_tgt: torch.Tensor
    tgt: torch.Tensor           # [E] int64 target of every edge (for the edge-embedding gradients)
    onehot_atom: torch.Tensor   # [V,120] / [V,4] / [E,8] / [E,4] fp32 one-hot matrices (embedding-table gradients as GEMMs)
    onehot_chir: torch.Tensor
    onehot_btype: torch.Tensor
    onehot_bdir: torch.Tensor


def build_gin_index(x, edge_index, edge_attr) -> GinIndex:
    lib = load()
    dev = x.device
    V, E = int(x.shape[0]), int(edge_index.shape[1])
    edge_index = edge_index.contiguous()
    i32 = dict(dtype=torch.int32, device=dev)
    out = []
    for by_source in (0, 1):
        rowptr = torch.empty(V + 1, **i32)
        col = torch.empty(max(E, 1), **i32)
        eid = torch.empty(max(E, 1), **i32)
        scratch = torch.empty(max(V, 1), **i32)
        check(lib.adkf_csr_from_edge_index(dptr(edge_index, torch.int64), E, V, by_source, dptr(rowptr), dptr(col), dptr(eid),
                                           dptr(scratch), stream_ptr()), "adkf_csr_from_edge_index")
        out.append((rowptr, col, eid))

    def onehot(idx, width):
        m = torch.zeros(idx.shape[0], width, dtype=torch.float32, device=dev)
        m.scatter_(1, idx.view(-1, 1), 1.0)
        return m

    return GinIndex(V, E, out[0][0], out[0][1], out[0][2], out[1][0], out[1][1], edge_index[1].contiguous(),
                    onehot(x[:, 0], NUM_ATOM_TYPE), onehot(x[:, 1], 4), onehot(edge_attr[:, 0], 8), onehot(edge_attr[:, 1], 4))


# ----------------------------------------------------------------------------------------------------------
# autograd wrappers of the gin.cu kernels
# ----------------------------------------------------------------------------------------------------------
def _table_grad(onehot: torch.Tensor, grad_rows: torch.Tensor, rows: int) -> torch.Tensor:
    """d table[r] = sum over the items that looked row r up = onehot^T @ grad_rows (tensor-core GEMM, deterministic)."""
    return ops.gemm_tf32x3(onehot, True, grad_rows.contiguous(), True)[:rows]


class _EmbedSum2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, T1, T2, idx, oh1, oh2):
        lib = load()
        D = T1.shape[1]
        out = torch.empty(idx.shape[0], D, dtype=torch.float32, device=T1.device)
        check(lib.adkf_embed_sum2(dptr(idx.contiguous(), torch.int64), idx.shape[0], D, dptr(T1.contiguous(), torch.float32),
                                  dptr(T2.contiguous(), torch.float32), dptr(out), stream_ptr()), "adkf_embed_sum2")
        ctx.save_for_backward(oh1, oh2)
        ctx.rows = (T1.shape[0], T2.shape[0])
        return out

    @staticmethod
    def backward(ctx, dout):
        oh1, oh2 = ctx.saved_tensors
        return _table_grad(oh1, dout, ctx.rows[0]), _table_grad(oh2, dout, ctx.rows[1]), None, None, None


class _GinAggregate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, E1, E2, edge_attr, gi: GinIndex):
        lib = load()
        h = h.contiguous()
        out = torch.empty_like(h)
        check(lib.adkf_gin_gather(dptr(h, torch.float32), gi.V, h.shape[1], dptr(gi.rowptr_t), dptr(gi.col_src),
                                  dptr(gi.col_eid_t), dptr(edge_attr.contiguous(), torch.int64),
                                  dptr(E1.contiguous(), torch.float32), dptr(E2.contiguous(), torch.float32), dptr(out),
                                  stream_ptr()), "adkf_gin_gather")
        ctx.gi = gi
        ctx.rows = (E1.shape[0], E2.shape[0])
        return out

    @staticmethod
    def backward(ctx, dout):
        import ctypes
        lib = load()
        gi = ctx.gi
        dout = dout.contiguous()
        dh = torch.empty_like(dout)
        null = ctypes.c_void_p(0)
        check(lib.adkf_gin_gather(dptr(dout, torch.float32), gi.V, dout.shape[1], dptr(gi.rowptr_s), dptr(gi.col_tgt), null, null,
                                  null, null, dptr(dh), stream_ptr()), "adkf_gin_gather")
        dE1 = dE2 = None
        if ctx.needs_input_grad[1] or ctx.needs_input_grad[2]:
            per_edge = dout.index_select(0, gi.tgt)           # gradient arriving at each edge's message
            self_sum = dout.sum(0)                            # the self-loop edges carry attr (4, 0)
            dE1 = _table_grad(gi.onehot_btype, per_edge, ctx.rows[0]).clone()
            dE2 = _table_grad(gi.onehot_bdir, per_edge, ctx.rows[1]).clone()
            dE1[4] += self_sum
            dE2[0] += self_sum
        return dh, dE1, dE2, None, None


class _SegBatchNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, seg_ptr, act, keep, seed, stats_out):
        lib = load()
        x = x.contiguous()
        D = x.shape[1]
        S = seg_ptr.numel() - 1
        f32 = dict(dtype=torch.float32, device=x.device)
        y = torch.empty_like(x)
        mean, rstd, uvar = (torch.empty(S, D, **f32) for _ in range(3))
        import ctypes
        null = ctypes.c_void_p(0)
        check(lib.adkf_seg_bn_fwd(dptr(x, torch.float32), D, dptr(seg_ptr, torch.int32), S, dptr(gamma.contiguous()),
                                  dptr(beta.contiguous()), null, null, BN_EPS, 0, int(act), float(keep), int(seed), dptr(y),
                                  dptr(mean), dptr(rstd), dptr(uvar), stream_ptr()), "adkf_seg_bn_fwd")
        ctx.save_for_backward(x, gamma, beta, seg_ptr, mean, rstd)
        ctx.meta = (int(act), float(keep), int(seed))
        stats_out.append((mean, uvar))
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = load()
        x, gamma, beta, seg_ptr, mean, rstd = ctx.saved_tensors
        act, keep, seed = ctx.meta
        D = x.shape[1]
        S = seg_ptr.numel() - 1
        dx = torch.empty_like(x)
        dg = torch.empty(S, D, dtype=torch.float32, device=x.device)
        db = torch.empty_like(dg)
        check(lib.adkf_seg_bn_bwd(dptr(x), dptr(dy.contiguous(), torch.float32), D, dptr(seg_ptr), S, dptr(gamma.contiguous()),
                                  dptr(beta.contiguous()), dptr(mean), dptr(rstd), act, keep, seed, dptr(dx), dptr(dg), dptr(db),
                                  stream_ptr()), "adkf_seg_bn_bwd")
        return dx, dg.sum(0), db.sum(0), None, None, None, None, None


def seg_bn_eval(x, gamma, beta, run_mean, run_var, seg_ptr, act):
    import ctypes
    lib = load()
    x = x.contiguous()
    y = torch.empty_like(x)
    null = ctypes.c_void_p(0)
    check(lib.adkf_seg_bn_fwd(dptr(x, torch.float32), x.shape[1], dptr(seg_ptr, torch.int32), seg_ptr.numel() - 1,
                              dptr(gamma.detach().contiguous()), dptr(beta.detach().contiguous()), dptr(run_mean.contiguous()),
                              dptr(run_var.contiguous()), BN_EPS, 1, int(act), 1.0, 0, dptr(y), null, null, null, stream_ptr()),
          "adkf_seg_bn_fwd")
    return y


class _SegMean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, graph_ptr, G):
        lib = load()
        x = x.contiguous()
        out = torch.empty(G, x.shape[1], dtype=torch.float32, device=x.device)
        check(lib.adkf_seg_mean_fwd(dptr(x, torch.float32), x.shape[1], dptr(graph_ptr, torch.int32), G, dptr(out), stream_ptr()),
              "adkf_seg_mean_fwd")
        ctx.save_for_backward(graph_ptr)
        ctx.meta = (G, x.shape[0])
        return out

    @staticmethod
    def backward(ctx, dout):
        lib = load()
        (graph_ptr,) = ctx.saved_tensors
        G, V = ctx.meta
        dout = dout.contiguous()
        dx = torch.zeros(V, dout.shape[1], dtype=torch.float32, device=dout.device)
        check(lib.adkf_seg_mean_bwd(dptr(dout, torch.float32), dout.shape[1], dptr(graph_ptr), G, dptr(dx), stream_ptr()),
              "adkf_seg_mean_bwd")
        return dx, None, None


# ----------------------------------------------------------------------------------------------------------
# encoder modules (reference state-dict names)
# ----------------------------------------------------------------------------------------------------------
class GINConv(nn.Module):
    def __init__(self, emb_dim: int, aggr: str = "add"):
        super().__init__()
        if aggr != "add":
            raise NotImplementedError("GINConv: only aggr='add' (the reference's setting) has a kernel")
        self.mlp = nn.Sequential(nn.Linear(emb_dim, 2 * emb_dim), nn.ReLU(), nn.Linear(2 * emb_dim, emb_dim))
        self.edge_embedding1 = nn.Embedding(NUM_BOND_TYPE, emb_dim)
        self.edge_embedding2 = nn.Embedding(NUM_BOND_DIRECTION, emb_dim)
        nn.init.xavier_uniform_(self.edge_embedding1.weight.data)
        nn.init.xavier_uniform_(self.edge_embedding2.weight.data)

    def forward(self, h, edge_attr, gi: GinIndex):
        agg = _GinAggregate.apply(h, self.edge_embedding1.weight, self.edge_embedding2.weight, edge_attr, gi)
        hid = ops.linear_tc(agg, self.mlp[0].weight, self.mlp[0].bias, "relu")
        return ops.linear_tc(hid, self.mlp[2].weight, self.mlp[2].bias)


class GNN(nn.Module):
    def __init__(self, num_layer, emb_dim, JK="last", drop_ratio=0.0, gnn_type="gin", batch_norm=True):
        super().__init__()
        if num_layer < 2:
            raise ValueError("Number of GNN layers must be greater than 1.")
        if gnn_type != "gin" or JK != "last" or not batch_norm:
            raise NotImplementedError("only gnn_type='gin', JK='last', batch_norm=True (main_adkf.py defaults) have sm_100a kernels")
        self.num_layer, self.drop_ratio, self.JK = num_layer, float(drop_ratio), JK
        self.x_embedding1 = nn.Embedding(NUM_ATOM_TYPE, emb_dim)
        self.x_embedding2 = nn.Embedding(NUM_CHIRALITY_TAG, emb_dim)
        nn.init.xavier_uniform_(self.x_embedding1.weight.data)
        nn.init.xavier_uniform_(self.x_embedding2.weight.data)
        self.gnns = nn.ModuleList(GINConv(emb_dim) for _ in range(num_layer))
        self.use_batch_norm = True
        self.batch_norms = nn.ModuleList(nn.BatchNorm1d(emb_dim) for _ in range(num_layer))
        self._dropout_step = 0

    def forward(self, x, edge_attr, gi: GinIndex, seg_ptr, seed: Optional[int] = None):
        """x [V,2] int64 -> node representations [V, emb_dim].  Training mode: BatchNorm statistics per segment of
        `seg_ptr`, running estimates updated once per segment in segment order (momentum 0.1, unbiased variance)."""
        h = _EmbedSum2.apply(self.x_embedding1.weight, self.x_embedding2.weight, x, gi.onehot_atom, gi.onehot_chir)
        keep = 1.0 - self.drop_ratio if self.training else 1.0
        if seed is None:
            self._dropout_step += 1
            seed = self._dropout_step
        for layer in range(self.num_layer):
            h = self.gnns[layer](h, edge_attr, gi)
            bn = self.batch_norms[layer]
            act = 0 if layer == self.num_layer - 1 else 1        # no ReLU after the last layer (encoder.py:276-281)
            if self.training:
                stats: list = []
                h = _SegBatchNorm.apply(h, bn.weight, bn.bias, seg_ptr, act, keep, seed * 131 + layer, stats)
                self._update_running(bn, *stats[0])
            else:
                h = seg_bn_eval(h, bn.weight, bn.bias, bn.running_mean, bn.running_var, seg_ptr, act)
        return h

    @staticmethod
    @torch.no_grad()
    def _update_running(bn: nn.BatchNorm1d, mean: torch.Tensor, uvar: torch.Tensor):
        """S sequential exponential updates r <- (1-m) r + m stat_s folded into one weighted sum."""
        S = mean.shape[0]
        m = BN_MOMENTUM
        w = m * (1.0 - m) ** torch.arange(S - 1, -1, -1, device=mean.device, dtype=torch.float32)
        decay = (1.0 - m) ** S
        bn.running_mean.mul_(decay).add_((w[:, None] * mean).sum(0))
        bn.running_var.mul_(decay).add_((w[:, None] * uvar).sum(0))
        bn.num_batches_tracked += S


class GNN_Encoder(nn.Module):
    def __init__(self, num_layer, emb_dim, JK="last", drop_ratio=0, graph_pooling="mean", gnn_type="gin", batch_norm=True):
        super().__init__()
        if graph_pooling != "mean":
            raise NotImplementedError("only graph_pooling='mean' has an sm_100a kernel")
        self.num_layer, self.drop_ratio, self.JK, self.emb_dim = num_layer, drop_ratio, JK, emb_dim
        self.gnn = GNN(num_layer, emb_dim, JK, drop_ratio, gnn_type=gnn_type, batch_norm=batch_norm)

    def from_pretrained(self, model_file, gpu_id=0):
        self.gnn.load_state_dict(torch.load(model_file, map_location="cpu"))

    def encode_merged(self, mb: MergedMolBatch, seed: Optional[int] = None):
        if mb.index is None:
            mb.index = build_gin_index(mb.x, mb.edge_index, mb.edge_attr)
        node = self.gnn(mb.x, mb.edge_attr, mb.index, mb.seg_ptr, seed)
        return _SegMean.apply(node, mb.graph_ptr, mb.num_graphs), node

    def forward(self, *argv):
        """(x, edge_index, edge_attr, batch) or a MolBatch -> (graph_representation, node_representation); one BatchNorm
        segment (one encoder call of the reference)."""
        if len(argv) == 1:
            d = argv[0]
            x, edge_index, edge_attr, batch = d.x, d.edge_index, d.edge_attr, d.batch
        elif len(argv) == 4:
            x, edge_index, edge_attr, batch = argv
        else:
            raise ValueError("unmatched number of arguments.")
        if not x.is_cuda:
            raise RuntimeError("adkf_ift_b200 kernels need CUDA tensors (no CPU fallback)")
        G = int(batch.max().item()) + 1 if batch.numel() else 0
        counts = torch.bincount(batch, minlength=G)
        gp_ = torch.zeros(G + 1, dtype=torch.int32, device=x.device)
        gp_[1:] = torch.cumsum(counts, 0).to(torch.int32)
        seg = torch.tensor([0, x.shape[0]], dtype=torch.int32, device=x.device)
        gi = build_gin_index(x, edge_index, edge_attr)
        node = self.gnn(x, edge_attr, gi, seg)
        return _SegMean.apply(node, gp_, G), node


# ----------------------------------------------------------------------------------------------------------
# model + trainer
# ----------------------------------------------------------------------------------------------------------
def default_args(**over):
    """The fields of MoleculeNet/parser.py that the model / trainer read, with the parser's defaults."""
    a = dict(emb_dim=300, enc_layer=5, JK="last", dropout=0.5, enc_pooling="mean", enc_gnn="gin", enc_batch_norm=1, pretrained=0,
             pretrained_weight_path="", gpu_id=0, meta_lr=0.001, weight_decay=5e-5, batch_task=9, update_step=1,
             update_step_test=1, n_shot_train=10, n_shot_test=10, n_query=16)
    a.update(over)
    return SimpleNamespace(**a)


class ADKFModel(nn.Module):
    def __init__(self, args):
        super().__init__()
        self.gp_kernel = "matern"
        self.emb_dim = args.emb_dim
        self.gpu_id = getattr(args, "gpu_id", 0)
        self.mol_encoder = GNN_Encoder(num_layer=args.enc_layer, emb_dim=args.emb_dim, JK=args.JK, drop_ratio=args.dropout,
                                       graph_pooling=args.enc_pooling, gnn_type=args.enc_gnn,
                                       batch_norm=bool(args.enc_batch_norm))
        if getattr(args, "pretrained", 0):
            self.mol_encoder.from_pretrained(args.pretrained_weight_path, self.gpu_id)
        self._gp_cfg = SimpleNamespace(gp_kernel=self.gp_kernel, use_numeric_labels=False)
        self._create_tail_gp()
        self._bound = None

    def feature_extractor_params(self) -> List[nn.Parameter]:
        return [p for n, p in self.named_parameters() if not n.startswith("gp_")]

    def gp_params(self) -> List[nn.Parameter]:
        return [p for n, p in self.named_parameters() if n.startswith("gp_")]

    @property
    def device(self):
        return next(self.parameters()).device

    def _create_tail_gp(self):
        dev = next(self.parameters()).device
        self.gp_likelihood = GaussianLikelihoodParams().to(dev)
        self.gp_model = ExactGPLayer(likelihood=self.gp_likelihood, kernel=self.gp_kernel, use_numeric_labels=False).to(dev)
        self.mll = MarginalLogLikelihood(self)

    def compute_median_lengthscale_init(self, gp_input):
        n = gp_input.shape[0]
        g = gpk.BatchedExactGP(gp_input.float(), torch.zeros(n, device=gp_input.device),
                               ops.TaskLayout.build([n], [0], gp_input.device), self.gp_kernel)
        return g.init().aux[0, ops.AUX_L0]

    def reinit_gp_params(self, gp_input, use_lengthscale_prior=True, targets=None):
        self._create_tail_gp()
        n = gp_input.shape[0]
        y = targets if targets is not None else torch.zeros(n, device=gp_input.device)
        g = gpk.BatchedExactGP(gp_input.float(), y.float(), ops.TaskLayout.build([n], [0], gp_input.device), self.gp_kernel,
                               False, use_lengthscale_prior)
        init = g.init()
        self._write_phi(init.phi[0])
        self.gp_model.covar_module.base_kernel.lengthscale_prior_loc = init.aux[0, ops.AUX_LS_LOC] if use_lengthscale_prior else None
        self._bound = dict(gp=g, fit=init)

    def _write_phi(self, phi3):
        with torch.no_grad():
            self.gp_likelihood.noise_covar.raw_noise.copy_(phi3[0:1])
            self.gp_model.covar_module.raw_outputscale.copy_(phi3[1])
            self.gp_model.covar_module.base_kernel.raw_lengthscale.copy_(phi3[2].view(1, 1))

    def _read_phi(self):
        return torch.stack([p.detach().reshape(()) for p in self.gp_params()]).view(1, 3).float().contiguous()

    def fit_gp(self, max_iter=50, gtol=2e-6):
        b = self._bound
        res = b["gp"].fit(start=gpk.FitResult(self._read_phi(), b["fit"].aux, b["fit"].status), max_iter=max_iter, gtol=gtol)
        self._write_phi(res.phi[0])
        b["fit"] = res
        return res

    @staticmethod
    def _convert_bool_labels(labels):
        return (labels.float() - 0.5) * 2.0

    def forward(self, s_data, q_data, train_loss: bool, s_label=None, q_pred_adj=False, predictive_val_loss: bool = False,
                is_functional_call: bool = False):
        """adkf_model.py:98-133 (training-mode branches) on the T = 1 case of the batched kernels."""
        assert self.training and train_loss is not None
        zs, _ = self.mol_encoder(s_data.x, s_data.edge_index, s_data.edge_attr, s_data.batch)
        ys = self._convert_bool_labels(s_label)
        n = zs.shape[0]
        phi = self.gp_params()
        ls_loc = self.gp_model.covar_module.base_kernel.lengthscale_prior_loc
        if train_loss:
            if is_functional_call:
                return _TaskGPLoss.apply(zs, phi[0], phi[1], phi[2], ys, ops.TaskLayout.build([n], [0], zs.device), self._gp_cfg,
                                         ls_loc, False)
            self.reinit_gp_params(zs.detach(), use_lengthscale_prior=True, targets=ys.detach())
            self.gp_model.set_train_data(inputs=zs.detach(), targets=ys.detach(), strict=False)
            return None
        assert is_functional_call
        if not predictive_val_loss:
            raise NotImplementedError
        zq, _ = self.mol_encoder(q_data.x, q_data.edge_index, q_data.edge_attr, q_data.batch)
        yq = self._convert_bool_labels(q_data.y)
        layout = ops.TaskLayout.build([n], [zq.shape[0]], zs.device)
        return _TaskGPLoss.apply(torch.cat([zs, zq]), phi[0], phi[1], phi[2], torch.cat([ys, yq]), layout, self._gp_cfg, ls_loc, True)

    def forward_query_loader(self, s_data, q_loader, train_loss, s_label=None, **_):
        """adkf_model.py:136-160: evaluation-mode posterior means of every query batch -> (sigmoid(mean), labels)."""
        assert not self.training and train_loss is None
        with torch.no_grad():
            zs, _ = self.mol_encoder(s_data.x, s_data.edge_index, s_data.edge_attr, s_data.batch)
            ys = self._convert_bool_labels(s_label)
            self.gp_model.set_train_data(inputs=zs, targets=ys, strict=False)
            phi = self._read_phi()
            outs, labels = [], []
            for q in q_loader:
                q = q.to(zs.device)
                labels.append(q.y)
                zq, _ = self.mol_encoder(q.x, q.edge_index, q.edge_attr, q.batch)
                layout = ops.TaskLayout.build([zs.shape[0]], [zq.shape[0]], zs.device)
                g = gpk.BatchedExactGP(torch.cat([zs, zq]), torch.cat([ys, torch.zeros(zq.shape[0], device=zs.device)]), layout,
                                       self.gp_kernel)
                outs.append(g.predict(phi)[0][zs.shape[0]:])
        return torch.sigmoid(torch.cat(outs, 0)), torch.cat(labels, 0)


@dataclass
class MolTaskBatchResult:
    phi: torch.Tensor
    status: torch.Tensor
    aux: torch.Tensor
    outer_loss: Optional[torch.Tensor] = None
    loss_per_sample: Optional[torch.Tensor] = None
    mean: Optional[torch.Tensor] = None
    variance: Optional[torch.Tensor] = None
    preds: Optional[torch.Tensor] = None


def run_on_mol_task_batch(model: ADKFModel, mb: MergedMolBatch, train: bool, grad_scale: Optional[float] = None,
                          phi: Optional[torch.Tensor] = None, seed: Optional[int] = None, fit_max_iter: int = 50,
                          fit_gtol: float = 2e-6) -> MolTaskBatchResult:
    """T MoleculeNet tasks in one pass.  train=True: inner fit + IFT hypergradient of every task, accumulated (x 1/T) into
    `.grad` of the encoder parameters (adkfift_trainer.py:163-213).  train=False: encoder in its CURRENT mode (the reference
    fits on train-mode features and predicts with eval-mode ones, :279-286 -- call twice if that is wanted), fit, predict."""
    if train:
        model.train()
        Z, _ = model.mol_encoder.encode_merged(mb, seed)
        g = gpk.BatchedExactGP(Z, mb.labels, mb.layout, model.gp_kernel, False, True, qq_block=True)
        if phi is None:
            fit = g.fit(max_iter=fit_max_iter, gtol=fit_gtol)
        else:
            fit = g.evaluate(phi.view(-1, 3).float().contiguous(), g.init().aux)
        dZ, aux, status = g.hypergradient(fit, mode=0)
        Z.backward(dZ * ((1.0 / mb.num_tasks) if grad_scale is None else grad_scale))
        F_t = aux[:, ops.AUX_FOUTER]
        n_q = torch.tensor([r - s for r, s in zip(mb.layout.n_rows_host, mb.layout.n_sup_host)], dtype=torch.float32,
                           device=Z.device)
        return MolTaskBatchResult(fit.phi, status | fit.status, aux, F_t, F_t / n_q.clamp_min(1))
    with torch.no_grad():
        Z, _ = model.mol_encoder.encode_merged(mb, seed)
        g = gpk.BatchedExactGP(Z, mb.labels, mb.layout, model.gp_kernel, False, True, qq_block=False)
        fit = g.fit(max_iter=fit_max_iter, gtol=fit_gtol) if phi is None else gpk.FitResult(phi.view(-1, 3).float().contiguous(), None, None)
        mean, var, st = g.predict(fit.phi)
    return MolTaskBatchResult(fit.phi, st, fit.aux, mean=mean, variance=var, preds=torch.sigmoid(mean))


class ADKF_Meta_Trainer(nn.Module):
    """Optimiser + outer step of adkfift_trainer.py (AdamW lr 1e-3, wd 5e-5, clip 1.0); data loading stays with the caller."""

    def __init__(self, args, model: ADKFModel):
        super().__init__()
        self.args, self.model = args, model
        self.optimizer = torch.optim.AdamW(model.feature_extractor_params(), lr=args.meta_lr, weight_decay=args.weight_decay)
        self.train_epoch = 0

    def train_step(self, mb: MergedMolBatch) -> MolTaskBatchResult:
        self.train_epoch += 1
        res = None
        for _ in range(self.args.update_step):
            self.optimizer.zero_grad(set_to_none=True)
            res = run_on_mol_task_batch(self.model, mb, train=True)
            torch.nn.utils.clip_grad_norm_(self.model.feature_extractor_params(), 1.0)
            self.optimizer.step()
            ops.weights_changed()
        return res
