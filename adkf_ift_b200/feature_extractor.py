"""Graph feature extractor of ADKF-IFT on the fused sm_100a kernels.

Call surface and state-dict names follow the reference so its checkpoints load unchanged:
  GNNConfig / GNNBlock / GNN              fs_mol/modules/gnn.py:16-28, 390-515, 518-556
  GraphReadoutConfig / combined readout   fs_mol/modules/graph_readout.py:14-28, 119-296
  GraphFeatureExtractor(Config)           fs_mol/modules/graph_feature_extractor.py:20-25, 43-98
  MLP                                     fs_mol/modules/mlp.py:6-22

What differs is how the forward runs.  The reference loops over 4 towers x 3 edge types with gather / cat /
Linear / relu / five torch_scatter calls each; here one GNN block is
  adkf_mp_pna_fwd (all towers, messages never leave the SM, CSR-ordered aggregation)  ->  msg_out_projection
  -> ReZero residual -> LayerNorm -> BOOM -> residual,
and the readout's scatter_softmax / index_add_ / scatter(max) are the segment kernels of readout_ecfp.cu.
The graph index (CSR by target node of the bidirectional edge set) is built once per batch on the device
(adkf_graph_csr_build) and shared by all blocks.  Only the default PNA / combined-readout configuration
of the CLI is implemented; anything else raises instead of silently taking a different code path.
"""
from dataclasses import dataclass, field
from typing import List, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .fsmol_batcher import NUM_EDGE_TYPES, NUM_NODE_FEATURES


@dataclass(frozen=True)
class GNNConfig:
    """Field names/defaults of fs_mol/modules/gnn.py:16-28 (the CLI overrides per_head_dim=64,
    intermediate_dim=1024, num_layers=10 -- gnn.py:47-63)."""
    type: str = "PNA"
    num_edge_types: int = NUM_EDGE_TYPES
    hidden_dim: int = 128
    num_heads: int = 4
    per_head_dim: int = 32
    intermediate_dim: int = 512
    message_function_depth: int = 1
    num_layers: int = 8
    dropout_rate: float = 0.0
    use_rezero_scaling: bool = True
    make_edges_bidirectional: bool = True


@dataclass(frozen=True)
class GraphReadoutConfig:
    readout_type: str = "combined"
    use_all_states: bool = True
    num_heads: int = 12
    head_dim: int = 64
    output_dim: int = 512


@dataclass(frozen=True)
class GraphFeatureExtractorConfig:
    initial_node_feature_dim: int = NUM_NODE_FEATURES
    gnn_config: GNNConfig = field(default_factory=GNNConfig)
    readout_config: GraphReadoutConfig = field(default_factory=GraphReadoutConfig)
    output_norm: str = "off"


def cli_default_config(num_layers: int = 10, gnn_type: str = "PNA") -> GraphFeatureExtractorConfig:
    """The configuration `adaptive_dkt_train.py` builds from its argparse defaults (`--gnn_type` PNA | MultiAggr | Plain)."""
    return GraphFeatureExtractorConfig(
        gnn_config=GNNConfig(type=gnn_type, hidden_dim=128, num_heads=4, per_head_dim=64, intermediate_dim=1024,
                             num_layers=num_layers),
        readout_config=GraphReadoutConfig())


class MLP(nn.Module):
    """Linear/activation stack; parameters live under `_layers.{0,2,..}` like the reference's."""

    def __init__(self, input_dim: int, out_dim: int, hidden_layer_dims: List[int]):
        super().__init__()
        mods: List[nn.Module] = []
        d = input_dim
        for h in hidden_layer_dims:
            mods += [nn.Linear(d, h), nn.ReLU()]
            d = h
        mods.append(nn.Linear(d, out_dim))
        self._layers = nn.Sequential(*mods)

    def forward(self, x):
        mods = list(self._layers)
        if len(mods) == 3 and isinstance(mods[1], nn.ReLU):      # one hidden layer: fused two-layer backward when eligible
            return ops.mlp2_tc(x, mods[0].weight, mods[0].bias, "relu", mods[2].weight, mods[2].bias)
        i = 0
        while i < len(mods):
            fuse_relu = i + 1 < len(mods) and isinstance(mods[i + 1], nn.ReLU)
            x = ops.linear_tc(x, mods[i].weight, mods[i].bias, "relu" if fuse_relu else None)
            i += 2 if fuse_relu else 1
        return x


class _TowerMessageFns(nn.Module):
    """Parameter holder for one tower: `message_fns.{t}._layers.0.{weight,bias}` ([192,64], [192])."""

    def __init__(self, tower_in: int, msg_dim: int, num_edge_types: int):
        super().__init__()
        self.message_fns = nn.ModuleList(MLP(2 * tower_in, msg_dim, []) for _ in range(num_edge_types))


class BOOMLayer(nn.Module):
    def __init__(self, inout_dim: int, intermediate_dim: int):
        super().__init__()
        self.linear1 = nn.Linear(inout_dim, intermediate_dim)
        self.linear2 = nn.Linear(intermediate_dim, inout_dim)

    def forward(self, x):
        return ops.mlp2_tc(x, self.linear1.weight, self.linear1.bias, "leaky_relu", self.linear2.weight, self.linear2.bias)


class RelationalMultiHeadAttentionMP(nn.Module):
    """Parameter holder with the reference's names (gnn.py:268-310): per edge type a message MLP Linear(2 hidden -> heads x
    head_dim), a bias-free target-node query projection and a bias-free neighbour-node key projection."""

    def __init__(self, hidden_dim: int, num_heads: int, per_head_dim: int, num_edge_types: int):
        super().__init__()
        out = num_heads * per_head_dim
        self.message_fns = nn.ModuleList(MLP(2 * hidden_dim, out, []) for _ in range(num_edge_types))
        self.target_node_query_projs = nn.ModuleList(nn.Linear(hidden_dim, out, bias=False) for _ in range(num_edge_types))
        self.neighbour_node_key_projs = nn.ModuleList(nn.Linear(hidden_dim, out, bias=False) for _ in range(num_edge_types))

    def forward(self, x, gi: ops.GraphIndex):
        Wqk = torch.cat([p.weight for p in self.target_node_query_projs] + [p.weight for p in self.neighbour_node_key_projs], 0)
        return ops.mha_message_passing(x, [f._layers[0].weight for f in self.message_fns],
                                       [f._layers[0].bias for f in self.message_fns], Wqk, gi)


class _ReZeroResidual(torch.autograd.Function):
    """h + alpha * x (gnn.py:497-498, 512-513): one streaming kernel forward; backward alpha * g and <g, x> in ONE pass over g
    (csrc/rezero.cu; torch needed addcmul with a broadcast scalar -- its slow non-vectorised path --, a mul and a cuBLAS dot).
    Every tensor is [V,128] and purely HBM-bound."""

    @staticmethod
    def forward(ctx, h, x, alpha):
        ctx.save_for_backward(x, alpha)
        return ops.rezero_fwd(h, x, alpha)

    @staticmethod
    def backward(ctx, g):
        x, alpha = ctx.saved_tensors
        dx, dalpha = ops.rezero_bwd(g, x, alpha, ctx.needs_input_grad[1], ctx.needs_input_grad[2])
        return g, dx, dalpha


class GNNBlock(nn.Module):
    """One residual block (gnn.py:477-515).  `mp_norm_layer` is registered but never applied, exactly like
    the reference, so the state dict keeps its keys."""

    def __init__(self, config: GNNConfig):
        super().__init__()
        kind = config.type.lower()
        if kind not in ("pna", "multiaggr", "plain", "multiheadattention"):
            raise ValueError(f"Unknown GNN type {config.type}.")          # gnn.py:447
        self._mp_mode = {"pna": ops.MP_PNA, "multiaggr": ops.MP_MULTIAGGR, "plain": ops.MP_PLAIN, "multiheadattention": -1}[kind]
        if (config.hidden_dim, config.num_heads, config.per_head_dim, config.num_edge_types) != (128, 4, 64, 3):
            raise NotImplementedError("the fused PNA kernel is specialised to hidden 128 / 4 towers / per-head 64 / 3 edge types")
        if config.message_function_depth != 1 or config.dropout_rate != 0.0 or not config.use_rezero_scaling:
            raise NotImplementedError("message_function_depth=1, dropout 0 and ReZero scaling are required")
        if config.intermediate_dim <= 0:
            raise NotImplementedError("BOOM layer cannot be disabled")
        self.config = config
        self.alpha = nn.Parameter(torch.full((1,), 1e-7))
        tower_in = config.hidden_dim // config.num_heads
        # gnn.py:421-447: PNA / MultiAggr towers emit 3 x per_head_dim messages aggregated as sum|mean|std|max (x 3 scalers for
        # PNA); Plain towers emit per_head_dim messages aggregated by sum
        msg_dim = config.per_head_dim if kind == "plain" else 3 * config.per_head_dim
        tower_out = {"pna": 12, "multiaggr": 4, "plain": 1, "multiheadattention": 1}[kind] * config.per_head_dim
        if kind == "multiheadattention":      # one attention layer over the whole hidden state (gnn.py:409-420)
            self.mp_layers = nn.ModuleList([RelationalMultiHeadAttentionMP(config.hidden_dim, config.num_heads, config.per_head_dim,
                                                                           config.num_edge_types)])
        else:
            self.mp_layers = nn.ModuleList(_TowerMessageFns(tower_in, msg_dim, config.num_edge_types)
                                           for _ in range(config.num_heads))
        self.msg_out_projection = nn.Linear(config.num_heads * tower_out, config.hidden_dim)
        self.mp_norm_layer = nn.LayerNorm(config.hidden_dim)
        self.boom_layer = BOOMLayer(config.hidden_dim, config.intermediate_dim)
        self.boom_norm_layer = nn.LayerNorm(config.hidden_dim)

    def _boom_norm(self, h):
        ln = self.boom_norm_layer
        return ops.layernorm128(h, ln.weight, ln.bias, ln.eps)

    def stacked_message_weights(self):
        W = torch.stack([torch.stack([fn._layers[0].weight for fn in tw.message_fns]) for tw in self.mp_layers])
        b = torch.stack([torch.stack([fn._layers[0].bias for fn in tw.message_fns]) for tw in self.mp_layers])
        return W, b  # [4,3,192,64], [4,3,192]

    def forward(self, h, gi: ops.GraphIndex):
        if self._mp_mode < 0:
            agg = self.mp_layers[0](h, gi)                                 # [V, 256]
            h = _ReZeroResidual.apply(h, ops.linear_tc(agg, self.msg_out_projection.weight, self.msg_out_projection.bias), self.alpha)
            return _ReZeroResidual.apply(h, self.boom_layer(self._boom_norm(h)), self.alpha)
        W, b = self.stacked_message_weights()
        if self._mp_mode == ops.MP_PNA and getattr(gi, "type_sorted", None) is not None:
            # degree scalers folded out of the aggregate: [V,1024] instead of [V,3072] between aggregation and projection
            proj = ops.pna_message_passing_projected(h, W, b, self.msg_out_projection.weight, self.msg_out_projection.bias, gi)
        else:
            agg = ops.pna_message_passing(h, W, b, gi) if self._mp_mode == ops.MP_PNA else ops.mp_message_passing(h, W, b, gi, self._mp_mode)   # [V, 3072 | 1024 | 256]
            proj = ops.linear_tc(agg, self.msg_out_projection.weight, self.msg_out_projection.bias)
        h = _ReZeroResidual.apply(h, proj, self.alpha)
        return _ReZeroResidual.apply(h, self.boom_layer(self._boom_norm(h)), self.alpha)


class GNN(nn.Module):
    def __init__(self, config: GNNConfig):
        super().__init__()
        if not config.make_edges_bidirectional:
            raise NotImplementedError("the CSR builder always expands bonds in both directions")
        self.config = config
        self.gnn_blocks = nn.ModuleList(GNNBlock(config) for _ in range(config.num_layers))

    def forward(self, h, gi: ops.GraphIndex) -> List[torch.Tensor]:
        states = [h]
        for blk in self.gnn_blocks:
            h = blk(h, gi)
            states.append(h)
        return states


class _WeightedPooler(nn.Module):
    def __init__(self, node_dim, out_dim, num_heads, head_dim, mode: int):
        super().__init__()
        self._mode, self._num_heads, self._head_dim = mode, num_heads, head_dim
        self._scoring_module = MLP(node_dim, num_heads, [num_heads * head_dim])
        self._transformation_mlp = MLP(node_dim, num_heads * head_dim, [num_heads * head_dim])
        self._combination_layer = nn.Linear(num_heads * head_dim, out_dim, bias=False)

    def forward(self, x, gi: ops.GraphIndex):
        pooled = ops.readout_weighted(self._scoring_module(x), self._transformation_mlp(x), gi.graph_ptr,
                                      gi.num_graphs, self._num_heads, self._head_dim, self._mode)
        return ops.linear_tc(pooled, self._combination_layer.weight)


class _MaxPooler(nn.Module):
    def __init__(self, node_dim, out_dim):
        super().__init__()
        self._combination_layer = nn.Linear(node_dim, out_dim, bias=False)

    def forward(self, x, gi: ops.GraphIndex):
        return ops.linear_tc(ops.readout_max(x, gi.graph_ptr, gi.num_graphs), self._combination_layer.weight)


class CombinedGraphReadout(nn.Module):
    """weighted-mean (per-graph softmax) + weighted-sum (sigmoid) + max pooling (graph_readout.py:119-177)."""

    def __init__(self, node_dim, out_dim, num_heads, head_dim):
        super().__init__()
        self._weighted_mean_pooler = _WeightedPooler(node_dim, out_dim, num_heads, head_dim, mode=0)
        self._weighted_sum_pooler = _WeightedPooler(node_dim, out_dim, num_heads, head_dim, mode=1)
        self._max_pooler = _MaxPooler(node_dim, out_dim)
        self._combination_layer = nn.Linear(3 * out_dim, out_dim, bias=False)

    def forward(self, x, gi: ops.GraphIndex):
        raw = torch.cat((self._weighted_mean_pooler(x, gi), self._weighted_sum_pooler(x, gi),
                         self._max_pooler(x, gi)), dim=1)
        return ops.linear_tc(F.relu(raw), self._combination_layer.weight)


class GraphFeatureExtractor(nn.Module):
    """forward(batch) -> [G, 512].  `batch` needs `.node_features [V,32] f32`, `.adjacency_lists` (3 x [E_t,2]
    int64, each bond once), `.node_to_graph [V] int64`, `.num_graphs` -- CUDA tensors, as the reference's
    FSMolBatch after torchify -- or a prebuilt `.graph_index` (see task_batch.py)."""

    def __init__(self, config: GraphFeatureExtractorConfig):
        super().__init__()
        if config.output_norm != "off":
            raise NotImplementedError("output_norm != 'off' is not on the ADKF-IFT path")
        rc = config.readout_config
        if rc.readout_type != "combined" or not rc.use_all_states:
            raise NotImplementedError("only the default combined / all-states readout has sm_100a kernels")
        self.config = config
        self.init_node_proj = nn.Linear(config.initial_node_feature_dim, config.gnn_config.hidden_dim, bias=False)
        self.gnn = GNN(config.gnn_config)
        node_dim = (config.gnn_config.num_layers + 1) * config.gnn_config.hidden_dim
        self.readout = CombinedGraphReadout(node_dim, rc.output_dim, rc.num_heads, rc.head_dim)
        self.final_norm_layer: Optional[nn.Module] = None

    def forward(self, batch) -> torch.Tensor:
        gi = getattr(batch, "graph_index", None)
        if gi is None:
            gi = ops.build_graph_index(list(batch.adjacency_lists), int(batch.node_features.shape[0]),
                                       batch.node_to_graph, int(batch.num_graphs))
        h0 = ops.linear_tc(batch.node_features, self.init_node_proj.weight, self.init_node_proj.bias)
        states = self.gnn(h0, gi)
        return self.readout(torch.cat(states, dim=-1), gi)
