"""Graph collation for FS-Mol-shaped molecules (host side, numpy, integer path bit-exact).

Mirrors the interface of the reference's `fs_mol/data/fsmol_batcher.py` (FSMolBatch :23-58,
fsmol_batch_finalizer :61-94, FSMolBatcher.batch :174-218) and the light-weight containers of
`fs_mol/data/fsmol_task.py:19-57` (GraphData, MoleculeDatapoint) / FSMolTaskSample, but collates
with vectorised offset arithmetic (cumsum + repeat) instead of per-graph list appends.  Results are
identical arrays: `adjacency + node offset` per edge type, `node_to_graph`, concatenated features;
a trailing batch holding a single graph is dropped exactly like the reference (:217).
"""
import math
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, Generic, Iterable, Iterator, List, Optional, Tuple, TypeVar

import numpy as np

NUM_EDGE_TYPES = 3  # fs_mol/data/fsmol_dataset.py:21
NUM_NODE_FEATURES = 32  # fs_mol/data/fsmol_dataset.py:22


@dataclass
class GraphData:
    node_features: np.ndarray  # [V, F] float32
    adjacency_lists: List[np.ndarray]  # NUM_EDGE_TYPES x [E_t, 2] int64 (each bond stored once)
    edge_features: List[np.ndarray] = field(default_factory=list)


@dataclass(frozen=True)
class MoleculeDatapoint:
    task_name: str
    smiles: str
    graph: GraphData
    numeric_label: float
    bool_label: bool
    fingerprint: Optional[np.ndarray]  # [2048] int32 counts
    descriptors: Optional[np.ndarray]  # [42] float32

    def get_fingerprint(self) -> np.ndarray:
        if self.fingerprint is None:
            raise ValueError("fingerprints must be precomputed (rdkit featurisation is out of scope)")
        return self.fingerprint

    def get_descriptors(self) -> np.ndarray:
        if self.descriptors is None:
            raise ValueError("descriptors must be precomputed (rdkit featurisation is out of scope)")
        return self.descriptors


@dataclass(frozen=True)
class FSMolTaskSample:
    name: str
    train_samples: List[MoleculeDatapoint]
    valid_samples: List[MoleculeDatapoint]
    test_samples: List[MoleculeDatapoint]


@dataclass(frozen=True)
class FSMolBatch:
    num_graphs: int
    num_nodes: int
    num_edges: int
    node_features: np.ndarray  # [V, F] float32
    adjacency_lists: List[np.ndarray]  # per edge type [E_t, 2] int64, node ids offset into the batch
    edge_features: List[np.ndarray]
    node_to_graph: np.ndarray  # [V] int64


BatchFeatureType = TypeVar("BatchFeatureType")
BatchLabelType = TypeVar("BatchLabelType")


def collate_graphs(graphs: List[GraphData]) -> FSMolBatch:
    """One disconnected graph from `graphs` (vectorised equivalent of the reference's
    per-datapoint appends + fsmol_batch_finalizer)."""
    num_graphs = len(graphs)
    node_counts = np.fromiter((len(g.node_features) for g in graphs), dtype=np.int64, count=num_graphs)
    node_offsets = np.zeros(num_graphs, dtype=np.int64)
    if num_graphs > 1:
        np.cumsum(node_counts[:-1], out=node_offsets[1:])
    adjacency_lists, edge_features = [], []
    total_edges = 0
    for t in range(NUM_EDGE_TYPES):
        parts = [np.asarray(g.adjacency_lists[t], dtype=np.int64).reshape(-1, 2) for g in graphs]
        counts = np.fromiter((len(p) for p in parts), dtype=np.int64, count=num_graphs)
        if num_graphs and counts.sum() > 0:
            adj = np.concatenate(parts, axis=0) + np.repeat(node_offsets, counts)[:, None]
        elif num_graphs:
            adj = np.concatenate(parts, axis=0) if parts else np.zeros((0, 2), dtype=np.int64)
        else:
            adj = np.zeros((0, 2), dtype=np.int64)
        total_edges += int(counts.sum())
        adjacency_lists.append(adj.astype(np.int64, copy=False))
        feats = [g.edge_features[t] for g in graphs if len(g.edge_features) > t]
        if len(feats) > 0:
            edge_features.append(np.concatenate(feats, axis=0))
        else:
            edge_features.append(np.zeros((adj.shape[0], 0), dtype=np.float32))
    return FSMolBatch(
        num_graphs=num_graphs,
        num_nodes=int(node_counts.sum()),
        num_edges=total_edges,
        node_features=np.concatenate([g.node_features for g in graphs], axis=0),
        adjacency_lists=adjacency_lists,
        edge_features=edge_features,
        node_to_graph=np.repeat(np.arange(num_graphs, dtype=np.int64), node_counts),
    )


def fsmol_batch_finalizer(batch_data: Dict[str, Any]) -> FSMolBatch:
    return collate_graphs(batch_data["graphs"])


class FSMolBatcher(Generic[BatchFeatureType, BatchLabelType]):
    """Splits a datapoint stream into batches bounded by graphs / nodes / edges
    (reference: fs_mol/data/fsmol_batcher.py:97-218)."""

    def __init__(
        self,
        max_num_graphs: Optional[int] = None,
        max_num_nodes: Optional[int] = None,
        max_num_edges: Optional[int] = None,
        init_callback: Optional[Callable[[Dict[str, Any]], None]] = None,
        per_datapoint_callback: Optional[Callable[[Dict[str, Any], int, MoleculeDatapoint], None]] = None,
        finalizer_callback: Optional[Callable[[Dict[str, Any]], Tuple[BatchFeatureType, BatchLabelType]]] = None,
    ):
        if max_num_graphs is None and max_num_nodes is None and max_num_edges is None:
            raise ValueError("Need to specify at least one of max_num_graphs, max_num_nodes or max_num_edges!")
        self._max_num_graphs = max_num_graphs or math.inf
        self._max_num_nodes = max_num_nodes or math.inf
        self._max_num_edges = max_num_edges or math.inf
        self._init_callback = init_callback
        self._per_datapoint_callback = per_datapoint_callback
        self._finalizer_callback = finalizer_callback

    def _new_batch(self) -> Dict[str, Any]:
        data: Dict[str, Any] = {"num_graphs": 0, "num_nodes": 0, "num_edges": 0, "graphs": [],
                                "graph_task": [], "bool_labels": [], "numeric_labels": []}
        if self._init_callback is not None:
            self._init_callback(data)
        return data

    def _finalize(self, data: Dict[str, Any]):
        if self._finalizer_callback is not None:
            return self._finalizer_callback(data)
        return fsmol_batch_finalizer(data), np.stack(data["bool_labels"], axis=0)

    def batch(self, datapoints: Iterable[MoleculeDatapoint]) -> Iterator[Tuple[BatchFeatureType, BatchLabelType]]:
        data = self._new_batch()
        for dp in datapoints:
            nv = len(dp.graph.node_features)
            ne = sum(len(a) for a in dp.graph.adjacency_lists)
            full = (data["num_graphs"] + 1 > self._max_num_graphs
                    or data["num_nodes"] + nv > self._max_num_nodes
                    or data["num_edges"] + ne > self._max_num_edges)
            if full:
                yield self._finalize(data)
                data = self._new_batch()
            slot = data["num_graphs"]
            data["graphs"].append(dp.graph)
            data["bool_labels"].append(dp.bool_label)
            data["numeric_labels"].append(dp.numeric_label)
            data["graph_task"].append(dp.task_name)
            data["num_graphs"] += 1
            data["num_nodes"] += nv
            data["num_edges"] += ne
            if self._per_datapoint_callback is not None:
                self._per_datapoint_callback(data, slot, dp)
        if data["num_graphs"] > 1:  # a trailing single-graph batch is dropped (reference :217)
            yield self._finalize(data)
