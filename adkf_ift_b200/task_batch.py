"""Multi-task collation: T few-shot tasks -> ONE disconnected graph batch + one GP row layout.

The reference processes the tasks of a meta-batch strictly one after the other
(fs_mol/utils/adaptive_dkt_utils.py:361-404) and, inside a task, runs the feature extractor once on the
support graphs and once on the query graphs (fs_mol/models/adaptive_dkt.py:141-143).  Here every graph of
every task goes through the extractor in a single pass: per task the support graphs come first and the query
graphs follow, so that row r of the feature matrix Z is graph r and task t owns rows
[z_off[t], z_off[t+1]) with its n_sup[t] support rows first -- the layout the batched GP kernels expect
(include/adkf_b200.h, "Batched GP tail").  The integer arithmetic is the reference collation
(fs_mol/data/fsmol_batcher.py:197-202: adjacency + node offset, node_to_graph fill) applied once more
at task granularity, so it stays bit-exact.

Host arrays are staged in pinned memory; `to_device` issues one async copy per field (7 copies per step).
"""
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import ops
from .dkt import DKTBatch


@dataclass
class HostTaskBatch:
    """Pinned host staging of T collated tasks."""
    num_tasks: int
    n_sup: List[int]
    n_query: List[int]
    num_graphs: int
    num_nodes: int
    node_features: torch.Tensor      # [V, 32] f32
    adjacency: List[torch.Tensor]    # 3 x [E_t, 2] int64
    node_to_graph: torch.Tensor      # [V] int64
    fingerprints: torch.Tensor       # [G, 2048] int32
    descriptors: torch.Tensor        # [G, 42] f32
    labels: torch.Tensor             # [G] f32: +-1 (classification) or standardised numeric
    bool_labels: torch.Tensor        # [G] bool (metrics)
    max_degree: int = -1             # largest in-degree over the bidirectional edge set (host-computed; -1 = unknown): lets the
                                     # extractor pick the fused message-passing kernel without a device round trip

    def nbytes(self) -> int:
        ts = [self.node_features, *self.adjacency, self.node_to_graph, self.fingerprints, self.descriptors,
              self.labels]
        return int(sum(t.numel() * t.element_size() for t in ts))


@dataclass
class DeviceTaskBatch:
    """What the model consumes: duck-types the reference's FSMolBatch fields plus the prebuilt index."""
    num_tasks: int
    num_graphs: int
    num_nodes: int
    node_features: torch.Tensor
    adjacency_lists: List[torch.Tensor]
    node_to_graph: torch.Tensor
    fingerprints: torch.Tensor
    descriptors: torch.Tensor
    labels: torch.Tensor
    layout: ops.TaskLayout
    graph_index: Optional[ops.GraphIndex] = None


def _pin(a: np.ndarray, pin: bool) -> torch.Tensor:
    t = torch.from_numpy(np.ascontiguousarray(a))
    if pin:
        try:
            t = t.pin_memory()
        except RuntimeError:  # no CUDA driver in the authoring container: plain host memory
            pass
    return t


def collate_tasks(batches: Sequence[DKTBatch], use_numeric_labels: bool, pin: bool = True) -> HostTaskBatch:
    """batches: one DKTBatch (host numpy, as produced by dkt.task_sample_to_dkt_task_sample) per task."""
    parts = []
    for b in batches:
        parts.append((b.support_features, b.support_numeric_labels if use_numeric_labels else b.support_labels))
        parts.append((b.query_features, b.query_numeric_labels if use_numeric_labels else b.query_labels))
    feats = [p[0] for p in parts]
    g_counts = np.array([f.num_graphs for f in feats], dtype=np.int64)
    v_counts = np.array([f.num_nodes for f in feats], dtype=np.int64)
    g_off = np.concatenate([[0], np.cumsum(g_counts)])
    v_off = np.concatenate([[0], np.cumsum(v_counts)])
    adjacency = []
    for t in range(len(feats[0].adjacency_lists)):
        adjacency.append(np.concatenate([np.asarray(f.adjacency_lists[t], dtype=np.int64).reshape(-1, 2) + v_off[i]
                                         for i, f in enumerate(feats)], axis=0))
    node_to_graph = np.concatenate([np.asarray(f.node_to_graph, dtype=np.int64) + g_off[i] for i, f in enumerate(feats)])
    if use_numeric_labels:
        labels = np.concatenate([np.asarray(p[1], dtype=np.float64) for p in parts]).astype(np.float32)
    else:  # True -> +1, False -> -1  (adaptive_dkt.py:207-209)
        labels = ((np.concatenate([np.asarray(p[1]) for p in parts]).astype(np.float32) - 0.5) * 2.0).astype(np.float32)
    bool_labels = np.concatenate([np.asarray(x, dtype=bool) for b in batches
                                  for x in (b.support_labels, b.query_labels)])
    ends = np.concatenate([a.reshape(-1) for a in adjacency]) if adjacency else np.zeros(0, dtype=np.int64)
    max_degree = int(np.bincount(ends).max()) if ends.size else 0      # every bond is a message in both directions
    return HostTaskBatch(
        max_degree=max_degree,
        num_tasks=len(batches),
        n_sup=[int(b.support_features.num_graphs) for b in batches],
        n_query=[int(b.query_features.num_graphs) for b in batches],
        num_graphs=int(g_off[-1]), num_nodes=int(v_off[-1]),
        node_features=_pin(np.concatenate([f.node_features for f in feats]).astype(np.float32, copy=False), pin),
        adjacency=[_pin(a, pin) for a in adjacency],
        node_to_graph=_pin(node_to_graph, pin),
        fingerprints=_pin(np.concatenate([f.fingerprints for f in feats]).astype(np.int32, copy=False), pin),
        descriptors=_pin(np.concatenate([f.descriptors for f in feats]).astype(np.float32, copy=False), pin),
        labels=_pin(labels, pin),
        bool_labels=torch.from_numpy(bool_labels))


def to_device(hb: HostTaskBatch, device, build_index: bool = True) -> DeviceTaskBatch:
    """Async H2D of every field on the current stream + device-side CSR build (no host sync)."""
    cp = lambda t: t.to(device, non_blocking=True)  # noqa: E731
    adjs = [cp(a) for a in hb.adjacency]
    n2g = cp(hb.node_to_graph)
    db = DeviceTaskBatch(num_tasks=hb.num_tasks, num_graphs=hb.num_graphs, num_nodes=hb.num_nodes,
                         node_features=cp(hb.node_features), adjacency_lists=adjs, node_to_graph=n2g,
                         fingerprints=cp(hb.fingerprints), descriptors=cp(hb.descriptors), labels=cp(hb.labels),
                         layout=ops.TaskLayout.build(hb.n_sup, hb.n_query, device))
    if build_index:
        db.graph_index = ops.build_graph_index(adjs, hb.num_nodes, n2g, hb.num_graphs,
                                               max_degree=hb.max_degree if hb.max_degree >= 0 else None)
    return db


class DevicePrefetcher:
    """Double-buffered staging of collated task batches: the H2D copies and the device-side CSR build of batch i+1 run on a
    side stream while the compute stream works on batch i (the reference copies field by field, synchronously, inside
    `torchify`, fs_mol/utils/torch_utils.py:7-28).  `next()` makes the compute stream wait for the copy event only.

        pf = DevicePrefetcher(iter_of_host_batches, device)
        for tb in pf: model.train_step(tb)
    """

    def __init__(self, host_batches, device, build_index: bool = True):
        self._it = iter(host_batches)
        self._device = torch.device(device)
        self._build_index = build_index
        self._stream = torch.cuda.Stream(device=self._device)
        self._ready = None
        self._issue()

    def _issue(self):
        try:
            hb = next(self._it)
        except StopIteration:
            self._ready = None
            return
        self._stream.wait_stream(torch.cuda.current_stream(self._device))   # buffers freed by the compute stream are safe to reuse
        with torch.cuda.stream(self._stream):
            tb = to_device(hb, self._device, build_index=self._build_index)
            ev = torch.cuda.Event()
            ev.record(self._stream)
        self._ready = (tb, ev)

    def __iter__(self):
        return self

    def __next__(self) -> DeviceTaskBatch:
        if self._ready is None:
            raise StopIteration
        tb, ev = self._ready
        cur = torch.cuda.current_stream(self._device)
        cur.wait_event(ev)
        for t in _tensors_of(tb):
            t.record_stream(cur)             # allocated on the side stream, consumed on the compute stream
        self._issue()                        # start staging the next batch behind the work about to be queued
        return tb


def _tensors_of(obj):
    import dataclasses
    if torch.is_tensor(obj):
        if obj.is_cuda:
            yield obj
    elif isinstance(obj, (list, tuple)):
        for o in obj:
            yield from _tensors_of(o)
    elif dataclasses.is_dataclass(obj) and not isinstance(obj, type):
        for f in dataclasses.fields(obj):
            yield from _tensors_of(getattr(obj, f.name))
