"""Support/query batching for one few-shot task (host side).

Mirrors `fs_mol/data/dkt.py` (MoleculeDKTFeatures :25-28, DKTBatch :31-46, FeaturisedDKTTaskSample :49-58,
batcher callbacks :61-81, task_sample_to_dkt_task_sample :84-138, get_dkt_batcher :141-153) and
`fs_mol/utils/torch_utils.py:7-28` (torchify).  Regression labels: log(y) standardised with the SUPPORT
set's mean and numpy population std (ddof=0), applied to support and query alike (dkt.py:91-95,110-113).
"""
import dataclasses
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import torch

from .fsmol_batcher import (FSMolBatch, FSMolBatcher, FSMolTaskSample, MoleculeDatapoint,
                            fsmol_batch_finalizer)


@dataclass(frozen=True)
class MoleculeDKTFeatures(FSMolBatch):
    fingerprints: np.ndarray  # [G, 2048] int32 counts
    descriptors: np.ndarray  # [G, 42] float32


@dataclass(frozen=True)
class DKTBatch:
    support_features: MoleculeDKTFeatures
    support_labels: np.ndarray
    support_numeric_labels: np.ndarray
    query_features: MoleculeDKTFeatures
    query_labels: np.ndarray
    query_numeric_labels: np.ndarray

    @property
    def num_support_samples(self) -> int:
        return self.support_features.num_graphs

    @property
    def num_query_samples(self) -> int:
        return self.query_features.num_graphs


@dataclass(frozen=True)
class FeaturisedDKTTaskSample:
    task_name: str
    num_support_samples: int
    num_positive_support_samples: int
    num_query_samples: int
    num_positive_query_samples: int
    batches: List[DKTBatch]
    batch_labels: List[np.ndarray]
    batch_numeric_labels: List[np.ndarray]


def batcher_init_fn(batch_data: Dict[str, Any]):
    batch_data["fingerprints"] = []
    batch_data["descriptors"] = []


def batcher_add_sample_fn(batch_data: Dict[str, Any], sample_id: int, sample: MoleculeDatapoint):
    batch_data["fingerprints"].append(sample.get_fingerprint())
    batch_data["descriptors"].append(sample.get_descriptors())


def batcher_finalizer_fn(batch_data: Dict[str, Any]) -> Tuple[MoleculeDKTFeatures, np.ndarray, np.ndarray]:
    plain = fsmol_batch_finalizer(batch_data)
    fields = {f.name: getattr(plain, f.name) for f in dataclasses.fields(plain)}
    feats = MoleculeDKTFeatures(
        fingerprints=np.stack(batch_data["fingerprints"], axis=0),
        descriptors=np.stack(batch_data["descriptors"], axis=0),
        **fields,
    )
    return feats, np.stack(batch_data["bool_labels"], axis=0), np.stack(batch_data["numeric_labels"], axis=0)


def get_dkt_batcher(max_num_graphs: Optional[int] = None, max_num_nodes: Optional[int] = None,
                    max_num_edges: Optional[int] = None) -> FSMolBatcher:
    return FSMolBatcher(max_num_graphs, max_num_nodes, max_num_edges,
                        init_callback=batcher_init_fn,
                        per_datapoint_callback=batcher_add_sample_fn,
                        finalizer_callback=batcher_finalizer_fn)


def task_sample_to_dkt_task_sample(task_sample: FSMolTaskSample, batcher: FSMolBatcher,
                                   filter_numeric_labels: bool) -> FeaturisedDKTTaskSample:
    support_batches = list(batcher.batch(task_sample.train_samples))
    if len(support_batches) > 1:
        raise ValueError("Support set too large to fit into a single batch!")
    support_features, support_labels, support_numeric = support_batches[0]
    if filter_numeric_labels:
        log_y = np.log(support_numeric)
        mu, sd = log_y.mean(), log_y.std()
        support_numeric_std = (log_y - mu) / sd
    else:
        support_numeric_std = support_numeric

    orig_max = batcher._max_num_graphs
    batches: List[DKTBatch] = []
    batch_labels: List[np.ndarray] = []
    batch_numeric: List[np.ndarray] = []
    try:
        # every DKTBatch carries the support set, so the query share is what is left of the budget
        batcher._max_num_graphs = orig_max - support_features.num_graphs
        for q_feats, q_labels, q_numeric in batcher.batch(task_sample.test_samples):
            q_numeric_std = (np.log(q_numeric) - mu) / sd if filter_numeric_labels else q_numeric
            batches.append(DKTBatch(support_features, support_labels, support_numeric_std,
                                    q_feats, q_labels, q_numeric_std))
            batch_labels.append(q_labels)
            batch_numeric.append(q_numeric_std)
    finally:
        batcher._max_num_graphs = orig_max

    return FeaturisedDKTTaskSample(
        task_name=task_sample.name,
        num_support_samples=len(task_sample.train_samples),
        num_positive_support_samples=sum(s.bool_label for s in task_sample.train_samples),
        num_query_samples=len(task_sample.test_samples),
        num_positive_query_samples=sum(s.bool_label for s in task_sample.test_samples),
        batches=batches,
        batch_labels=batch_labels,
        batch_numeric_labels=batch_numeric,
    )


def torchify(data, device: torch.device):
    """Recursive numpy -> torch + `.to(device)` (fs_mol/utils/torch_utils.py:7-28)."""
    if isinstance(data, (int, float, str)) or data is None:
        return data
    if torch.is_tensor(data):
        return data.to(device)
    if isinstance(data, tuple):
        return tuple(torchify(e, device) for e in data)
    if isinstance(data, list):
        return [torchify(e, device) for e in data]
    if isinstance(data, dict):
        return {k: torchify(v, device) for k, v in data.items()}
    if isinstance(data, np.ndarray):
        return torch.from_numpy(data).to(device)
    if isinstance(data, np.generic):
        return data.item()
    if dataclasses.is_dataclass(data):
        new = {f.name: torchify(getattr(data, f.name), device) for f in dataclasses.fields(data)}
        return dataclasses.replace(data, **new)
    raise ValueError(f"Trying to torchify unknown value type {type(data)}!")


def get_dkt_task_sample_iterable(dataset, data_fold, num_samples: int, support_size: int, query_size: Optional[int],
                                 max_num_graphs: Optional[int] = None, max_num_nodes: Optional[int] = None,
                                 max_num_edges: Optional[int] = None, repeat: bool = False,
                                 filter_numeric_labels: bool = False):
    """fs_mol/data/dkt.py:154-207: stratified (support, query) draws of every task of a fold, featurised into DKT batches.
    `dataset` needs `get_task_reading_iterable(data_fold, repeat=...)` yielding tasks; the draw seed is task index + draw
    index as in the reference's reader pipeline."""
    from .eval_utils import StratifiedTaskSampler, numeric_task_is_filtered
    sampler = StratifiedTaskSampler(train_size_or_ratio=support_size, test_size_or_ratio=query_size)
    batcher = get_dkt_batcher(max_num_graphs=max_num_graphs, max_num_nodes=max_num_nodes, max_num_edges=max_num_edges)
    for idx, task in enumerate(dataset.get_task_reading_iterable(data_fold, repeat=repeat)):
        if filter_numeric_labels and numeric_task_is_filtered(task):
            continue
        drawn = 0
        for _ in range(num_samples):
            try:
                sample = sampler.sample(task, seed=idx + drawn)
                drawn += 1
            except Exception:      # noqa: BLE001  (the reference skips any failed draw, dkt.py:191-196)
                continue
            yield task_sample_to_dkt_task_sample(sample, batcher, filter_numeric_labels)
