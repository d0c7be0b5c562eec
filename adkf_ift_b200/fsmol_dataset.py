"""FS-Mol on-disk format (the step before the hot path): one `<task>.jsonl.gz` per task under
`train/ valid/ test/`, one JSON object per molecule, optional task-list JSON selecting the assays of every fold
(fs_mol/data/fsmol_task.py:100-144, fs_mol/data/fsmol_dataset.py:30-142, file_reader_iterable.py).  Plain gzip + json:
no dpu_utils / rdkit (fingerprints and descriptors must be in the files, as the released FS-Mol data has them).

Reading is prefetched by a small thread pool so that parsing overlaps the GPU step; the training fold is visited in a fresh
random order every epoch, validation and test folds in sorted order (as the reference)."""
import glob
import gzip
import json
import os
from concurrent.futures import ThreadPoolExecutor
from typing import Dict, Iterable, Iterator, List, Optional, Sequence

import numpy as np

from .eval_utils import DataFold, FSMolTask
from .fsmol_batcher import GraphData, MoleculeDatapoint

_FOLD_DIR = {DataFold.TRAIN: "train", DataFold.VALIDATION: "valid", DataFold.TEST: "test"}


def get_task_name_from_path(path: str) -> str:
    name = os.path.basename(str(path))
    return name[: -len(".jsonl.gz")] if name.endswith(".jsonl.gz") else name


def _datapoint_from_json(raw: Dict, task_name: str) -> MoleculeDatapoint:
    """Field handling of FSMolTask.load_from_file (fsmol_task.py:108-143): int32 count fingerprints, float32 descriptors,
    empty adjacency lists become [0,2] int64, bool label = bool(float(Property)), numeric label = RegressionProperty or nan."""
    g = raw.get("graph")
    fp, desc = raw.get("fingerprints"), raw.get("descriptors")
    adjacency_lists = [np.array(a, dtype=np.int64) if len(a) > 0 else np.zeros((0, 2), dtype=np.int64) for a in g["adjacency_lists"]]
    return MoleculeDatapoint(
        task_name=task_name, smiles=raw["SMILES"], bool_label=bool(float(raw["Property"])),
        numeric_label=float(raw.get("RegressionProperty") or "nan"),
        fingerprint=np.array(fp, dtype=np.int32) if fp is not None else None,
        descriptors=np.array(desc, dtype=np.float32) if desc is not None else None,
        graph=GraphData(node_features=np.array(g["node_features"], dtype=np.float32), adjacency_lists=adjacency_lists,
                        edge_features=[np.array(e, dtype=np.float32) for e in (g.get("edge_features") or [])]))


def load_task_from_file(path: str) -> FSMolTask:
    name = get_task_name_from_path(path)
    samples: List[MoleculeDatapoint] = []
    with gzip.open(path, "rt", encoding="utf-8") as f:
        for line in f:
            if line.strip():
                samples.append(_datapoint_from_json(json.loads(line), name))
    return FSMolTask(name, samples)


def write_task_to_file(task: FSMolTask, path: str) -> None:
    """Inverse of load_task_from_file (used to export synthetic tasks and by the tests)."""
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with gzip.open(path, "wt", encoding="utf-8") as f:
        for m in task.samples:
            obj = {"SMILES": m.smiles, "Property": "1.0" if m.bool_label else "0.0", "Assay_ID": task.name,
                   "RegressionProperty": repr(float(m.numeric_label)),
                   "fingerprints": None if m.fingerprint is None else np.asarray(m.fingerprint).astype(int).tolist(),
                   "descriptors": None if m.descriptors is None else [float(x) for x in np.asarray(m.descriptors, dtype=np.float32)],
                   "graph": {"adjacency_lists": [np.asarray(a).astype(int).tolist() for a in m.graph.adjacency_lists],
                             "node_features": [[float(x) for x in row] for row in np.asarray(m.graph.node_features, dtype=np.float32)],
                             "edge_features": [np.asarray(e, dtype=np.float32).tolist() for e in m.graph.edge_features]}}
            f.write(json.dumps(obj) + "\n")


class FSMolDataset:
    """fs_mol/data/fsmol_dataset.py:39-142 over plain path strings."""

    def __init__(self, train_data_paths: Sequence[str] = (), valid_data_paths: Sequence[str] = (),
                 test_data_paths: Sequence[str] = (), num_workers: Optional[int] = None, seed: Optional[int] = None):
        self._fold_to_data_paths: Dict[DataFold, List[str]] = {DataFold.TRAIN: list(train_data_paths),
                                                               DataFold.VALIDATION: list(valid_data_paths),
                                                               DataFold.TEST: list(test_data_paths)}
        self._num_workers = num_workers if num_workers is not None else min(8, os.cpu_count() or 1)
        self._rng = np.random.default_rng(seed)

    def get_num_fold_tasks(self, fold: DataFold) -> int:
        return len(self._fold_to_data_paths[DataFold(fold)])

    def get_task_names(self, data_fold: DataFold) -> List[str]:
        return [get_task_name_from_path(p) for p in self._fold_to_data_paths[DataFold(data_fold)]]

    @staticmethod
    def from_directory(directory: str, task_list_file: Optional[str] = None, **kwargs) -> "FSMolDataset":
        task_list = None
        if task_list_file is not None:
            with open(task_list_file) as f:
                task_list = json.load(f)

        def fold_files(fold_name: str) -> List[str]:
            files = glob.glob(os.path.join(directory, fold_name, "*.jsonl.gz"))
            if task_list is not None:
                wanted = {f"{t}.jsonl.gz" for t in task_list[fold_name]}
                files = [p for p in files if os.path.basename(p) in wanted]
            return files

        return FSMolDataset(train_data_paths=fold_files("train"), valid_data_paths=sorted(fold_files("valid")),
                            test_data_paths=sorted(fold_files("test")), **kwargs)

    def get_task_reading_iterable(self, data_fold: DataFold, repeat: bool = False, shard: int = 0,
                                  num_shards: int = 1) -> Iterable[FSMolTask]:
        """`shard` / `num_shards`: every num_shards-th task of the (shuffled, repeating) stream starting at `shard`; the other
        files are not read.  The shuffle comes from the dataset's own generator, whose state travels with the pickled dataset,
        so the worker processes of host_pipeline agree on the order (fs_mol/data/fsmol_dataset.py:103-142 shards the FILE
        LIST over its reader workers instead)."""
        return _TaskIterable(self, DataFold(data_fold), repeat, shard, num_shards)


class _TaskIterable:
    def __init__(self, ds: FSMolDataset, fold: DataFold, repeat: bool, shard: int = 0, num_shards: int = 1):
        self._ds, self._fold, self._repeat, self._shard, self._num_shards = ds, fold, repeat, shard, max(1, num_shards)

    def __iter__(self) -> Iterator[FSMolTask]:
        ds = self._ds
        paths = list(ds._fold_to_data_paths[self._fold])
        if not paths:
            return
        workers = max(1, ds._num_workers)
        rng = np.random.Generator(type(ds._rng.bit_generator)())
        rng.bit_generator.state = ds._rng.bit_generator.state      # private copy: iterating does not disturb other iterators
        g = 0
        with ThreadPoolExecutor(max_workers=workers) as pool:
            while True:
                order = list(paths)
                if self._fold == DataFold.TRAIN:
                    rng.shuffle(order)
                mine = []
                for p in order:
                    if g % self._num_shards == self._shard:
                        mine.append(p)
                    g += 1
                pending = []
                for p in mine:                                # bounded prefetch: at most `workers` parsed tasks waiting
                    pending.append(pool.submit(load_task_from_file, p))
                    if len(pending) > workers:
                        yield pending.pop(0).result()
                for fut in pending:
                    yield fut.result()
                if not self._repeat:
                    return
