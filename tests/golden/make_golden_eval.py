"""Generates tests/golden/eval_harness.npz: the reference's OWN evaluation harness (fs_mol/utils/test_utils.py:236-350
`eval_model`, :165-233 `write_csv_summary*`, fs_mol/data/fsmol_task_sampler.py:299-440 `StratifiedTaskSampler`,
fs_mol/utils/metrics.py), run unmodified through oracle/shims.py on seeded synthetic tasks with a deterministic stand-in
model.  Stored per (task, support size, seed): the SMILES of the support and query molecules IN ORDER, and per task the
exact bytes of the result CSV -- classification and --use-numeric-labels.  tests/test_host_logic.py asserts that
adkf_ift_b200.eval_utils / metrics produce the same draws and the same bytes.

    python tests/golden/make_golden_eval.py        (authoring container only)
"""
import os
import sys
import tempfile
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_runner as rr  # noqa: E402

REPORT = rr.install()

from adkf_ift_b200 import synthetic  # noqa: E402

TASK_IDS = (7001, 7002, 7003)
MOLS = 90
SIZES = [16, 32, 64]            # 64 leaves 26 query molecules; all three can be drawn
NUM_SAMPLES = 3
SEED = 11
NUMERIC_SCALE = 40.0            # pushes the numeric labels out of the 0..100 "percentage" filter (test_utils.py:281-290)


def fake_score(smiles: str) -> float:
    """Deterministic pseudo-prediction in (0, 1) from the SMILES string: the 'model' both harnesses evaluate."""
    return ((zlib.crc32(smiles.encode()) & 0xFFFFFF) + 0.5) / float(1 << 24)


def reference_tasks(scale):
    import dataclasses
    from fs_mol.data.fsmol_task import FSMolTask, GraphData, MoleculeDatapoint
    tasks = []
    for tid in TASK_IDS:
        mols = synthetic.make_task_molecules(tid, MOLS)
        conv = []
        for m in mols:
            g = GraphData(node_features=m.graph.node_features, adjacency_lists=list(m.graph.adjacency_lists),
                          edge_features=list(m.graph.edge_features))
            conv.append(MoleculeDatapoint(task_name=m.task_name, smiles=m.smiles, graph=g, numeric_label=m.numeric_label * scale,
                                          bool_label=m.bool_label, fingerprint=m.fingerprint, descriptors=m.descriptors))
        tasks.append(FSMolTask(name=mols[0].task_name, samples=conv))
    return tasks


class RefDataset:
    def __init__(self, tasks):
        self._tasks = tasks

    def get_task_reading_iterable(self, fold, **kw):
        return iter(self._tasks)


def run(numeric: bool):
    from fs_mol.data.fsmol_dataset import DataFold
    from fs_mol.utils.metrics import compute_binary_task_metrics, compute_numeric_task_metrics
    from fs_mol.utils.test_utils import eval_model
    draws = {}

    def test_model_fn(task_sample, temp_out_folder, seed):
        key = f"{task_sample.name}|{len(task_sample.train_samples)}|{seed}"
        draws[key + "|support"] = np.array([s.smiles for s in task_sample.train_samples])
        draws[key + "|query"] = np.array([s.smiles for s in task_sample.test_samples])
        p = np.array([fake_score(s.smiles) for s in task_sample.test_samples])
        if numeric:
            y = np.array([float(s.numeric_label) for s in task_sample.test_samples])
            return compute_numeric_task_metrics(predictions=np.log(y) * (0.5 + p), labels=np.log(y))
        y = np.array([bool(s.bool_label) for s in task_sample.test_samples])
        return compute_binary_task_metrics(predictions=p, labels=y)

    out = {}
    with tempfile.TemporaryDirectory() as d:
        res = eval_model(test_model_fn, RefDataset(reference_tasks(NUMERIC_SCALE if numeric else 1.0)), SIZES, out_dir=d,
                         num_samples=NUM_SAMPLES, test_size_or_ratio=None, fold=DataFold.TEST, seed=SEED,
                         filter_numeric_labels=numeric)
        for name in res:
            with open(os.path.join(d, f"{name}_eval_results.csv"), "rb") as f:
                out[f"csv|{name}"] = np.frombuffer(f.read(), dtype=np.uint8)
        out["task_names"] = np.array(list(res.keys()))
        out["num_results"] = np.array([len(v) for v in res.values()])
    out.update(draws)
    return out


if __name__ == "__main__":
    print(REPORT)
    out = {}
    for numeric in (False, True):
        tag = "numeric" if numeric else "binary"
        r = run(numeric)
        print(tag, "tasks", list(r["task_names"]), "results per task", list(r["num_results"]))
        out.update({f"{tag}::{k}": v for k, v in r.items()})
    out["meta"] = np.array([MOLS, NUM_SAMPLES, SEED])
    out["sizes"] = np.array(SIZES)
    out["task_ids"] = np.array(TASK_IDS)
    out["numeric_scale"] = np.array(NUMERIC_SCALE)
    path = os.path.join(HERE, "eval_harness.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")
