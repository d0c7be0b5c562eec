"""Seeded synthetic FS-Mol-shaped tasks.  There is no network / dataset in the
build or bench environment, so every test, the oracle and bench.py draw their inputs from here.

Molecule: V = clip(round(N(26, 7)), 8, 60) atoms; bonds = random spanning tree (atom i -> uniform j < i)
plus floor(V/6) extra random pairs (each bond stored once, as FS-Mol does); bond type in {0,1,2} w.p.
(0.80, 0.19, 0.01); node features [V, 32] float32 = concatenated one-hot blocks; ECFP count fingerprint
[2048] int32 with ~45 non-zeros of value 1..4; 42 phys-chem descriptors ~ N(0, 1).
Each task is a handful of chemical series ("scaffolds") sharing fingerprint bits; the hidden score is
a per-series effect plus noise -> numeric label exp(1 + score) (positive, like FS-Mol activity values)
and bool label = score above the task median, so the features cluster and the GP fit has signal.  `task_rng = np.random.default_rng(1234 + task_id)`.
"""
import math
from typing import List, Tuple

import numpy as np

from .fsmol_batcher import (NUM_EDGE_TYPES, NUM_NODE_FEATURES, FSMolTaskSample, GraphData,
                            MoleculeDatapoint)

FINGERPRINT_DIM = 2048
PHYS_CHEM_DESCRIPTORS_DIM = 42
_ONE_HOT_BLOCKS = (15, 6, 5, 4, 2)  # sums to NUM_NODE_FEATURES
assert sum(_ONE_HOT_BLOCKS) == NUM_NODE_FEATURES


def _random_molecule_graph(rng: np.random.Generator) -> GraphData:
    v = int(np.clip(np.rint(rng.normal(26.0, 7.0)), 8, 60))
    # spanning tree + a few ring-closing bonds
    src = np.arange(1, v, dtype=np.int64)
    tgt = (rng.random(v - 1) * src).astype(np.int64)  # uniform j < i
    n_extra = v // 6
    ex_a = rng.integers(0, v, size=n_extra)
    ex_b = rng.integers(0, v, size=n_extra)
    keep = ex_a != ex_b
    src = np.concatenate([src, ex_a[keep]])
    tgt = np.concatenate([tgt, ex_b[keep]])
    btype = rng.choice(NUM_EDGE_TYPES, size=src.shape[0], p=(0.80, 0.19, 0.01))
    adjacency = [np.stack([src[btype == t], tgt[btype == t]], axis=1).astype(np.int64)
                 for t in range(NUM_EDGE_TYPES)]
    feats = np.zeros((v, NUM_NODE_FEATURES), dtype=np.float32)
    col = 0
    for width in _ONE_HOT_BLOCKS:
        # skewed categorical so that some atom "types" dominate, as in real molecules
        p = 0.5 ** np.arange(1, width + 1)
        p /= p.sum()
        pick = rng.choice(width, size=v, p=p)
        feats[np.arange(v), col + pick] = 1.0
        col += width
    return GraphData(node_features=feats, adjacency_lists=adjacency, edge_features=[])


def _random_fingerprint(rng: np.random.Generator, scaffold: np.ndarray) -> np.ndarray:
    """Scaffold bits (shared by the molecules of one chemical series) + molecule-specific decoration."""
    fp = np.zeros(FINGERPRINT_DIM, dtype=np.int32)
    keep = scaffold[rng.random(scaffold.shape[0]) < 0.9]
    fp[keep] = rng.integers(1, 5, size=keep.shape[0]).astype(np.int32)
    n_own = int(np.clip(np.rint(rng.normal(12, 3)), 3, 40))
    own = rng.choice(FINGERPRINT_DIM, size=n_own, replace=False)
    fp[own] = rng.integers(1, 4, size=n_own).astype(np.int32)
    return fp


def make_task_molecules(task_id: int, num_molecules: int) -> List[MoleculeDatapoint]:
    rng = np.random.default_rng(1234 + task_id)
    name = f"SYNTH{task_id:06d}"
    graphs = [_random_molecule_graph(rng) for _ in range(num_molecules)]
    # a task = a handful of chemical series ("scaffolds"); activity depends on the series + noise
    n_scaffolds = 4
    scaffolds = [rng.choice(FINGERPRINT_DIM, size=int(rng.integers(28, 40)), replace=False)
                 for _ in range(n_scaffolds)]
    effect = rng.normal(size=n_scaffolds)
    effect = (effect - effect.mean()) / (effect.std() + 1e-12)
    member = rng.integers(0, n_scaffolds, size=num_molecules)
    fps = [_random_fingerprint(rng, scaffolds[k]) for k in member]
    descs = rng.normal(size=(num_molecules, PHYS_CHEM_DESCRIPTORS_DIM)).astype(np.float32)
    score = effect[member] + 0.5 * rng.normal(size=num_molecules)
    numeric = np.exp(1.0 + score)
    thresh = np.median(score)
    out = []
    for i in range(num_molecules):
        out.append(MoleculeDatapoint(task_name=name, smiles=f"SYN{task_id}_{i}", graph=graphs[i],
                                     numeric_label=float(numeric[i]), bool_label=bool(score[i] > thresh),
                                     fingerprint=fps[i], descriptors=descs[i]))
    return out


def make_task_sample(task_id: int, support_size: int, query_size: int) -> FSMolTaskSample:
    """One few-shot task: `support_size` train molecules (>=1 of each class) and `query_size` test
    molecules.  Deterministic in (task_id, support_size, query_size)."""
    mols = make_task_molecules(task_id, support_size + query_size)
    support, query = mols[:support_size], mols[support_size:]
    labels = [m.bool_label for m in support]
    if support_size >= 2 and (all(labels) or not any(labels)):
        # force both classes into the support set by swapping with a query molecule
        want = not labels[0]
        for j, q in enumerate(query):
            if q.bool_label == want:
                support[-1], query[j] = query[j], support[-1]
                break
    return FSMolTaskSample(name=mols[0].task_name, train_samples=support, valid_samples=[],
                           test_samples=query)


def make_task_samples(first_task_id: int, num_tasks: int, support_size: int,
                      query_size: int) -> List[FSMolTaskSample]:
    return [make_task_sample(first_task_id + t, support_size, query_size) for t in range(num_tasks)]


def trained_like_(model, alpha: float = 0.5, seed: int = 0) -> None:
    """Make a freshly initialised feature extractor behave like a trained one for numerics tests:
    the ReZero gates `alpha` are 1e-7 at init (fs_mol/modules/gnn.py:405-406), which would make every
    GNN test vacuous; set them to `alpha`."""
    import torch
    with torch.no_grad():
        for name, p in model.named_parameters():
            if name.endswith(".alpha"):
                p.fill_(alpha)


def deterministic_init_(module_or_state, seed: int = 0, alpha: float = 0.5, gain: float = 1.0) -> None:
    """Fill every non-GP tensor of a model (nn.Module or state-dict mapping) with values that depend
    only on (tensor name, shape, seed) -- NOT on module construction order -- so that the reference
    model (authoring container), the CPU oracle and the CUDA product get bit-identical weights and
    the golden fixtures need not store 100 MB of parameters.  Linear weights ~ U(+-gain/sqrt(fan_in)),
    biases ~ U(+-0.05), LayerNorm weight 1 +- 0.1, ReZero alpha = `alpha` (trained-like)."""
    import zlib

    import torch

    state = module_or_state.state_dict() if hasattr(module_or_state, "state_dict") else module_or_state
    with torch.no_grad():
        for name in sorted(state.keys()):
            t = state[name]
            if name.startswith("gp_") or name.startswith("mll.") or not torch.is_floating_point(t):
                continue
            g = torch.Generator(device="cpu")
            g.manual_seed((zlib.crc32(name.encode()) ^ (seed * 2654435761)) & 0x7FFFFFFF)
            u = torch.rand(t.shape, generator=g, dtype=torch.float64) * 2.0 - 1.0
            if name.endswith(".alpha"):
                val = torch.full(t.shape, float(alpha), dtype=torch.float64)
            elif "norm_layer.weight" in name or ("batch_norms" in name and name.endswith(".weight")):
                val = 1.0 + 0.1 * u
            elif "norm_layer.bias" in name or ("batch_norms" in name and name.endswith(".bias")):
                val = 0.1 * u
            elif name.endswith("running_mean"):
                val = torch.zeros(t.shape, dtype=torch.float64)
            elif name.endswith("running_var"):
                val = torch.ones(t.shape, dtype=torch.float64)
            elif "embedding" in name:            # MoleculeNet atom / bond embedding tables
                val = 0.5 * u
            elif t.dim() == 2:
                val = u * (gain / math.sqrt(t.shape[1]))
            else:
                val = 0.05 * u
            t.copy_(val.to(t.dtype))


def bo_pool(seed: int = 5, pool: int = 2312, dim: int = 64, binary: bool = False):
    """dockstring-shaped candidate pool of BASELINE config 5: `pool` molecules with `dim` features and
    docking-score-like targets y ~ N(-8, 1.5) SORTED ascending (bo_utils.py:341 expects that order).  Features carry
    signal about y (a smooth function of three latent coordinates) so that the GP fit is non-degenerate; binary=True
    gives fingerprint-like 0/1 features for the Tanimoto kernel.  Returns (x [pool, dim] f32, y [pool] f32)."""
    rng = np.random.default_rng(seed)
    lat = rng.normal(size=(pool, 3))
    score = lat[:, 0] + 0.5 * np.sin(2.0 * lat[:, 1]) + 0.3 * lat[:, 2] ** 2
    order = np.argsort(score)
    lat = lat[order]
    y = np.sort(rng.normal(-8.0, 1.5, size=pool))
    proj = rng.normal(size=(3, dim)) / np.sqrt(3.0)
    x = np.tanh(lat @ proj) + 0.15 * rng.normal(size=(pool, dim))
    if binary:
        x = (x > 0.6).astype(np.float32)
    return x.astype(np.float32), y.astype(np.float32)


def make_mol_graphs(seed: int, num_graphs: int):
    """MoleculeNet-shaped graphs in the PyG layout of MoleculeNet/chem_lib: atom type U{0..118},
    chirality U{0..2}; random spanning tree + V//6 extra bonds, every bond stored in both directions with the same
    attributes (bond type U{0..3}, direction U{0..2}); labels Bernoulli(0.5) with both classes present.
    Returns dict(x [V,2], edge_index [2,E], edge_attr [E,2], batch [V], y [G]) of int64 numpy arrays."""
    rng = np.random.default_rng(77_000 + seed)
    xs, srcs, tgts, attrs, batch = [], [], [], [], []
    v_off = 0
    for g in range(num_graphs):
        V = int(np.clip(round(rng.normal(24, 6)), 6, 50))
        xs.append(np.stack([rng.integers(0, 119, size=V), rng.integers(0, 3, size=V)], 1))
        bonds = [(i, int(rng.integers(0, i))) for i in range(1, V)]
        have = set(bonds)
        for _ in range(V // 6):
            a, b = (int(v) for v in rng.integers(0, V, size=2))
            if a != b and (a, b) not in have and (b, a) not in have:
                bonds.append((a, b)); have.add((a, b))
        for (a, b) in bonds:
            at = (int(rng.integers(0, 4)), int(rng.integers(0, 3)))
            srcs += [a + v_off, b + v_off]; tgts += [b + v_off, a + v_off]; attrs += [at, at]
        batch += [g] * V
        v_off += V
    y = rng.integers(0, 2, size=num_graphs)
    y[0], y[1 % num_graphs] = 1, 0
    return dict(x=np.concatenate(xs).astype(np.int64), edge_index=np.array([srcs, tgts], dtype=np.int64),
                edge_attr=np.array(attrs, dtype=np.int64).reshape(-1, 2), batch=np.array(batch, dtype=np.int64),
                y=y.astype(np.int64))


class SyntheticFSMolDataset:
    """Stand-in for fs_mol.data.FSMolDataset (fsmol_dataset.py:30-142) over seeded synthetic tasks: the only method the task
    loops use is `get_task_reading_iterable(data_fold)`.  `numeric_scale` > 5 pushes the numeric labels out of the
    percentage range that `--use-numeric-labels` filters (test_utils.py:281-290)."""

    def __init__(self, num_train_tasks: int = 8, num_valid_tasks: int = 2, num_test_tasks: int = 2, molecules_per_task: int = 96,
                 first_task_id: int = 5000, numeric_scale: float = 1.0):
        self._counts = (num_train_tasks, num_valid_tasks, num_test_tasks)
        self._mols, self._first, self._scale = molecules_per_task, first_task_id, numeric_scale
        self._cache = {}

    def materialize(self, data_fold) -> "SyntheticFSMolDataset":
        """Generate and keep every task of a fold in memory (a stand-in for tasks already read from disk): the cache travels
        with the pickled dataset to host_pipeline's worker processes, which then only sample / featurise / collate."""
        fold = int(data_fold)
        start = self._first + sum(self._counts[:fold])
        for i in range(self._counts[fold]):
            self._cache[start + i] = self._task(start + i)
        return self

    def get_num_fold_tasks(self, data_fold) -> int:
        return self._counts[int(data_fold)]

    def _task(self, task_id: int):
        import dataclasses
        from .eval_utils import FSMolTask
        if task_id in self._cache:
            return self._cache[task_id]
        mols = make_task_molecules(task_id, self._mols)
        if self._scale != 1.0:
            mols = [dataclasses.replace(m, numeric_label=m.numeric_label * self._scale) for m in mols]
        return FSMolTask(name=mols[0].task_name, samples=mols)

    def get_task_reading_iterable(self, data_fold, repeat: bool = False, shard: int = 0, num_shards: int = 1):
        """`shard` / `num_shards`: every num_shards-th task of the (repeating) stream starting at `shard`, without generating
        the others (host_pipeline workers each take one shard)."""
        fold = int(data_fold)
        start = self._first + sum(self._counts[:fold])
        g = 0
        while True:
            for i in range(self._counts[fold]):
                if g % num_shards == shard:
                    yield self._task(start + i)
                g += 1
            if not repeat or self._counts[fold] == 0:
                return
