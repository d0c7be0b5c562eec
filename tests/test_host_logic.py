"""Host-side logic on CPU: graph collation (bit-exact vs the reference's own batcher output), support/query DKT
batching incl. regression label standardisation, multi-task collation, model surface / state-dict names,
checkpoint loading."""
import os

import numpy as np
import pytest
import torch

from adkf_ift_b200 import dkt, synthetic, task_batch
from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer
from adkf_ift_b200.feature_extractor import GNNConfig, GraphFeatureExtractorConfig, cli_default_config


def test_batcher_bit_exact_vs_reference(golden_dir):
    """fsmol_batcher.py:174-218 + dkt.py:84-138 run unmodified in the authoring container -> batcher.npz."""
    g = np.load(os.path.join(golden_dir, "batcher.npz"))
    task_id, n_s, n_q, max_graphs = (int(x) for x in g["meta"])
    ts = synthetic.make_task_sample(task_id, n_s, n_q)
    host = dkt.task_sample_to_dkt_task_sample(ts, dkt.get_dkt_batcher(max_num_graphs=max_graphs), True)
    assert len(host.batches) == int(g["num_batches"]) == 2     # query split into two batches
    for bi, b in enumerate(host.batches):
        for side in ("support", "query"):
            f = getattr(b, side + "_features")
            for t in range(3):
                a = f.adjacency_lists[t]
                assert a.dtype == np.int64 and np.array_equal(a, g[f"b{bi}_{side}_adj{t}"])
                assert np.array_equal(np.concatenate([a, a[:, ::-1]], 0), g[f"b{bi}_{side}_bidir{t}"])
            assert np.array_equal(f.node_to_graph, g[f"b{bi}_{side}_node_to_graph"])
            assert [f.num_graphs, f.num_nodes, f.num_edges] == list(g[f"b{bi}_{side}_counts"])
            assert np.array_equal(f.node_features, g[f"b{bi}_{side}_node_features"])
            assert np.array_equal(f.fingerprints, g[f"b{bi}_{side}_fingerprints"])
            assert np.array_equal(getattr(b, side + "_labels"), g[f"b{bi}_{side}_labels"])
            # regression labels: log(y) standardised with the SUPPORT mean / population std (dkt.py:91-95,110-113)
            assert np.array_equal(getattr(b, side + "_numeric_labels"), g[f"b{bi}_{side}_numeric_labels"])


def test_trailing_single_graph_batch_is_dropped():
    ts = synthetic.make_task_sample(1, 4, 7)       # 7 query graphs, budget 3 per batch -> 3 + 3 + (1 dropped)
    host = dkt.task_sample_to_dkt_task_sample(ts, dkt.get_dkt_batcher(max_num_graphs=7), False)
    assert [b.query_features.num_graphs for b in host.batches] == [3, 3]
    with pytest.raises(ValueError):
        dkt.task_sample_to_dkt_task_sample(ts, dkt.get_dkt_batcher(max_num_graphs=3), False)  # support does not fit


def test_multi_task_collation_layout():
    sizes = [(5, 9), (8, 3), (4, 4)]
    hosts = [dkt.task_sample_to_dkt_task_sample(synthetic.make_task_sample(10 + i, a, b),
                                                dkt.get_dkt_batcher(max_num_graphs=64), False).batches[0]
             for i, (a, b) in enumerate(sizes)]
    hb = task_batch.collate_tasks(hosts, False, pin=False)
    assert hb.n_sup == [5, 8, 4] and hb.n_query == [9, 3, 4] and hb.num_graphs == 33
    n2g = hb.node_to_graph.numpy()
    assert np.all(np.diff(n2g) >= 0) and n2g[-1] == 32 and len(n2g) == hb.num_nodes
    # graph r of the merged batch == the r-th graph in (task0 support, task0 query, task1 support, ...) order
    g_off = v_off = 0
    fps = []
    for h in hosts:
        for f in (h.support_features, h.query_features):
            for t in range(3):
                rows = (hb.adjacency[t].numpy() >= v_off).all(1) & (hb.adjacency[t].numpy() < v_off + f.num_nodes).all(1)
                assert np.array_equal(hb.adjacency[t].numpy()[rows] - v_off, f.adjacency_lists[t])
            assert np.array_equal(n2g[v_off:v_off + f.num_nodes] - g_off, f.node_to_graph)
            fps.append(f.fingerprints)
            g_off += f.num_graphs
            v_off += f.num_nodes
    assert np.array_equal(hb.fingerprints.numpy(), np.concatenate(fps))
    want = np.concatenate([np.concatenate([h.support_labels, h.query_labels]) for h in hosts])
    assert np.array_equal(hb.labels.numpy(), (want.astype(np.float32) - 0.5) * 2)
    assert hb.fingerprints.dtype == torch.int32 and hb.adjacency[0].dtype == torch.int64
    # edges never cross graphs
    for t in range(3):
        a = hb.adjacency[t].numpy()
        assert np.array_equal(n2g[a[:, 0]], n2g[a[:, 1]])


def test_model_surface_and_state_dict_names(golden_dir):
    g = np.load(os.path.join(golden_dir, "c1_classification.npz"))
    model = ADKTModelTrainer(ADKTModelTrainerConfig())
    fe = [n for n, _ in model.named_parameters() if not n.startswith("gp_")]
    assert fe == [str(x) for x in g["grad_names"]]          # reference registration order, 375 tensors
    assert sum(p.numel() for p in model.feature_extractor_params()) == 25_336_354
    assert [n for n, _ in model.named_parameters() if n.startswith("gp_")] == [
        "gp_likelihood.noise_covar.raw_noise", "gp_model.covar_module.raw_outputscale",
        "gp_model.covar_module.base_kernel.raw_lengthscale"]
    assert [tuple(p.shape) for p in model.gp_params()] == [(1,), (), (1, 1)]
    sd = model.state_dict()
    assert sd["graph_feature_extractor.gnn.gnn_blocks.0.mp_layers.0.message_fns.0._layers.0.weight"].shape == (192, 64)
    assert sd["graph_feature_extractor.gnn.gnn_blocks.9.mp_norm_layer.weight"].shape == (128,)   # unused, kept
    assert sd["fc.0.weight"].shape == (2048, 2560) and sd["fc.2.weight"].shape == (2048, 2048)
    assert abs(float(model.gp_likelihood.noise) - 0.1) < 1e-6                       # gp_utils.py:16-17
    assert abs(float(model.gp_model.covar_module.outputscale) - np.log(2.0)) < 1e-6  # softplus(0)
    for attr in ("config", "use_fc", "fc", "fc_out_dim", "gp_likelihood", "gp_model", "mll", "normalizing_features"):
        assert hasattr(model, attr)


def test_unsupported_configurations_fail_loudly():
    for bad in ("tanimoto", "spectral"):        # gp_utils.py:44-45: ValueError for kernels ExactGPLayer does not know
        with pytest.raises(ValueError):
            ADKTModelTrainer(ADKTModelTrainerConfig(gp_kernel=bad))
    assert ADKTModelTrainer(ADKTModelTrainerConfig(gp_kernel="cossim", graph_feature_extractor_config=cli_default_config(1))
                            ).normalizing_features
    # --use-ard: one raw lengthscale per GP input dimension (adaptive_dkt.py:107-110, gp_utils.py:27,30)
    ard = ADKTModelTrainer(ADKTModelTrainerConfig(use_ard=True, graph_feature_extractor_config=cli_default_config(1)))
    assert [tuple(p.shape) for p in ard.gp_params()] == [(1,), (), (1, 2048)] and ard.use_ard
    with pytest.raises(AttributeError):               # like the reference: fc_out_dim only exists with '+fc'
        ADKTModelTrainer(ADKTModelTrainerConfig(use_ard=True, used_features="gnn+ecfp",
                                                graph_feature_extractor_config=cli_default_config(1)))
    # every GNN type of the reference is built (MultiHeadAttention included); unknown names raise like gnn.py:447
    ADKTModelTrainer(ADKTModelTrainerConfig(graph_feature_extractor_config=GraphFeatureExtractorConfig(
        gnn_config=GNNConfig(type="MultiHeadAttention", per_head_dim=64, intermediate_dim=1024, num_layers=2))))
    with pytest.raises(ValueError):
        ADKTModelTrainer(ADKTModelTrainerConfig(graph_feature_extractor_config=GraphFeatureExtractorConfig(
            gnn_config=GNNConfig(type="NoSuchMP", num_layers=2))))


def test_checkpoint_roundtrip_and_reference_gp_keys(tmp_path):
    cfg = ADKTModelTrainerConfig(graph_feature_extractor_config=cli_default_config(1))
    a = ADKTModelTrainer(cfg)
    synthetic.deterministic_init_(a, seed=3)
    path = str(tmp_path / "best_validation.pt")
    a.save_model(path)
    # a reference checkpoint additionally carries gpytorch's own entries: tolerated and skipped
    ck = torch.load(path, weights_only=False)
    ck["model_state_dict"]["mll.likelihood.noise_covar.raw_noise"] = torch.zeros(1)
    ck["model_state_dict"]["gp_model.likelihood.noise_covar.noise_prior.loc"] = torch.zeros(())
    torch.save(ck, path)
    b = ADKTModelTrainer.build_from_model_file(path)
    for (n1, p1), (n2, p2) in zip(a.named_parameters(), b.named_parameters()):
        assert n1 == n2 and torch.equal(p1, p2)
    ck["model_state_dict"]["fc.9.weight"] = torch.zeros(3)
    torch.save(ck, path)
    with pytest.raises(KeyError):
        ADKTModelTrainer.build_from_model_file(path)


def test_synthetic_tasks_are_deterministic_and_well_formed():
    a, b = synthetic.make_task_sample(7, 16, 32), synthetic.make_task_sample(7, 16, 32)
    assert all(np.array_equal(x.fingerprint, y.fingerprint) for x, y in zip(a.train_samples, b.train_samples))
    labels = [m.bool_label for m in a.train_samples]
    assert any(labels) and not all(labels)
    m = a.test_samples[0]
    assert m.graph.node_features.shape[1] == 32 and m.fingerprint.dtype == np.int32 and m.fingerprint.shape == (2048,)
    assert 8 <= len(m.graph.node_features) <= 60 and len(m.graph.adjacency_lists) == 3


def test_metrics_match_scikit_learn():
    """adkf_ift_b200.metrics restates fs_mol/utils/metrics.py without scikit-learn; scikit-learn is the checker here."""
    import warnings
    from sklearn import metrics as S
    from adkf_ift_b200 import metrics as M
    rng = np.random.default_rng(0)
    for trial in range(120):
        n = int(rng.integers(4, 60))
        y = rng.integers(0, 2, n).astype(float)
        if trial % 9 == 0:
            y[:] = 1.0
        p = np.round(rng.random(n), 1 if trial % 3 == 0 else 6)      # ties in the scores every third trial
        m = M.compute_binary_task_metrics(p, y)
        yh = p >= 0.5
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            ref = dict(acc=S.accuracy_score(y, yh), balanced_acc=S.balanced_accuracy_score(y, yh),
                       f1=S.f1_score(y, yh, zero_division=1), prec=S.precision_score(y, yh, zero_division=1),
                       recall=S.recall_score(y, yh, zero_division=1),
                       roc_auc=0.0 if y.sum() in (0, n) else S.roc_auc_score(y, p),
                       avg_precision=S.average_precision_score(y, p), kappa=S.cohen_kappa_score(y, yh))
        for k, v in ref.items():
            got = getattr(m, k)
            assert np.isclose(got, v, atol=1e-12) or (np.isnan(got) and np.isnan(v)), (trial, k, got, v)
    y, p = rng.normal(size=50), rng.normal(size=50)
    nm = M.compute_numeric_task_metrics(p, y)
    assert np.isclose(nm.mse, S.mean_squared_error(y, p)) and np.isclose(nm.mae, S.mean_absolute_error(y, p))
    assert np.isclose(nm.r2, 1.0 - ((y - p) ** 2).sum() / (y ** 2).sum())       # out-of-sample R2 against a zero baseline


def test_stratified_sampler_and_eval_draws(tmp_path):
    """fsmol_task_sampler.py:299-440 / test_utils.py:236-350: split sizes, stratification, determinism in the seed, skipped
    draws, the numeric-label task filter and the CSV column contract."""
    import csv
    from adkf_ift_b200 import eval_utils as E, synthetic
    from adkf_ift_b200.metrics import BinaryEvalMetrics
    ds = synthetic.SyntheticFSMolDataset(num_train_tasks=2, num_valid_tasks=1, num_test_tasks=2, molecules_per_task=80)
    task = next(iter(ds.get_task_reading_iterable(E.DataFold.TEST)))
    s1 = E.StratifiedTaskSampler(16, 0.0, None).sample(task, seed=3)
    s2 = E.StratifiedTaskSampler(16, 0.0, None).sample(task, seed=3)
    s3 = E.StratifiedTaskSampler(16, 0.0, None).sample(task, seed=4)
    assert len(s1.train_samples) == 16 and len(s1.test_samples) == 64
    assert [m.smiles for m in s1.train_samples] == [m.smiles for m in s2.train_samples]
    assert [m.smiles for m in s1.train_samples] != [m.smiles for m in s3.train_samples]
    assert not {m.smiles for m in s1.train_samples} & {m.smiles for m in s1.test_samples}
    frac = np.mean([m.bool_label for m in task.samples])
    assert abs(np.mean([m.bool_label for m in s1.train_samples]) - frac) <= 1.0 / 16 + 1e-9      # stratified
    assert len(E.StratifiedTaskSampler(16, 0.0, 32).sample(task, seed=0).test_samples) == 32
    assert len(E.StratifiedTaskSampler(16, 0.0, (8, 256)).sample(task, seed=0).test_samples) == 64
    with pytest.raises(E.DatasetTooSmallException):
        E.StratifiedTaskSampler(79, 0.0, None).sample(task, seed=0)
    draws = list(E.iter_eval_draws(ds, [16, 79, 32], num_samples=2, seed=7))
    assert len(draws) == 2 and [(d.train_size, d.seed) for d in draws[0][1]] == [(16, 7), (16, 8), (32, 7), (32, 8)]
    # numeric filter: the synthetic labels lie in (0, 100) -> percentage-like -> filtered; scaled by 50 they are kept
    assert list(E.iter_eval_draws(ds, [16], 1, filter_numeric_labels=True)) == []
    ds50 = synthetic.SyntheticFSMolDataset(2, 1, 2, 80, numeric_scale=50.0)
    assert len(list(E.iter_eval_draws(ds50, [16], 1, filter_numeric_labels=True))) == 2

    def fake_model(sample, tmp, seed):
        return BinaryEvalMetrics(size=len(sample.test_samples), acc=0.5, balanced_acc=0.5, f1=0.5, prec=0.5, recall=0.5,
                                 roc_auc=0.5, avg_precision=0.6, kappa=0.0)
    res = E.eval_model(fake_model, ds, [16, 32], out_dir=str(tmp_path), num_samples=2)
    assert sorted(res) == sorted(t.name for t in ds.get_task_reading_iterable(E.DataFold.TEST))
    name = sorted(res)[0]
    assert [r.num_train for r in res[name]] == [16, 16, 32, 32] and res[name][0].num_test == 64
    rows = list(csv.DictReader(open(tmp_path / f"{name}_eval_results.csv")))
    assert list(rows[0].keys())[:6] == ["num_train_requested", "num_train", "fraction_positive_train", "num_test",
                                        "fraction_positive_test", "seed"]
    assert abs(float(rows[0]["delta_auprc"]) - (0.6 - res[name][0].fraction_pos_test)) < 1e-12


def test_fsmol_on_disk_format_round_trip(tmp_path):
    """`<task>.jsonl.gz` files (fsmol_task.py:108-144) -> FSMolDataset.from_directory (fsmol_dataset.py:66-118) -> the same
    arrays, dtypes and batcher output as the in-memory tasks; fold ordering, task-list subset selection, repeat."""
    import json
    from adkf_ift_b200 import dkt, eval_utils as E, fsmol_dataset as D, synthetic
    src = synthetic.SyntheticFSMolDataset(num_train_tasks=3, num_valid_tasks=2, num_test_tasks=2, molecules_per_task=20)
    names = {}
    for fold, sub in ((E.DataFold.TRAIN, "train"), (E.DataFold.VALIDATION, "valid"), (E.DataFold.TEST, "test")):
        names[sub] = []
        for task in src.get_task_reading_iterable(fold):
            D.write_task_to_file(task, str(tmp_path / sub / f"{task.name}.jsonl.gz"))
            names[sub].append(task.name)
    ds = D.FSMolDataset.from_directory(str(tmp_path), num_workers=2, seed=0)
    assert [ds.get_num_fold_tasks(f) for f in E.DataFold] == [3, 2, 2]
    assert ds.get_task_names(E.DataFold.TEST) == sorted(names["test"])
    got = {t.name: t for t in ds.get_task_reading_iterable(E.DataFold.TEST)}
    for want in src.get_task_reading_iterable(E.DataFold.TEST):
        t = got[want.name]
        assert len(t.samples) == len(want.samples)
        for a, b in zip(t.samples, want.samples):
            assert (a.smiles, a.bool_label, a.task_name) == (b.smiles, b.bool_label, b.task_name)
            assert a.numeric_label == b.numeric_label
            assert a.fingerprint.dtype == np.int32 and np.array_equal(a.fingerprint, b.fingerprint)
            assert a.descriptors.dtype == np.float32 and np.array_equal(a.descriptors, b.descriptors)
            assert a.graph.node_features.dtype == np.float32 and np.array_equal(a.graph.node_features, b.graph.node_features)
            assert all(x.dtype == np.int64 and x.shape[1] == 2 and np.array_equal(x, y)
                       for x, y in zip(a.graph.adjacency_lists, b.graph.adjacency_lists))
        # through the sampler and the DKT batcher: identical integer arrays and features
        s1 = E.StratifiedTaskSampler(8, 0.0, None).sample(t, seed=1)
        s2 = E.StratifiedTaskSampler(8, 0.0, None).sample(want, seed=1)
        b1 = dkt.task_sample_to_dkt_task_sample(s1, dkt.get_dkt_batcher(max_num_graphs=64), False).batches[0]
        b2 = dkt.task_sample_to_dkt_task_sample(s2, dkt.get_dkt_batcher(max_num_graphs=64), False).batches[0]
        for x, y in zip(b1.support_features.adjacency_lists, b2.support_features.adjacency_lists):
            assert np.array_equal(x, y)
        assert np.array_equal(b1.query_features.fingerprints, b2.query_features.fingerprints)
        assert np.array_equal(b1.support_features.node_features, b2.support_features.node_features)
    # training fold: every task once per epoch, repeat keeps going
    it = iter(ds.get_task_reading_iterable(E.DataFold.TRAIN, repeat=True))
    seen = [next(it).name for _ in range(6)]
    assert sorted(seen[:3]) == sorted(names["train"]) and sorted(seen[3:]) == sorted(names["train"])
    # task-list subset (datasets/fsmol-0.1.json style)
    with open(tmp_path / "tasks.json", "w") as f:
        json.dump({"train": names["train"][:1], "valid": [], "test": names["test"][1:]}, f)
    sub = D.FSMolDataset.from_directory(str(tmp_path), task_list_file=str(tmp_path / "tasks.json"))
    assert sub.get_task_names(E.DataFold.TRAIN) == names["train"][:1] and sub.get_num_fold_tasks(E.DataFold.VALIDATION) == 0
    assert E.FSMolTask.load_from_file(tmp_path / "test" / f"{names['test'][0]}.jsonl.gz").name == names["test"][0]
    # an empty adjacency list stays a [0,2] int64 array
    empty = [m for t in got.values() for m in t.samples for a in m.graph.adjacency_lists if a.shape[0] == 0]
    assert all(a.shape == (0, 2) for m in empty for a in m.graph.adjacency_lists if a.shape[0] == 0)


def test_functional_call_restores_parameters():
    """Semantics of functional_call (fs_mol/utils/_stateless.py:79-119): substitution for the call only, restoration
    afterwards (also after an error), KeyError for unknown names."""
    from adkf_ift_b200.hypergradient import functional_call
    m = torch.nn.Sequential(torch.nn.Linear(3, 2), torch.nn.Linear(2, 1))
    w = torch.randn(2, 3, requires_grad=True)
    before = m[0].weight
    y = functional_call(m, {"0.weight": w}, (torch.ones(4, 3),))
    y.sum().backward()
    assert w.grad is not None and m[0].weight is before
    with pytest.raises(KeyError):
        functional_call(m, {"0.nope": w}, (torch.ones(4, 3),))
    assert m[0].weight is before


@pytest.mark.parametrize("numeric", [False, True])
def test_eval_harness_equals_reference_draws_and_csv_bytes(golden_dir, tmp_path, numeric):
    """adkf_ift_b200.eval_utils.eval_model / StratifiedTaskSampler / write_csv_summary* / metrics against the
    reference's OWN harness (fs_mol/utils/test_utils.py:165-350, fs_mol/data/fsmol_task_sampler.py:299-440,
    fs_mol/utils/metrics.py) run unmodified on the same synthetic tasks with the same stand-in model
    (tests/golden/make_golden_eval.py): identical support / query molecules per (task, support size, seed), identical CSV bytes."""
    import dataclasses
    import zlib
    from adkf_ift_b200 import eval_utils, metrics
    g = np.load(os.path.join(golden_dir, "eval_harness.npz"))
    tag = "numeric" if numeric else "binary"
    mols, num_samples, seed = (int(x) for x in g["meta"])
    scale = float(g["numeric_scale"]) if numeric else 1.0

    class Dataset:
        def get_task_reading_iterable(self, fold, **kw):
            for tid in g["task_ids"]:
                ms = synthetic.make_task_molecules(int(tid), mols)
                ms = [dataclasses.replace(m, numeric_label=m.numeric_label * scale) for m in ms]
                yield eval_utils.FSMolTask(name=ms[0].task_name, samples=ms)

    def fake_score(smiles):
        return ((zlib.crc32(smiles.encode()) & 0xFFFFFF) + 0.5) / float(1 << 24)

    seen = {}

    def test_model_fn(task_sample, temp_out_folder, s):
        key = f"{task_sample.name}|{len(task_sample.train_samples)}|{s}"
        seen[key + "|support"] = [m.smiles for m in task_sample.train_samples]
        seen[key + "|query"] = [m.smiles for m in task_sample.test_samples]
        p = np.array([fake_score(m.smiles) for m in task_sample.test_samples])
        if numeric:
            y = np.array([float(m.numeric_label) for m in task_sample.test_samples])
            return metrics.compute_numeric_task_metrics(predictions=np.log(y) * (0.5 + p), labels=np.log(y))
        y = np.array([bool(m.bool_label) for m in task_sample.test_samples])
        return metrics.compute_binary_task_metrics(predictions=p, labels=y)

    res = eval_utils.eval_model(test_model_fn, Dataset(), [int(x) for x in g["sizes"]], out_dir=str(tmp_path),
                                num_samples=num_samples, test_size_or_ratio=None, fold=eval_utils.DataFold.TEST, seed=seed,
                                filter_numeric_labels=numeric)
    assert list(res.keys()) == [str(x) for x in g[f"{tag}::task_names"]]
    assert [len(v) for v in res.values()] == [int(x) for x in g[f"{tag}::num_results"]]
    draw_keys = [k for k in g.files if k.startswith(tag + "::") and (k.endswith("|support") or k.endswith("|query"))]
    assert len(draw_keys) == len(seen) and len(seen) > 0
    for k in draw_keys:
        assert seen[k[len(tag) + 2:]] == [str(x) for x in g[k]], k
    for name in res:
        with open(os.path.join(str(tmp_path), f"{name}_eval_results.csv"), "rb") as f:
            assert f.read() == g[f"{tag}::csv|{name}"].tobytes(), name


def test_ranking_metrics_bit_identical_to_scikit_learn():
    """roc_auc / average precision follow scikit-learn's order of operations: EXACT equality (the reference's metrics are
    scikit-learn calls, fs_mol/utils/metrics.py:38-58, and the result CSVs are compared byte for byte)."""
    sk = pytest.importorskip("sklearn.metrics")
    from adkf_ift_b200 import metrics
    rng = np.random.default_rng(0)
    for t in range(300):
        n = int(rng.integers(4, 200))
        y = rng.random(n) < rng.uniform(0.2, 0.8)
        if y.all() or (~y).all():
            continue
        p = rng.random(n)
        if t % 3 == 0:
            p = np.round(p, 1)          # ties
        assert metrics.roc_auc_score(y, p) == float(sk.roc_auc_score(y, p))
        assert metrics.average_precision_score(y, p) == float(sk.average_precision_score(y, p))


def test_host_pipeline_workers_equal_inline():
    """host_pipeline.HostBatchPipeline: worker PROCESSES produce exactly the batches, in exactly the order, of the inline
    round-robin over the same shard generators; ranks see disjoint shards of the task stream (CPU: no pinning)."""
    from adkf_ift_b200.eval_utils import DataFold
    from adkf_ift_b200.host_pipeline import HostBatchPipeline, StreamSpec
    ds = synthetic.SyntheticFSMolDataset(10, 0, 0, molecules_per_task=48)
    spec = StreamSpec(dataset=ds, data_fold=int(DataFold.TRAIN), support_size=8, query_size=16, max_num_graphs=64,
                      use_numeric_labels=False, tasks_per_step=2, num_shards=1, repeat=True)
    inline = HostBatchPipeline(spec, num_workers=0, emulate_workers=2)
    procs = HostBatchPipeline(spec, num_workers=2)
    try:
        for _ in range(5):
            a, b = next(inline), next(procs)
            assert a.n_sup == b.n_sup and a.n_query == b.n_query and a.num_nodes == b.num_nodes
            assert torch.equal(a.node_features, b.node_features) and torch.equal(a.fingerprints, b.fingerprints)
            assert all(torch.equal(x, y) for x, y in zip(a.adjacency, b.adjacency)) and torch.equal(a.labels, b.labels)
    finally:
        procs.close()
    # two ranks: disjoint task shards
    r0 = HostBatchPipeline(spec, num_workers=0, emulate_workers=1, rank=0, world=2)
    r1 = HostBatchPipeline(spec, num_workers=0, emulate_workers=1, rank=1, world=2)
    f0, f1 = next(r0).fingerprints, next(r1).fingerprints
    assert f0.shape != f1.shape or not torch.equal(f0, f1)


def test_weight_gradient_orientation_rule():
    """ops.weight_grad_gemm computes dW in the orientation whose padded tile area (256 x 128 pair tiles above 128 rows, 128 x 128
    below) is smaller: the [384, 1024] projection gradient goes transposed, square and already-fitting shapes stay."""
    from adkf_ift_b200 import ops
    assert ops._padded_tile_area(384, 1024) == 512 * 1024 and ops._padded_tile_area(1024, 384) == 1024 * 384
    assert ops._padded_tile_area(128, 1024) == ops._padded_tile_area(1024, 128) == 128 * 1024
    assert ops._padded_tile_area(768, 1408) == 768 * 1408 and ops._padded_tile_area(1408, 768) == 1536 * 768
    e0 = ops._WEIGHT_EPOCH[0]
    ops.weights_changed()
    assert ops._WEIGHT_EPOCH[0] == e0 + 1
