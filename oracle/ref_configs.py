"""oracle/ref_configs.py -- ORACLE / TEST INFRASTRUCTURE ONLY (bench.py's cpu_baseline legs).

Times the reference's OWN files (unmodified, /root/reference or its staged copy oracle/_ref; third-party libraries through
oracle/shims.py) on the host cores for the BASELINE configs that are reported beside the headline:

  C1  ADKTModel classification meta-test, one task (support 16, query 32): run_on_batches body (adaptive_dkt_utils.py:87-104)
  C4  MoleculeNet ADKF-IFT, one task of the meta-training step (adkfift_trainer.py:163-213 per-task body: fit + cauchy_hypergradient)
  C5  bayes_opt: create_gp + fit on 64 points, joint predictive NLL over the ~2.2k other pool molecules and the reference's
      one-point-at-a-time expected improvement over the pool (bo_utils.py:342-397, predictive_performance/adkt.py:64-96)

Each function returns seconds of one unit of work; recipes as in tests/golden/make_golden*.py.
"""
import ast
import importlib
import os
import sys
import time
import types
from types import SimpleNamespace

import numpy as np
import torch

from . import ref_runner as rr
from . import shims


def c1_setup():
    from adkf_ift_b200 import synthetic
    rr.install()
    model = rr.build_reference_model(rr.reference_config(num_layers=10, gp_kernel="matern"))

    def run(task_id):
        _, dev = rr.reference_dkt_sample(synthetic.make_task_sample(task_id, 16, 32), False)
        t0 = time.perf_counter()
        rr.reference_meta_test(model, dev)
        return time.perf_counter() - t0

    return run


def c3_setup():
    from adkf_ift_b200 import synthetic
    rr.install()
    model = rr.build_reference_model(rr.reference_config(num_layers=10, gp_kernel="matern", use_numeric_labels=True))

    def run(task_id):
        _, dev = rr.reference_dkt_sample(synthetic.make_task_sample(task_id, 128, 192), True)
        t0 = time.perf_counter()
        rr.reference_meta_test(model, dev)
        return time.perf_counter() - t0

    return run


def c4_setup(n_s=20, n_q=32):
    from adkf_ift_b200 import synthetic
    from . import torch_geometric_lite
    shims.install()
    torch_geometric_lite.install()
    from botorch.optim.fit import fit_gpytorch_scipy
    from fs_mol.utils._stateless import functional_call
    from fs_mol.utils.cauchy_hypergradient import cauchy_hypergradient
    path = os.path.join(shims.REFERENCE_ROOT, "MoleculeNet", "chem_lib", "models")
    pkg = types.ModuleType("ref_chem_models")
    pkg.__path__ = [path]
    sys.modules["ref_chem_models"] = pkg
    mod = importlib.import_module("ref_chem_models.adkf_model")
    args = SimpleNamespace(emb_dim=300, enc_layer=5, JK="last", dropout=0.0, enc_pooling="mean", enc_gnn="gin",
                           enc_batch_norm=1, pretrained=0, gpu_id=0)
    model = mod.ADKFModel(args)
    synthetic.deterministic_init_(model, seed=1)
    fe_names = [n for n, _ in model.named_parameters() if not n.startswith("gp_")]

    def data_of(d):
        return SimpleNamespace(x=torch.from_numpy(d["x"]), edge_index=torch.from_numpy(d["edge_index"]),
                               edge_attr=torch.from_numpy(d["edge_attr"]), batch=torch.from_numpy(d["batch"]),
                               y=torch.from_numpy(d["y"]))

    def run(t):
        s_data, q_data = data_of(synthetic.make_mol_graphs(100 + 2 * t, n_s)), data_of(synthetic.make_mol_graphs(101 + 2 * t, n_q))
        t0 = time.perf_counter()
        model.train()
        _ = model(s_data=s_data, q_data=None, s_label=s_data.y, train_loss=True)
        fit_gpytorch_scipy(model.mll)
        model.train()
        gp_names = [n for n, _ in model.named_parameters() if n.startswith("gp_")]

        def f_inner(po, pi):
            d = {**dict(zip(fe_names, po)), **dict(zip(gp_names, pi))}
            return functional_call(model, d, (s_data, q_data), kwargs={"s_label": s_data.y, "train_loss": True, "is_functional_call": True})

        def f_outer(po, pi):
            d = {**dict(zip(fe_names, po)), **dict(zip(gp_names, pi))}
            return functional_call(model, d, (s_data, q_data), kwargs={"s_label": s_data.y, "train_loss": False,
                                                                       "predictive_val_loss": True, "is_functional_call": True})

        cauchy_hypergradient(f_outer, f_inner, tuple(model.feature_extractor_params()), tuple(model.gp_params()),
                             torch.device("cpu"), ignore_grad_correction=False)
        return time.perf_counter() - t0

    return run


def c5_setup(pool=2312, dim=2048, n_fit=64):
    from adkf_ift_b200 import synthetic
    shims.install()
    import botorch
    import gpytorch
    from botorch.optim.fit import fit_gpytorch_scipy
    from fs_mol.utils.gp_utils import TanimotoKernel
    path = os.path.join(shims.REFERENCE_ROOT, "bayes_opt", "bo_utils.py")
    tree = ast.parse(open(path).read())
    ns = {"torch": torch, "np": np, "gpytorch": gpytorch, "botorch": botorch, "TanimotoKernel": TanimotoKernel,
          "fit_gpytorch_scipy": fit_gpytorch_scipy}
    for node in tree.body:
        if isinstance(node, (ast.FunctionDef, ast.ClassDef)) and node.name in {"create_gp", "CustomKernelGP", "compute_median_lengthscale_init"}:
            exec(compile(ast.Module([node], []), path, "exec"), ns)
    create_gp = ns["create_gp"]
    x_all, y_all = synthetic.bo_pool(seed=5, pool=pool, dim=dim)
    X, Y = torch.from_numpy(x_all), torch.from_numpy(y_all)

    def run(seed, ei_points=None):
        """-> (seconds for fit + joint NLL, seconds for the EI sweep over `ei_points` pool molecules, points swept)."""
        rng = np.random.RandomState(seed)
        tr = np.sort(rng.choice(np.arange(pool), size=n_fit, replace=False))
        te = np.array([i for i in range(pool) if i not in set(tr.tolist())])
        y_tr = (Y[tr] - Y[tr].mean()) / Y[tr].std()
        y_te = (Y[te] - Y[tr].mean()) / Y[tr].std()
        t0 = time.perf_counter()
        likelihood, model, mll = create_gp(X[tr], y_tr, "matern", torch.device("cpu"), 0.01, True)
        model.train(); likelihood.train()
        fit_gpytorch_scipy(mll)
        model.eval(); likelihood.eval()
        with torch.no_grad():
            pred = likelihood(model(X[te]))
            _ = (-pred.log_prob(y_te) / y_te.shape[0]).item()
        t1 = time.perf_counter()
        n_ei = pool if ei_points is None else ei_points
        with torch.no_grad():
            acq = botorch.acquisition.analytic.ExpectedImprovement(model, y_tr.min().item(), maximize=False)
            for j in range(n_ei):
                acq(X[j:j + 1, :])
        t2 = time.perf_counter()
        return t1 - t0, t2 - t1, n_ei

    return run
