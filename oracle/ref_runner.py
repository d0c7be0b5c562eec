"""oracle/ref_runner.py -- ORACLE / TEST INFRASTRUCTURE ONLY; authoring container only.

Drives the reference's OWN, UNMODIFIED Python files (under /root/reference, through oracle/shims.py)
on the seeded synthetic tasks, to generate the golden fixtures in tests/golden/ and to pin
oracle/adkf_oracle.py.  Nothing here can run on the GPU box (no /root/reference there) and nothing in
`-m gpu` tests, smoke() or bench.py imports it.
"""
from typing import Dict, List

import numpy as np
import torch

from . import shims


def install():
    return shims.install()


def reference_config(num_layers=10, used_features="gnn+ecfp+fc", gp_kernel="matern",
                     use_numeric_labels=False, use_lengthscale_prior=True, gnn_type="PNA", **extra):
    """CLI-default model (fs_mol/adaptive_dkt_train.py + gnn.py:31-63) as an ADKTModelTrainerConfig."""
    from fs_mol.modules.gnn import GNNConfig
    from fs_mol.modules.graph_feature_extractor import GraphFeatureExtractorConfig
    from fs_mol.modules.graph_readout import GraphReadoutConfig
    from fs_mol.utils.adaptive_dkt_utils import ADKTModelTrainerConfig
    return ADKTModelTrainerConfig(
        graph_feature_extractor_config=GraphFeatureExtractorConfig(
            gnn_config=GNNConfig(type=gnn_type, hidden_dim=128, num_heads=4, per_head_dim=64,
                                 intermediate_dim=1024, num_layers=num_layers),
            readout_config=GraphReadoutConfig()),
        used_features=used_features, gp_kernel=gp_kernel, use_numeric_labels=use_numeric_labels,
        use_lengthscale_prior=use_lengthscale_prior, **extra)


def build_reference_model(config, seed=0, alpha=0.5, dtype=torch.float32):
    from fs_mol.utils.adaptive_dkt_utils import ADKTModelTrainer
    from adkf_ift_b200.synthetic import deterministic_init_
    if dtype == torch.float64:
        torch.set_default_dtype(torch.float64)  # model.double() fails in the readout
    model = ADKTModelTrainer(config)
    deterministic_init_(model, seed=seed, alpha=alpha)
    return model


def to_reference_task_sample(sample):
    """adkf_ift_b200 FSMolTaskSample -> the reference's own dataclasses."""
    from fs_mol.data.fsmol_task import FSMolTaskSample, GraphData, MoleculeDatapoint

    def conv(m):
        g = GraphData(node_features=m.graph.node_features, adjacency_lists=list(m.graph.adjacency_lists),
                      edge_features=list(m.graph.edge_features))
        return MoleculeDatapoint(task_name=m.task_name, smiles=m.smiles, graph=g,
                                 numeric_label=m.numeric_label, bool_label=m.bool_label,
                                 fingerprint=m.fingerprint, descriptors=m.descriptors)

    return FSMolTaskSample(name=sample.name, train_samples=[conv(m) for m in sample.train_samples],
                           valid_samples=[], test_samples=[conv(m) for m in sample.test_samples])


def reference_dkt_sample(sample, use_numeric_labels, max_num_graphs=320, dtype=torch.float32):
    from fs_mol.data.dkt import get_dkt_batcher, task_sample_to_dkt_task_sample
    from fs_mol.utils.torch_utils import torchify
    ref_sample = to_reference_task_sample(sample)
    batcher = get_dkt_batcher(max_num_graphs=max_num_graphs)
    host = task_sample_to_dkt_task_sample(ref_sample, batcher, use_numeric_labels)
    dev = torchify(host, device=torch.device("cpu"))
    if dtype == torch.float64:
        dev = _cast_batches(dev, dtype)
    return host, dev


def _cast_batches(dkt_sample, dtype):
    import dataclasses

    def cast(x):
        if torch.is_tensor(x) and x.dtype == torch.float32:
            return x.to(dtype)
        if isinstance(x, list):
            return [cast(e) for e in x]
        if dataclasses.is_dataclass(x):
            return dataclasses.replace(x, **{f.name: cast(getattr(x, f.name)) for f in dataclasses.fields(x)})
        return x

    return cast(dkt_sample)


def reference_meta_test(model, dkt_sample):
    """run_on_batches(train=False) body, instrumented: per query batch returns the MVN and phi."""
    from botorch.optim.fit import fit_gpytorch_scipy
    outs = []
    for batch in dkt_sample.batches:
        model.train()
        _ = model(batch, train_loss=True)
        phi0 = [p.detach().clone() for p in model.gp_params()]
        fit_gpytorch_scipy(model.mll)
        model.eval()
        with torch.no_grad():
            mvn = model(batch, train_loss=None)
        outs.append(dict(mean=mvn.mean.detach().clone(), variance=mvn.variance.detach().clone(),
                         phi_init=phi0, phi_star=[p.detach().clone() for p in model.gp_params()]))
    return outs


def reference_meta_train_task(model, batch, fit=True, phi=None):
    """One task of train_loop (adaptive_dkt_utils.py:361-404) with the reference's own
    cauchy_hypergradient and functional_call.  Returns F and the per-parameter gradients.
    `phi` (3 raw values, with fit=False) replaces the inner fit by given GP parameters (fixed-phi goldens)."""
    from botorch.optim.fit import fit_gpytorch_scipy
    from fs_mol.utils._stateless import functional_call
    from fs_mol.utils.cauchy_hypergradient import cauchy_hypergradient
    model.train()
    _ = model(batch, train_loss=True)
    if fit:
        fit_gpytorch_scipy(model.mll)
    elif phi is not None:
        with torch.no_grad():
            for p, v in zip(model.gp_params(), phi):
                p.copy_(torch.as_tensor(v, dtype=p.dtype).reshape(p.shape))
    model.train()
    fe_names = [n for n, _ in model.named_parameters() if not n.startswith("gp_")]
    gp_names = [n for n, _ in model.named_parameters() if n.startswith("gp_")]

    def f_inner(po, pi):
        d = {**dict(zip(fe_names, po)), **dict(zip(gp_names, pi))}
        return functional_call(model, d, batch, kwargs={"train_loss": True, "is_functional_call": True})

    def f_outer(po, pi):
        d = {**dict(zip(fe_names, po)), **dict(zip(gp_names, pi))}
        return functional_call(model, d, batch, kwargs={"train_loss": False, "predictive_val_loss": True,
                                                        "is_functional_call": True})

    Fv = cauchy_hypergradient(f_outer, f_inner, tuple(model.feature_extractor_params()),
                              tuple(model.gp_params()), model.device,
                              ignore_grad_correction=model.config.ignore_grad_correction)
    grads = {n: p.grad.detach().clone() for n, p in model.named_parameters() if not n.startswith("gp_")}
    gp_grads = [p.grad.detach().clone() for p in model.gp_params()]
    return dict(F=Fv.detach().clone(), grads=grads, gp_grads=gp_grads,
                phi_star=[p.detach().clone() for p in model.gp_params()])
