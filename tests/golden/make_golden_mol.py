"""Generates tests/golden/c4_moleculenet.npz (BASELINE config 4) by running the reference's OWN MoleculeNet files
unmodified -- chem_lib/models/encoder.py (GNN_Encoder / GIN) and chem_lib/models/adkf_model.py (ADKFModel), plus
fs_mol/utils/cauchy_hypergradient.py and _stateless.functional_call -- through oracle/shims.py (gpytorch / botorch
restated) and oracle/torch_geometric_lite.py (MessagePassing / pooling restated).  The loop body below is the per-task
body of ADKF_Meta_Trainer.train_step / test_step (adkfift_trainer.py:163-213, 279-286); the trainer module itself cannot
be imported here (torchmetrics, rdkit-backed datasets).  Authoring container only:

    python tests/golden/make_golden_mol.py
"""
import importlib
import os
import sys
import types
import zlib
from types import SimpleNamespace

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import shims, torch_geometric_lite  # noqa: E402

REPORT = shims.install()
REPORT["torch_geometric"] = torch_geometric_lite.install()

from botorch.optim.fit import fit_gpytorch_scipy  # noqa: E402
from fs_mol.utils._stateless import functional_call  # noqa: E402
from fs_mol.utils.cauchy_hypergradient import cauchy_hypergradient  # noqa: E402

from adkf_ift_b200 import synthetic  # noqa: E402

T, N_S, N_Q = 3, 20, 32


def reference_models_package():
    """Import encoder.py / adkf_model.py as members of a synthetic package so that chem_lib/models/__init__.py (which pulls
    in the trainers and their unavailable dependencies) is not executed."""
    path = os.path.join(shims.REFERENCE_ROOT, "MoleculeNet", "chem_lib", "models")
    pkg = types.ModuleType("ref_chem_models")
    pkg.__path__ = [path]
    sys.modules["ref_chem_models"] = pkg
    return importlib.import_module("ref_chem_models.adkf_model")


def data_of(d):
    return SimpleNamespace(x=torch.from_numpy(d["x"]), edge_index=torch.from_numpy(d["edge_index"]),
                           edge_attr=torch.from_numpy(d["edge_attr"]), batch=torch.from_numpy(d["batch"]),
                           y=torch.from_numpy(d["y"]))


def probe_vector(name, shape):
    g = torch.Generator().manual_seed(zlib.crc32(("probe:" + name).encode()) & 0x7FFFFFFF)
    return torch.randn(shape, generator=g, dtype=torch.float64)


def main():
    mod = reference_models_package()
    args = SimpleNamespace(emb_dim=300, enc_layer=5, JK="last", dropout=0.0, enc_pooling="mean", enc_gnn="gin",
                           enc_batch_norm=1, pretrained=0, gpu_id=0)
    torch.manual_seed(0)
    model = mod.ADKFModel(args)
    synthetic.deterministic_init_(model, seed=1)
    device = torch.device("cpu")
    fe_names = [n for n, _ in model.named_parameters() if not n.startswith("gp_")]
    out = {"grad_names": np.array(fe_names)}
    accum = [torch.zeros_like(p) for p in model.feature_extractor_params()]
    for t in range(T):
        s_data, q_data = data_of(synthetic.make_mol_graphs(100 + 2 * t, N_S)), data_of(synthetic.make_mol_graphs(101 + 2 * t, N_Q))
        train_data = {"s_data": s_data, "q_data": q_data, "s_label": s_data.y}
        model.train()
        with torch.no_grad():
            zs, _ = model.mol_encoder(s_data.x, s_data.edge_index, s_data.edge_attr, s_data.batch)
            zq, _ = model.mol_encoder(q_data.x, q_data.edge_index, q_data.edge_attr, q_data.batch)
        out[f"t{t}_z_support"], out[f"t{t}_z_query"] = zs.numpy(), zq.numpy()
        _ = model(s_data=s_data, q_data=None, s_label=s_data.y, train_loss=True)
        out[f"t{t}_phi_init"] = np.array([p.detach().reshape(-1)[0].item() for p in model.gp_params()])
        fit_gpytorch_scipy(model.mll)
        model.train()
        gp_names = [n for n, _ in model.named_parameters() if n.startswith("gp_")]

        def f_inner(params_outer, params_inner):
            d = {**dict(zip(fe_names, params_outer)), **dict(zip(gp_names, params_inner))}
            return functional_call(model, d, (train_data["s_data"], train_data["q_data"]),
                                   kwargs={"s_label": train_data["s_label"], "train_loss": True, "is_functional_call": True})

        def f_outer(params_outer, params_inner):
            d = {**dict(zip(fe_names, params_outer)), **dict(zip(gp_names, params_inner))}
            return functional_call(model, d, (train_data["s_data"], train_data["q_data"]),
                                   kwargs={"s_label": train_data["s_label"], "train_loss": False, "predictive_val_loss": True,
                                           "is_functional_call": True})

        F = cauchy_hypergradient(f_outer, f_inner, tuple(model.feature_extractor_params()), tuple(model.gp_params()), device,
                                 ignore_grad_correction=False)
        out[f"t{t}_F"] = np.array(F.item())
        out[f"t{t}_phi_star"] = np.array([p.detach().reshape(-1)[0].item() for p in model.gp_params()])
        for i, p in enumerate(model.feature_extractor_params()):
            accum[i] += p.grad.data.clone() / T
        print("task", t, "F", F.item(), "phi*", out[f"t{t}_phi_star"])
    out["grad_norms"] = np.array([g.double().norm().item() for g in accum])
    out["grad_dots"] = np.array([(g.double() * probe_vector(n, g.shape)).sum().item() for n, g in zip(fe_names, accum)])
    out["grad::mol_encoder.gnn.batch_norms.4.bias"] = accum[fe_names.index("mol_encoder.gnn.batch_norms.4.bias")].numpy()
    out["grad::mol_encoder.gnn.gnns.0.edge_embedding1.weight"] = accum[fe_names.index("mol_encoder.gnn.gnns.0.edge_embedding1.weight")].numpy()
    out["grad::mol_encoder.gnn.x_embedding2.weight"] = accum[fe_names.index("mol_encoder.gnn.x_embedding2.weight")].numpy()

    # ---- test_step on a fresh model: fit on train-mode support features, predict in eval mode (running statistics) ----
    torch.manual_seed(0)
    model = mod.ADKFModel(args)
    synthetic.deterministic_init_(model, seed=1)
    s_data, q_data = data_of(synthetic.make_mol_graphs(900, N_S)), data_of(synthetic.make_mol_graphs(901, 48))
    model.train()
    _ = model(s_data=s_data, q_data=None, s_label=s_data.y, train_loss=True)
    fit_gpytorch_scipy(model.mll)
    model.eval()
    with torch.no_grad():
        preds, labels = model.forward_query_loader(s_data, [SimpleNamespace(**vars(q_data), to=lambda d, q=q_data: q)],
                                                   train_loss=None, s_label=s_data.y)
    out["test_phi_star"] = np.array([p.detach().reshape(-1)[0].item() for p in model.gp_params()])
    out["test_preds"] = preds.numpy()
    out["test_running_mean_l0"] = model.mol_encoder.gnn.batch_norms[0].running_mean.numpy()
    out["test_running_var_l4"] = model.mol_encoder.gnn.batch_norms[4].running_var.numpy()
    out["meta"] = np.array([T, N_S, N_Q])
    path = os.path.join(HERE, "c4_moleculenet.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; test phi*", out["test_phi_star"])


if __name__ == "__main__":
    print(REPORT)
    main()
