"""Generates tests/golden/dkt_dkl.npz by running the reference's OWN DKT / DKL model files
(fs_mol/models/dkt.py, dkl.py; unmodified, /root/reference) through oracle/shims.py.  Authoring container only:

    python tests/golden/make_golden_dkt.py

gpytorch / botorch maths is the restated one (oracle/gpytorch_lite): "parity unpinned" at that boundary.
"""
import os
import sys
import zlib

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_runner as rr  # noqa: E402

rr.install()

from adkf_ift_b200 import synthetic  # noqa: E402

NUM_LAYERS, TASK, N_S, N_Q = 2, 5, 16, 32


def probe_vector(name, shape):
    g = torch.Generator().manual_seed(zlib.crc32(("probe:" + name).encode()) & 0x7FFFFFFF)
    return torch.randn(shape, generator=g, dtype=torch.float64)


def ref_cfg(cls, **kw):
    base = rr.reference_config(num_layers=NUM_LAYERS)
    return cls(graph_feature_extractor_config=base.graph_feature_extractor_config, used_features="gnn+ecfp+fc", **kw)


def grads_of(model, out, tag):
    names = [n for n, p in model.named_parameters() if p.grad is not None]
    out[tag + "_grad_names"] = np.array(names)
    g = dict(model.named_parameters())
    out[tag + "_grad_norms"] = np.array([g[n].grad.double().norm().item() for n in names])
    out[tag + "_grad_dots"] = np.array([(g[n].grad.double() * probe_vector(n, g[n].grad.shape)).sum().item()
                                        for n in names])


def main(path):
    from fs_mol.models.dkl import DKLModel
    from fs_mol.models.dkt import DKTModel
    from fs_mol.utils.dkt_utils import DKTModelTrainerConfig
    from fs_mol.utils.dkl_utils import DKLModelTrainerConfig
    from botorch.optim.fit import fit_gpytorch_scipy
    out = {"meta": np.array([TASK, N_S, N_Q, NUM_LAYERS])}
    ts = synthetic.make_task_sample(TASK, N_S, N_Q)
    for use_numeric, lsp, tag in ((False, True, "cls"), (True, False, "reg")):
        _, dev = rr.reference_dkt_sample(ts, use_numeric)
        batch = dev.batches[0]
        # ---- DKT: train loss on support+query at the initial shared phi, then eval-mode posterior -----------
        torch.manual_seed(0)
        m = DKTModel(ref_cfg(DKTModelTrainerConfig, use_numeric_labels=use_numeric, use_lengthscale_prior=lsp))
        synthetic.deterministic_init_(m, seed=0, alpha=0.5)
        out[f"dkt_{tag}_phi0"] = np.array([p.detach().reshape(-1)[0].item() for n, p in m.named_parameters()
                                           if n.startswith("gp_")])
        m.train()
        m.zero_grad()
        loss = m.compute_loss(m(batch))
        loss.backward()
        out[f"dkt_{tag}_loss"] = np.array(loss.item())
        grads_of(m, out, f"dkt_{tag}")
        m.eval()
        mvn = m(batch)
        out[f"dkt_{tag}_mean"] = mvn.mean.detach().numpy()
        out[f"dkt_{tag}_variance"] = mvn.variance.detach().numpy()
        # test-time adaptation: re-fit the shared phi on the support set from the saved values (dkt.py:158-164)
        m.save_gp_params()
        m.test_time_adaptation = True
        mvn = m(batch)
        out[f"dkt_{tag}_tta_mean"] = mvn.mean.detach().numpy()
        out[f"dkt_{tag}_tta_variance"] = mvn.variance.detach().numpy()
        out[f"dkt_{tag}_tta_phi"] = np.array([p.detach().reshape(-1)[0].item() for n, p in m.named_parameters()
                                              if n.startswith("gp_")])
        # ---- DKL: train loss on the support set only, eval posterior -------------------------------------------
        torch.manual_seed(0)
        m = DKLModel(ref_cfg(DKLModelTrainerConfig, use_numeric_labels=use_numeric, use_lengthscale_prior=lsp))
        synthetic.deterministic_init_(m, seed=0, alpha=0.5)
        m.train()
        m.zero_grad()
        loss = m.compute_loss(m(batch, train=True))
        loss.backward()
        out[f"dkl_{tag}_loss"] = np.array(loss.item())
        grads_of(m, out, f"dkl_{tag}")
        m.eval()
        mvn = m(batch, train=False)
        out[f"dkl_{tag}_mean"] = mvn.mean.detach().numpy()
        out[f"dkl_{tag}_variance"] = mvn.variance.detach().numpy()
    np.savez_compressed(path, **out)
    print("wrote", path, {k: (v.shape, float(np.abs(v).max()) if v.dtype.kind == "f" else None) for k, v in out.items()
                          if not k.endswith("names")})


if __name__ == "__main__":
    main(os.path.join(HERE, "dkt_dkl.npz"))
