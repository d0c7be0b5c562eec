"""Generates tests/golden/theta_grads_f64.npz: FP64 whole-model hypergradients from the reference's OWN
cauchy_hypergradient / functional_call / ADKTModel files (unmodified, /root/reference, through oracle/shims.py),
evaluated at the SAME weights (fp32 values held in fp64 tensors), the same synthetic task and the same fixed GP
parameters phi* as the fp32 fixtures c1_classification / c1_regression / s64_rbf_l2 (tests/golden/make_golden.py).

Why: the fp32 fixtures cannot separate the CUDA path's error from the reference's own fp32 rounding.  With an fp64
truth the -m gpu tests assert  err(cuda vs f64) <= max(1e-4, 3 x err(reference-fp32 vs f64))  -- the north star's 1e-4
wherever the reference itself achieves it ("Next round" 1b; SURVEY App. D recipe: set_default_dtype(float64)).

    python tests/golden/make_golden_f64.py            (authoring container only; ~minutes on 8 cores)
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

REPORT = rr.install()
torch.set_default_dtype(torch.float64)

from adkf_ift_b200 import synthetic  # noqa: E402

CASES = [("c1_classification", "matern"), ("c1_regression", "matern"), ("s64_rbf_l2", "rbf")]


def probe_vector(name, shape):
    g = torch.Generator().manual_seed(zlib.crc32(("probe:" + name).encode()) & 0x7FFFFFFF)
    return torch.randn(shape, generator=g, dtype=torch.float64)


def one_case(name, kernel):
    g32 = np.load(os.path.join(HERE, name + ".npz"))
    task_id, n_s, n_q, numeric, layers = (int(x) for x in g32["meta"])
    cfg = rr.reference_config(num_layers=layers, gp_kernel=kernel, use_numeric_labels=bool(numeric))
    model = rr.build_reference_model(cfg, dtype=torch.float64)
    with torch.no_grad():                               # the fp32 weights of every other fixture, exactly
        for n, p in model.named_parameters():
            if not n.startswith("gp_"):
                p.copy_(p.float().double())
    assert all(p.dtype == torch.float64 for p in model.parameters())
    ts = synthetic.make_task_sample(task_id, n_s, n_q)
    _, dev = rr.reference_dkt_sample(ts, bool(numeric), dtype=torch.float64)
    batch = dev.batches[0]
    phi = [torch.tensor(float(v), dtype=torch.float64) for v in g32["train_phi_star"]]
    tr = rr.reference_meta_train_task(model, batch, fit=False, phi=phi)
    names = [str(x) for x in g32["grad_names"]]
    out = {
        f"{name}::F": np.float64(tr["F"].item()),
        f"{name}::dF_dphi": np.array([float(x.reshape(-1)[0]) for x in tr["gp_grads"]]),
        f"{name}::grad_norms": np.array([tr["grads"][n].norm().item() for n in names]),
        f"{name}::grad_dots": np.array([(tr["grads"][n] * probe_vector(n, tr["grads"][n].shape)).sum().item() for n in names]),
    }
    for n in ("fc.2.bias", "graph_feature_extractor.gnn.gnn_blocks.0.alpha", "graph_feature_extractor.init_node_proj.weight",
              "graph_feature_extractor.gnn.gnn_blocks.0.mp_layers.0.message_fns.0._layers.0.weight", "fc.0.bias"):
        out[f"{name}::grad::{n}"] = tr["grads"][n].numpy()
    # how far the reference's own fp32 run is from this fp64 truth (printed; the tests recompute it from the two fixtures)
    n32, n64 = g32["grad_norms"], out[f"{name}::grad_norms"]
    big = n64 > 1e-3 * n64.max()
    print(name, "F64", out[f"{name}::F"], "F32", float(g32["train_F"]),
          "fp32-vs-fp64 max rel grad-norm err", float((np.abs(n32 - n64)[big] / n64[big]).max()))
    return out


if __name__ == "__main__":
    print(REPORT)
    out = {}
    for name, kernel in CASES:
        out.update(one_case(name, kernel))
    path = os.path.join(HERE, "theta_grads_f64.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")
