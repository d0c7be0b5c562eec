"""Generates tests/golden/gnn_types.npz: the reference's OWN GraphFeatureExtractor (fs_mol/modules/gnn.py, unmodified,
through oracle/shims.py) with `--gnn_type MultiAggr`, `Plain` and `MultiHeadAttention` on a seeded synthetic task: features of the support
graphs and the gradient of a fixed probe loss sum(features * probe) w.r.t. every extractor parameter (per-tensor norms and
random-probe dot products).  Authoring container only:  python tests/golden/make_golden_gnn_types.py"""
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


def probe_vector(name, shape):
    g = torch.Generator().manual_seed(zlib.crc32(("probe:" + name).encode()) & 0x7FFFFFFF)
    return torch.randn(shape, generator=g, dtype=torch.float64)


def main():
    out = {}
    task_id, n_s, n_q, layers = 3, 24, 8, 2
    for gnn_type in ("MultiAggr", "Plain", "MultiHeadAttention"):
        torch.manual_seed(0)
        model = rr.build_reference_model(rr.reference_config(num_layers=layers, gnn_type=gnn_type))
        ts = synthetic.make_task_sample(task_id, n_s, n_q)
        host, dev = rr.reference_dkt_sample(ts, False)
        feats = dev.batches[0].support_features
        fe = model.graph_feature_extractor
        fe.zero_grad()
        z = fe(feats)
        probe = probe_vector("features", z.shape).to(z.dtype)
        (z * probe).sum().backward()
        names = [n for n, p in fe.named_parameters() if p.grad is not None]
        out[f"{gnn_type}::features"] = z.detach().numpy()
        out[f"{gnn_type}::grad_names"] = np.array(names)
        params = dict(fe.named_parameters())
        out[f"{gnn_type}::grad_norms"] = np.array([params[n].grad.double().norm().item() for n in names])
        out[f"{gnn_type}::grad_dots"] = np.array([(params[n].grad.double() * probe_vector(n, params[n].shape)).sum().item()
                                                  for n in names])
        tower = "0" if gnn_type == "MultiHeadAttention" else "2"
        for n in ("gnn.gnn_blocks.0.alpha", "init_node_proj.weight", f"gnn.gnn_blocks.1.mp_layers.{tower}.message_fns.0._layers.0.weight"):
            out[f"{gnn_type}::grad::{n}"] = params[n].grad.numpy()
        print(gnn_type, z.shape, len(names), "tensors; msg_out_projection", tuple(params["gnn.gnn_blocks.0.msg_out_projection.weight"].shape))
    out["meta"] = np.array([task_id, n_s, n_q, layers])
    np.savez_compressed(os.path.join(HERE, "gnn_types.npz"), **out)


if __name__ == "__main__":
    main()
