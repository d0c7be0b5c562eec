"""Authoring-container helper: the reference's max-pool readout (graph_readout.py:262-296) routes the gradient of each
(graph, channel) to the arg-max node, so where the two largest node values of a graph differ by rounding noise (~1e-7
relative) the reference's OWN gradient depends on its rounding.  Golden gradient cases are therefore chosen among task ids
whose smallest relative top-2 gap is well above fp32 noise.  usage: python tests/golden/scan_ties.py [layers] [n_s] [n_q]"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref_runner as rr  # noqa: E402

rr.install()
from adkf_ift_b200 import synthetic  # noqa: E402


def min_top2_gap(model, feats):
    """smallest non-zero (top1 - top2) / |top1| over every (graph, channel) of the max-pool readout input."""
    got = {}
    pooler = model.graph_feature_extractor.readout._max_pooler

    def hook(mod, inp):
        got["x"], got["ids"], got["G"] = inp[0].detach(), inp[1].detach(), int(inp[2])
    h = pooler.register_forward_pre_hook(hook)
    with torch.no_grad():
        model.graph_feature_extractor(feats)
    h.remove()
    x, ids, G = got["x"], got["ids"], got["G"]
    worst = np.inf
    for g in range(G):
        rows = x[ids == g]
        if rows.shape[0] < 2:
            continue
        top = torch.topk(rows, 2, dim=0).values
        gaps = (top[0] - top[1]) / top[0].abs().clamp_min(1e-30)
        gaps = gaps[gaps > 0]          # exact ties (identical atoms / symmetric twins) are resolved by index, not by rounding
        if gaps.numel():
            worst = min(worst, gaps.min().item())
    return worst


def main():
    layers = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    n_s = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    n_q = int(sys.argv[3]) if len(sys.argv) > 3 else 64
    kernel = "rbf" if layers == 2 else "matern"
    torch.manual_seed(0)
    model = rr.build_reference_model(rr.reference_config(num_layers=layers, gp_kernel=kernel, use_numeric_labels=False))
    for task_id in range(0, 40):
        ts = synthetic.make_task_sample(task_id, n_s, n_q)
        host, dev = rr.reference_dkt_sample(ts, False)
        b = dev.batches[0]
        gap = min(min_top2_gap(model, b.support_features), min_top2_gap(model, b.query_features))
        print(f"task {task_id}: min relative top-2 gap of the max-pool readout {gap:.3e}", flush=True)


if __name__ == "__main__":
    main()
