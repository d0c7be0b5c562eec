"""Generates tests/golden/ignore_grad_correction.npz: one meta-training task through the reference's OWN cauchy_hypergradient with
`ignore_grad_correction=True` (ADKTModelTrainerConfig flag, adaptive_dkt_utils.py:66; cauchy_hypergradient.py:66-70: the
hypergradient degenerates to the direct gradient dF/dtheta).  Authoring container only:
    python tests/golden/make_golden_flags.py"""
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

NUM_LAYERS, TASK, N_S, N_Q = 2, 21, 16, 24


def probe_vector(name, shape):
    g = torch.Generator().manual_seed(zlib.crc32(("probe:" + name).encode()) & 0x7FFFFFFF)
    return torch.randn(shape, generator=g, dtype=torch.float64)


def main(path):
    torch.manual_seed(0)
    cfg = rr.reference_config(num_layers=NUM_LAYERS, ignore_grad_correction=True)
    model = rr.build_reference_model(cfg)
    ts = synthetic.make_task_sample(TASK, N_S, N_Q)
    _, dev = rr.reference_dkt_sample(ts, False)
    tr = rr.reference_meta_train_task(model, dev.batches[0])
    names = [n for n, _ in model.named_parameters() if not n.startswith("gp_")]
    out = dict(meta=np.array([TASK, N_S, N_Q, NUM_LAYERS]), F=tr["F"].numpy(),
               phi_star=np.array([p.reshape(-1)[0].item() for p in tr["phi_star"]]),
               grad_names=np.array(names),
               grad_norms=np.array([tr["grads"][n].double().norm().item() for n in names]),
               grad_dots=np.array([(tr["grads"][n].double() * probe_vector(n, tr["grads"][n].shape)).sum().item() for n in names]))
    np.savez_compressed(path, **out)
    print("wrote", path, "F", out["F"], "phi*", out["phi_star"], "max |g|", out["grad_norms"].max())


if __name__ == "__main__":
    main(os.path.join(HERE, "ignore_grad_correction.npz"))
