"""Generates tests/golden/used_features.npz: the reference's OWN ADKTModel (fs_mol/models/adaptive_dkt.py, unmodified, through
oracle/shims.py) for every `used_features` variant it accepts -- GP inputs of the support set, meta-test phi*, predictive
mean / variance.  Authoring container only:   python tests/golden/make_golden_features.py"""
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

VARIANTS = ["gnn+ecfp+pc-descs+fc", "ecfp+fc", "pc-descs+fc", "gnn+ecfp", "ecfp", "gnn", "pc-descs"]
NUM_LAYERS, TASK, N_S, N_Q = 2, 11, 16, 24


def main(path):
    out = {"meta": np.array([TASK, N_S, N_Q, NUM_LAYERS]), "variants": np.array(VARIANTS)}
    ts = synthetic.make_task_sample(TASK, N_S, N_Q)
    for v in VARIANTS:
        key = v.replace("+", "_").replace("-", "")
        try:
            torch.manual_seed(0)
            cfg = rr.reference_config(num_layers=NUM_LAYERS, used_features=v)
            model = rr.build_reference_model(cfg)
            _, dev = rr.reference_dkt_sample(ts, False)
            mt = rr.reference_meta_test(model, dev)[0]
            out[key + "_z_support"] = model.gp_model.train_inputs[0].detach().numpy()
            out[key + "_phi_star"] = np.array([p.reshape(-1)[0].item() for p in mt["phi_star"]])
            out[key + "_mean"] = mt["mean"].numpy()
            out[key + "_variance"] = mt["variance"].numpy()
            out[key + "_ok"] = np.array(1)
            print(v, "d =", out[key + "_z_support"].shape[1], "phi*", out[key + "_phi_star"])
        except Exception as e:   # the reference itself rejects the variant: recorded, the port must fail too
            out[key + "_ok"] = np.array(0)
            out[key + "_error"] = np.array(type(e).__name__ + ": " + str(e)[:200])
            print(v, "REFERENCE FAILS:", type(e).__name__, str(e)[:200])
    np.savez_compressed(path, **out)
    print("wrote", path)


if __name__ == "__main__":
    main(os.path.join(HERE, "used_features.npz"))
