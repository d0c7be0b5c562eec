"""Development probe: per-tensor difference of the whole-model hypergradient between tf32 and bf16 correction products of the
pair GEMM (adkf_gemm_configure), plus the same for the masked and the batched entry points on their real operands."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from adkf_ift_b200 import _cabi, ops  # noqa: E402

lib = _cabi.load()
dev = torch.device("cuda:0")
from test_gpu_model import build  # noqa: E402
from adkf_ift_b200.adaptive_dkt_utils import run_on_task_batch  # noqa: E402

g, model, host, tb = build(os.path.join(ROOT, "tests", "golden"), "c1_classification", "matern", dev)
phi = torch.from_numpy(g["train_phi_star"].astype(np.float32)).view(1, 3)
grads, feats = {}, {}
for corr in (0, 1):
    lib.adkf_gemm_configure(2, corr)
    model.zero_grad(set_to_none=True)
    with torch.no_grad():
        feats[corr] = model.embed(tb).double()
    run_on_task_batch(model, tb, train=True, grad_scale=1.0, phi=phi)
    grads[corr] = {n: p.grad.double().clone() for n, p in model.named_parameters() if p.grad is not None}
lib.adkf_gemm_configure(2, 0)
print("features: rel diff", float((feats[1] - feats[0]).abs().max() / feats[0].abs().max()))
top = max(float(v.norm()) for v in grads[0].values())
rows = []
for n in grads[0]:
    a, b = grads[0][n], grads[1][n]
    rows.append((float((a - b).norm() / max(float(a.norm()), 1e-3 * top)), float(abs(a.norm() - b.norm()) / max(float(a.norm()), 1e-3 * top)), n, tuple(a.shape)))
rows.sort(reverse=True)
for r in rows:
    print(f"{r[0]:.2e} norm-diff {r[1]:.2e} {r[2]} {r[3]}")
