"""Development probe: which class of GEMM launches makes the whole-model hypergradient differ when its correction products run
in bf16.  The c1_classification training task is run with bf16 corrections enabled for ONE class of launches at a time and the
per-tensor gradient difference against the all-tf32 run is reported."""
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
orig, orig_m, orig_b = ops.gemm_tf32x3, ops.gemm_tf32x3_masked, ops.gemm_tf32x3_batched
active = {"cls": None, "fwd": True}


def form(ta, tb_):
    return ("T" if ta else "K") + ("T" if tb_ else "K")


def w_plain(a, trans_a, b, trans_b, bias=None, act=None, out=None):
    on = active["cls"] in ("all", "plain", "plain" + form(trans_a, trans_b))
    if on and active["cls"].startswith("plain") and active.get("minK") and a.shape[1 if not trans_a else 0] < active["minK"]:
        on = False
    lib.adkf_gemm_configure(2, 1 if on else 0)
    y = orig(a, trans_a, b, trans_b, bias, act, out)
    lib.adkf_gemm_configure(2, 0)
    return y


def w_masked(a, b, mask, mask_act):
    lib.adkf_gemm_configure(2, 1 if active["cls"] in ("all", "masked") else 0)
    y = orig_m(a, b, mask, mask_act)
    lib.adkf_gemm_configure(2, 0)
    return y


def w_batched(a, trans_a, *rest, **kw):
    trans_b = rest[3]
    lib.adkf_gemm_configure(2, 1 if active["cls"] in ("all", "batched", "batched" + form(trans_a, trans_b)) else 0)
    y = orig_b(a, trans_a, *rest, **kw)
    lib.adkf_gemm_configure(2, 0)
    return y


ops.gemm_tf32x3, ops.gemm_tf32x3_masked, ops.gemm_tf32x3_batched = w_plain, w_masked, w_batched


def run(cls):
    active["cls"] = cls
    model.zero_grad(set_to_none=True)
    run_on_task_batch(model, tb, train=True, grad_scale=1.0, phi=phi)
    return {n: p.grad.double().clone() for n, p in model.named_parameters() if p.grad is not None}


base = run(None)
top = max(float(v.norm()) for v in base.values())
for cls in ("all", "plain", "plainKK", "plainKT", "plainTT", "masked", "batched", "batchedKK", "batchedKT", "batchedTT"):
    gr = run(cls)
    rows = sorted(((float((gr[n] - base[n]).norm() / max(float(base[n].norm()), 1e-3 * top)), n) for n in base), reverse=True)
    med = float(np.median([r[0] for r in rows]))
    print(f"{cls:10s} worst {rows[0][0]:.2e} ({rows[0][1][-60:]})  median {med:.2e}  #>1e-4: {sum(r[0] > 1e-4 for r in rows)}")
