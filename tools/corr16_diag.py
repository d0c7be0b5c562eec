"""Development probe: which GEMMs of a real meta-training step lose accuracy when the pair kernel's correction products run in
bf16 (adkf_gemm_configure(2, 1)) instead of tf32.  Every gemm_tf32x3 / masked call of one fixed-phi training task of the
c1_classification fixture is re-run in both modes and compared with an fp64 matmul of the same operands."""
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
rows = []
orig = ops.gemm_tf32x3


def wrapped(a, trans_a, b, trans_b, bias=None, act=None, out=None):
    A = a.double().t() if trans_a else a.double()
    B = b.double() if trans_b else b.double().t()
    ref = A @ B
    if bias is not None:
        ref = ref + bias.double()
    if ops.ACT_IDS[act] == 1:
        ref = torch.relu(ref)
    elif ops.ACT_IDS[act] == 2:
        ref = torch.nn.functional.leaky_relu(ref, 0.01)
    errs = []
    for corr in (0, 1):
        lib.adkf_gemm_configure(2, corr)
        y = orig(a, trans_a, b, trans_b, bias, act, None)
        errs.append(float((y.double() - ref).abs().max() / ref.abs().max().clamp_min(1e-300)))
        errs.append(float((y.double() - ref).norm() / ref.norm().clamp_min(1e-300)))
    lib.adkf_gemm_configure(2, 0)
    M, N = ref.shape
    K = A.shape[1]
    rows.append((f"{'T' if trans_a else 'K'}{'T' if trans_b else 'K'}", M, N, K, *errs,
                 float(a.abs().max()), float(a.abs().mean()), float(b.abs().max()), float(b.abs().mean())))
    return orig(a, trans_a, b, trans_b, bias, act, out)


ops.gemm_tf32x3 = wrapped
orig_masked = ops.gemm_tf32x3_masked


def wrapped_masked(a, b, mask, mask_act):
    ref = (a.double() @ b.double()) * torch.where(mask.double() > 0, 1.0, 0.0 if mask_act == 1 else 0.01)
    errs = []
    for corr in (0, 1):
        lib.adkf_gemm_configure(2, corr)
        y = orig_masked(a, b, mask, mask_act)
        errs.append(float((y.double() - ref).abs().max() / ref.abs().max().clamp_min(1e-300)))
        errs.append(float((y.double() - ref).norm() / ref.norm().clamp_min(1e-300)))
    lib.adkf_gemm_configure(2, 0)
    rows.append(("mKT", a.shape[0], b.shape[1], a.shape[1], *errs, float(a.abs().max()), float(a.abs().mean()), float(b.abs().max()),
                 float(b.abs().mean())))
    return orig_masked(a, b, mask, mask_act)


ops.gemm_tf32x3_masked = wrapped_masked
orig_batched = ops.gemm_tf32x3_batched


def wrapped_batched(a, trans_a, a_mn, a_k, b, trans_b, b_mn, b_k, M, N, K, batch, out, ldd, d_stride, bias=None, bias_stride=0, act=None,
                    allow_splitk=False):
    outs = []
    for corr in (0, 1):
        lib.adkf_gemm_configure(2, corr)
        o = torch.zeros_like(out)
        orig_batched(a, trans_a, a_mn, a_k, b, trans_b, b_mn, b_k, M, N, K, batch, o, ldd, d_stride, bias, bias_stride, act, allow_splitk)
        outs.append(o.double().clone())
    lib.adkf_gemm_configure(2, 0)
    # fp64 reference of batch 0 only
    A0 = (a.double()[:K, :M].t() if trans_a else a.double()[:M, :K])
    B0 = (b.double()[:K, :N] if trans_b else b.double()[:N, :K].t())
    ref = A0 @ B0
    if bias is not None:
        ref = ref + bias.double()[:N]
    got = [o.view(-1)[:M * ldd].view(M, ldd)[:, :N] if ldd != N else o.view(-1)[:M * N].view(M, N) for o in outs]
    errs = []
    for y in got:
        errs.append(float((y - ref).abs().max() / ref.abs().max().clamp_min(1e-300)))
        errs.append(float((y - ref).norm() / ref.norm().clamp_min(1e-300)))
    rows.append((f"b{'T' if trans_a else 'K'}{'T' if trans_b else 'K'}", M, N, K, *errs, float(a.abs().max()), float(a.abs().mean()),
                 float(b.abs().max()), float(b.abs().mean())))
    rows.append(("bdif", M, N, K, float((outs[1] - outs[0]).abs().max() / outs[0].abs().max().clamp_min(1e-300)), 0.0,
                 float((outs[1] - outs[0]).norm() / outs[0].norm().clamp_min(1e-300)), 0.0, 0.0, 0.0, 0.0, 0.0))
    return orig_batched(a, trans_a, a_mn, a_k, b, trans_b, b_mn, b_k, M, N, K, batch, out, ldd, d_stride, bias, bias_stride, act, allow_splitk)


ops.gemm_tf32x3_batched = wrapped_batched
from test_gpu_model import build  # noqa: E402
from adkf_ift_b200.adaptive_dkt_utils import run_on_task_batch  # noqa: E402

g, model, host, tb = build(os.path.join(ROOT, "tests", "golden"), "c1_classification", "matern", dev)
phi = torch.from_numpy(g["train_phi_star"].astype(np.float32)).view(1, 3)
model.zero_grad(set_to_none=True)
run_on_task_batch(model, tb, train=True, grad_scale=1.0, phi=phi)
torch.cuda.synchronize()
print(f"{'form':4s} {'M':>7s} {'N':>6s} {'K':>7s} | tf32corr max / l2 | bf16corr max / l2 | |a|max mean |b|max mean")
worst = sorted(rows, key=lambda r: -r[6])
for r in worst[:25]:
    print(f"{r[0]:4s} {r[1]:7d} {r[2]:6d} {r[3]:7d} | {r[4]:.2e} {r[5]:.2e} | {r[6]:.2e} {r[7]:.2e} | {r[8]:.2e} {r[9]:.2e} {r[10]:.2e} {r[11]:.2e}")
print("calls", len(rows), "median bf16 max err", float(np.median([r[6] for r in rows])), "median tf32 max err", float(np.median([r[4] for r in rows])))
