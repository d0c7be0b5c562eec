"""Development aid: every gemm_tf32x3 call of one training task of a golden case is re-computed in fp64 (and in cuBLAS fp32)
on the same operands; prints the error of each.  usage: python tools/linear_diag.py [case] [kernel]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import test_gpu_model as T  # noqa: E402
from adkf_ift_b200 import ops  # noqa: E402
from adkf_ift_b200.adaptive_dkt_utils import run_on_task_batch  # noqa: E402

orig = ops.gemm_tf32x3
rows = []


def wrapped(a, trans_a, b, trans_b, bias=None, act=None, out=None):
    y = orig(a, trans_a, b, trans_b, bias, act, out)
    A = a.double().t() if trans_a else a.double()
    B = b.double() if trans_b else b.double().t()
    ref = A @ B
    y32 = (a.t() if trans_a else a) @ (b if trans_b else b.t())
    if bias is not None:
        ref = ref + bias.double()
        y32 = y32 + bias
    if ops.ACT_IDS[act] == 1:
        ref, y32 = ref.clamp_min(0), y32.clamp_min(0)
    elif ops.ACT_IDS[act] == 2:
        ref, y32 = torch.where(ref > 0, ref, 0.01 * ref), torch.where(y32 > 0, y32, 0.01 * y32)
    den = ref.abs().max().clamp_min(1e-300)
    # scale-aware error: |err| / (|A| |B|) elementwise bound would be ideal; report max/max and the l2 ratio
    e_tc = float((y.double() - ref).abs().max() / den)
    e_32 = float((y32.double() - ref).abs().max() / den)
    l_tc = float((y.double() - ref).norm() / ref.norm().clamp_min(1e-300))
    l_32 = float((y32.double() - ref).norm() / ref.norm().clamp_min(1e-300))
    bias_tc = float(((y.double() - ref) * ref.sign()).sum() / ref.abs().sum().clamp_min(1e-300))
    rows.append((tuple(ref.shape), A.shape[1], int(trans_a), int(trans_b), e_tc, e_32, l_tc, l_32, bias_tc))
    return y


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "s64_rbf_l2"
    kernel = sys.argv[2] if len(sys.argv) > 2 else "rbf"
    dev = torch.device("cuda:0")
    torch.backends.cuda.matmul.allow_tf32 = False
    g, model, host, tb = T.build(os.path.join(ROOT, "tests", "golden"), name, kernel, dev)
    phi = torch.from_numpy(g["train_phi_star"].astype(np.float32)).view(1, 3)
    ops.gemm_tf32x3 = wrapped
    model.zero_grad(set_to_none=True)
    run_on_task_batch(model, tb, train=True, grad_scale=1.0, phi=phi)
    torch.cuda.synchronize()
    print("shape K tA tB | max-err tc, fp32 | l2-err tc, fp32 | signed shrink tc")
    for r in rows:
        print(f"{str(r[0]):16s} K={r[1]:6d} {r[2]}{r[3]}  {r[4]:.2e} {r[5]:.2e}   {r[6]:.2e} {r[7]:.2e}   {r[8]:+.2e}")


if __name__ == "__main__":
    main()
