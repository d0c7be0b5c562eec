"""Development probe: short-K GEMM timings under ADKF_GEMM_DEBUG bits (1 no split, 2 no MMA, 4 no stores, 8 no bias, 16 no TMA)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from adkf_ift_b200 import ops
dev = torch.device("cuda:0")
V, e = 133728, 236867
def bench(fn, reps=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
XG, W2, b1 = torch.randn(e, 256, device=dev), torch.randn(12 * 192, 64, device=dev), torch.randn(12 * 192, device=dev)
pre = torch.empty(e, 768, device=dev)
x128, w1, bb = torch.randn(V, 128, device=dev), torch.randn(1024, 128, device=dev), torch.randn(1024, device=dev)
wT = torch.randn(128, 3072, device=dev)
x1408, wr, br = torch.randn(V, 1408, device=dev), torch.randn(768, 1408, device=dev), torch.randn(768, device=dev)
t1 = bench(lambda: ops.gemm_tf32x3_batched(XG, False, 0, 64, W2, False, 3 * 192, 0, e, 192, 64, 4, pre, 768, 192, bias=b1, bias_stride=3 * 192))
t2 = bench(lambda: ops.gemm_tf32x3(x128, False, w1, False, bb, "leaky_relu"))
t3 = bench(lambda: ops.gemm_tf32x3(x128, False, wT, True))
t4 = bench(lambda: ops.gemm_tf32x3(x1408, False, wr, False, br, "relu"))
x3072, wp = torch.randn(V, 3072, device=dev), torch.randn(128, 3072, device=dev)
t5 = bench(lambda: ops.gemm_tf32x3(x3072, False, wp, False, None, None))
print(f"DEBUG={os.environ.get('ADKF_GEMM_DEBUG','0'):>3s} pna_fwd {t1:.3f}  boom1_fwd {t2:.3f}  proj_dX {t3:.3f}  readout_hidden {t4:.3f}  proj_fwd {t5:.3f} ms")
