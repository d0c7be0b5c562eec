"""Development probe: the short-K GEMM shapes of the step under the ADKF_GEMM_DEBUG bits (2 no MMAs, 4 no global stores,
32 no TMEM drains, 64 no epilogue) -- where does a short-K tile spend its time?"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from adkf_ift_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
V = 133728


def bench(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


x128, w1, b1 = torch.randn(V, 128, device=dev), torch.randn(1024, 128, device=dev), torch.randn(1024, device=dev)
wT = torch.randn(128, 3072, device=dev)
x1408, wr, br = torch.randn(V, 1408, device=dev), torch.randn(768, 1408, device=dev), torch.randn(768, device=dev)
x1024, w2 = torch.randn(V, 1024, device=dev), torch.randn(128, 1024, device=dev)
print(f"ADKF_GEMM_DEBUG={os.environ.get('ADKF_GEMM_DEBUG', '0'):>3s}"
      f"  fwd N=1024 K=128 {bench(lambda: ops.gemm_tf32x3(x128, False, w1, False, b1, 'leaky_relu')):.3f}"
      f"  dX N=3072 K=128 {bench(lambda: ops.gemm_tf32x3(x128, False, wT, True)):.3f}"
      f"  fwd N=128 K=1024 {bench(lambda: ops.gemm_tf32x3(x1024, False, w2, False)):.3f}"
      f"  fwd N=768 K=1408 {bench(lambda: ops.gemm_tf32x3(x1408, False, wr, False, br, 'relu')):.3f} ms")
