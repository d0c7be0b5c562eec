"""ncu target: four representative GEMM launches of the C2 step (short-K message MLP, short-K BOOM expand, short-K input
gradient, long-K readout hidden), each run 3 times."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from adkf_ift_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
V, e = 133728, 236867
XG, W2, b1 = torch.randn(e, 256, device=dev), torch.randn(12 * 192, 64, device=dev), torch.randn(12 * 192, device=dev)
pre = torch.empty(e, 768, device=dev)
x128, w1, bb = torch.randn(V, 128, device=dev), torch.randn(1024, 128, device=dev), torch.randn(1024, device=dev)
wT = torch.randn(128, 3072, device=dev)
x1408, wr, br = torch.randn(V, 1408, device=dev), torch.randn(768, 1408, device=dev), torch.randn(768, device=dev)
for _ in range(3):
    ops.gemm_tf32x3_batched(XG, False, 0, 64, W2, False, 3 * 192, 0, e, 192, 64, 4, pre, 768, 192, bias=b1, bias_stride=3 * 192)
    ops.gemm_tf32x3(x128, False, w1, False, bb, "leaky_relu")
    ops.gemm_tf32x3(x128, False, wT, True)
    ops.gemm_tf32x3(x1408, False, wr, False, br, "relu")
torch.cuda.synchronize()
print("ok")
