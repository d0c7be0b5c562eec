"""Development probe: where the 3xTF32 GEMM spends its time (ADKF_GEMM_DEBUG=0|1|2|3 set by the caller)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from adkf_ift_b200 import ops
dev = torch.device("cuda:0")
def bench(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
V = 133728
out = []
for (K, N) in [(1408, 768), (3072, 128), (128, 1024)]:
    x = torch.randn(V, K, device=dev); w = torch.randn(N, K, device=dev) / K ** 0.5
    t = bench(lambda: ops.linear_tf32x3_forward(x, w, None))
    out.append(f"fwd K={K} N={N}: {t:.3f} ms {2.0*V*K*N/t/1e9:.0f} TF/s")
dy = torch.randn(V, 768, device=dev); w = torch.randn(768, 1408, device=dev); x = torch.randn(V, 1408, device=dev)
t = bench(lambda: ops.gemm_tf32x3(dy, False, w, True)); out.append(f"dX 768->1408: {t:.3f} ms {2.0*V*768*1408/t/1e9:.0f} TF/s")
t = bench(lambda: ops.gemm_tf32x3(dy, True, x, True)); out.append(f"dW 768x1408: {t:.3f} ms {2.0*V*768*1408/t/1e9:.0f} TF/s")
print("DEBUG", os.environ.get("ADKF_GEMM_DEBUG", "0"), " | ".join(out))
