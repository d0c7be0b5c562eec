"""Development probe: accuracy (signed bias) and speed of the tcgen05 3xTF32 GEMM vs cuBLAS fp32."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from adkf_ift_b200 import ops

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
for (M, K, N) in [(1024, 1408, 768), (1024, 128, 1024), (1024, 3072, 128)]:
    x = torch.randn(M, K, generator=g).to(dev)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(dev)
    ref = x.double() @ w.double().t()
    y = ops.linear_tf32x3_forward(x, w, None)
    y32 = x @ w.t()
    err = (y.double() - ref)
    bias_toward_zero = float((err * torch.sign(ref)).mean() / ref.abs().mean())
    print(f"M={M} K={K} N={N}: max rel {float(err.abs().max() / ref.abs().max()):.2e} (cuBLAS {float((y32.double() - ref).abs().max() / ref.abs().max()):.2e}); "
          f"mean signed err toward |ref| growth {bias_toward_zero:+.2e}; rms rel {float(err.pow(2).mean().sqrt() / ref.pow(2).mean().sqrt()):.2e}")

def bench(fn, reps=20):
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

V = 133728
for (K, N) in [(3072, 128), (128, 1024), (1024, 128), (1408, 768), (768, 768), (2560, 2048)]:
    M = V if K != 2560 else 5120
    x = torch.randn(M, K, device=dev)
    w = torch.randn(N, K, device=dev) / K ** 0.5
    b = torch.randn(N, device=dev)
    t_tc = bench(lambda: ops.linear_tf32x3_forward(x, w, b, "relu"))
    t_cb = bench(lambda: torch.relu(torch.nn.functional.linear(x, w, b)))
    fl = 2.0 * M * K * N
    print(f"M={M} K={K} N={N}: tcgen05 3xTF32 {t_tc:.3f} ms ({fl / t_tc / 1e9:.1f} TFLOP/s fp32-equivalent), cuBLAS fp32+relu {t_cb:.3f} ms ({fl / t_cb / 1e9:.1f} TFLOP/s)")
