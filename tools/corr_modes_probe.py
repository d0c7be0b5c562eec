"""Correction-operand schemes of the pair GEMM (gemm_pair.cu) side by side on a forward (K,K) product: error against fp64 and time.
  tf32     two tf32 correction MMAs per k-step
  bf16     one kind::f16 chain over bf16 [lo | hi] x [hi | lo]
  fp16     the same chain over fp16 with the lo parts scaled by 2^11 (exact), accumulator scaled back at the drain
Data sets: N(0,1) activations x N(0,1/K) weights (the layer shapes), and a wide-dynamic-range set (magnitudes 1e-7 .. 1e3,
log-uniform) that leaves the fp16 range on both sides.
usage: python tools/corr_modes_probe.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


def main():
    from adkf_ift_b200 import _cabi, ops
    lib = _cabi.load()
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    modes = [("tf32", 0), ("bf16", 1), ("fp16", 4)]
    M, K, N = 8192, 1408, 768
    sets = {}
    x = torch.randn(M, K, device=dev, generator=g)
    w = torch.randn(N, K, device=dev, generator=g) / K ** 0.5
    sets["normal"] = (x, w)
    mag = 10.0 ** (torch.rand(M, K, device=dev, generator=g) * 10.0 - 7.0)
    sets["wide 1e-7..1e3"] = (torch.randn(M, K, device=dev, generator=g).sign() * mag, w)
    sets["tiny 1e-6 scale"] = (x * 1e-6, w)
    sets["huge 1e5 scale"] = (x * 1e5, w)
    for name, (a, b) in sets.items():
        ref = a.double() @ b.double().t()
        scale = float(ref.abs().max())
        y32 = a @ b.t()
        line = [f"{name:18s} cuBLAS fp32 {float((y32.double() - ref).abs().max()) / scale:.2e}"]
        for mname, m in modes:
            lib.adkf_gemm_configure(2, m)
            y = ops.gemm_tf32x3(a, False, b, False)
            torch.cuda.synchronize()
            line.append(f"{mname} {float((y.double() - ref).abs().max()) / scale:.2e}")
        print(" | ".join(line))
    # time: the long-K forward launch of the C2 step
    M = 133728
    a = torch.randn(M, K, device=dev, generator=g)
    out = torch.empty(M, N, device=dev)
    bias = torch.zeros(N, device=dev)
    for mname, m in modes:
        lib.adkf_gemm_configure(2, m)
        ms = timed(lambda: ops.gemm_tf32x3(a, False, w, False, bias, "relu", out=out))
        print(f"forward M={M} N={N} K={K} {mname}: {ms:.3f} ms = {2.0 * M * N * K / ms / 1e9:.1f} TFLOP/s")
    lib.adkf_gemm_configure(2, 3)


if __name__ == "__main__":
    main()
