"""Per-role cycle counters of the pair GEMM (development build: ADKF_NVCC_EXTRA=-DADKF_PAIR_PROFILE python -m adkf_ift_b200.build
--force).  For a few launches of the C2 step: where each warp role of a CTA spends its cycles (waiting on which barrier / working).
usage: ADKF_NVCC_EXTRA=-DADKF_PAIR_PROFILE python -m adkf_ift_b200.build --force && python tools/gemm_pair_profile.py"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    from adkf_ift_b200 import _cabi, ops
    lib = _cabi.load()
    raw = ctypes.CDLL(_cabi.LIB_PATH)
    if not hasattr(raw, "adkf_gemm_pair_profile_read"):
        raise SystemExit("library was not built with -DADKF_PAIR_PROFILE")
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    V, E = 133728, 236867
    rnd = lambda *s: torch.randn(*s, device=dev, generator=g)  # noqa: E731
    x1408, wr, br = rnd(V, 1408), rnd(768, 1408), rnd(768)
    x128, w1, b1 = rnd(V, 128), rnd(1024, 128), rnd(1024)
    dy384, wT = rnd(V, 128), rnd(128, 3072)
    XG, W2, bb = rnd(E, 256), rnd(12 * 192, 64), rnd(12 * 192)
    pre = torch.empty(E, 768, device=dev)
    dyT, xT = rnd(V, 768), rnd(V, 1408)
    cases = [
        ("fwd readout hidden  M=V N=768 K=1408 (K,K)", 44, lambda: ops.gemm_tf32x3(x1408, False, wr, False, br, "relu")),
        ("fwd readout hidden  PRE-SPLIT weight images   ", 44, lambda: ops.gemm_weight_pre(x1408, wr, False, br, "relu")),
        ("fwd boom1           M=V N=1024 K=128 (K,K)", 4, lambda: ops.gemm_tf32x3(x128, False, w1, False, b1, "leaky_relu")),
        ("dX  proj-like       M=V N=3072 K=128 (K,T)", 4, lambda: ops.gemm_tf32x3(dy384, False, wT, True)),
        ("fwd pna type0 4x    M=E N=192 K=64 batched", 2, lambda: ops.gemm_tf32x3_batched(XG, False, 0, 64, W2, False, 3 * 192, 0, E, 192, 64, 4, pre, 768, 192,
                                                                                           bias=bb, bias_stride=3 * 192)),
        ("dW  readout hidden  M=768 N=1408 K=V (T,T)", None, lambda: ops.gemm_tf32x3(dyT, True, xT, True)),
    ]
    buf = (ctypes.c_ulonglong * (160 * 32))()
    for name, nk, fn in cases:
        fn()
        torch.cuda.synchronize()
        raw.adkf_gemm_pair_profile_read(buf, 1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        raw.adkf_gemm_pair_profile_read(buf, 1)
        d = np.frombuffer(buf, dtype=np.uint64).reshape(160, 32)[:148].astype(np.float64) / 1e3
        lead, peer = d[0::2].mean(0), d[1::2].mean(0)
        print(f"{name}: {a.elapsed_time(b):.3f} ms (profiling build)   [kcycles per CTA, mean over leaders | peers]")
        print(f"   TMA       wait-slot {lead[0]:7.0f} issue {lead[1]:6.0f} total {lead[2]:7.0f} | {peer[0]:7.0f} {peer[1]:6.0f} {peer[2]:7.0f}")
        print(f"   MMA       wait-acc {lead[3]:7.0f} wait-stage {lead[4]:7.0f} wait-corr-acc {lead[5]:6.0f} issue {lead[6]:6.0f} total {lead[7]:7.0f}")
        print(f"   A split   wait-raw {lead[8 + 0]:7.0f} wait-stage {lead[9]:7.0f} work {lead[10]:7.0f} total {lead[11]:7.0f} | "
              f"{peer[8]:7.0f} {peer[9]:7.0f} {peer[10]:7.0f} {peer[11]:7.0f}")
        print(f"   B split   wait-raw {lead[12]:7.0f} wait-stage {lead[13]:7.0f} work {lead[14]:7.0f} total {lead[15]:7.0f} | "
              f"{peer[12]:7.0f} {peer[13]:7.0f} {peer[14]:7.0f} {peer[15]:7.0f}")
        print(f"   drain     wait-acc {lead[16]:7.0f} drain {lead[17]:7.0f} epilogue {lead[18]:7.0f} total {lead[19]:7.0f} | "
              f"{peer[16]:7.0f} {peer[17]:7.0f} {peer[18]:7.0f} {peer[19]:7.0f}")


if __name__ == "__main__":
    main()
