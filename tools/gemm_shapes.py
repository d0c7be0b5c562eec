"""Development probe: every tensor-core GEMM shape of one C2 meta-training step (16 tasks: V = 133,728 nodes, E = 296,084
directed edges, G = 5,120 graphs), timed alone with CUDA events.  Prints ms, fp32-equivalent TFLOP/s (2MNK), the HBM floor
(operands + result once) and the count per step, then the per-step total."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from adkf_ift_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
V, G = 133728, 5120
ET = [236867, 56256, 2961]   # edges per type (0.80 / 0.19 / 0.01 of 296,084)


def bench(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def rnd(*shape):
    return torch.randn(*shape, device=dev)


rows = []


def add(name, count, M, N, K, fn, batch=1):
    t = bench(fn)
    fl = 2.0 * M * N * K * batch
    by = 4.0 * (M * K + N * K + M * N) * batch
    rows.append((name, count, t, fl / t / 1e9, by / 6.53e9 / 1e3 * 1e3, count * t))


def fwd(name, count, M, N, K, act=None):
    x, w, b = rnd(M, K), rnd(N, K), rnd(N)
    add(f"fwd {name} M={M} N={N} K={K}", count, M, N, K, lambda: ops.gemm_tf32x3(x, False, w, False, b, act))


def dx(name, count, M, N, K):      # dX[M,N] = dY[M,K] W[K,N]
    dy, w = rnd(M, K), rnd(K, N)
    add(f"dX  {name} M={M} N={N} K={K}", count, M, N, K, lambda: ops.gemm_tf32x3(dy, False, w, True))


def dw(name, count, M, N, K):      # dW[M,N] = dY[K,M]^T X[K,N]
    dy, x = rnd(K, M), rnd(K, N)
    add(f"dW  {name} M={M} N={N} K={K}", count, M, N, K, lambda: ops.gemm_tf32x3(dy, True, x, True))


def layer(name, count, rows_, n_in, n_out, act=None, need_dx=True):
    fwd(name, count, rows_, n_out, n_in, act)
    if need_dx:
        dx(name, count, rows_, n_in, n_out)
    dw(name, count, n_out, n_in, rows_)


def pna(count):
    for t, e in enumerate(ET):
        XG, W2, b1 = rnd(e, 256), rnd(12 * 192, 64), rnd(12 * 192)
        pre = torch.empty(e, 768, device=dev)
        add(f"fwd pna type{t} 4x M={e} N=192 K=64", count, e, 192, 64,
            lambda: ops.gemm_tf32x3_batched(XG, False, 0, 64, W2[t * 192:], False, 3 * 192, 0, e, 192, 64, 4, pre, 768, 192,
                                            bias=b1[t * 192:], bias_stride=3 * 192), batch=4)
        dpre, dXG = rnd(e, 768), torch.empty(e, 256, device=dev)
        add(f"dX  pna type{t} 4x M={e} N=64 K=192", count, e, 64, 192,
            lambda: ops.gemm_tf32x3_batched(dpre, False, 0, 192, W2[t * 192:], True, 0, 3 * 192, e, 64, 192, 4, dXG, 256, 64), batch=4)
        dW = torch.zeros(4, 3, 192, 64, device=dev)
        add(f"dW  pna type{t} 4x M=192 N=64 K={e}", count, 192, 64, e,
            lambda: ops.gemm_tf32x3_batched(dpre, True, 192, 0, XG, True, 64, 0, 192, 64, e, 4, dW.view(-1)[t * 192 * 64:], 64,
                                            3 * 192 * 64, allow_splitk=True), batch=4)


pna(10)
layer("proj (degree scalers folded: [384,1024])", 10, V, 1024, 384)
layer("boom1", 10, V, 128, 1024, "leaky_relu")
layer("boom2", 10, V, 1024, 128)
layer("readout hidden", 4, V, 1408, 768, "relu")
layer("readout values", 2, V, 768, 768)
layer("readout comb", 2, G, 768, 512)
layer("readout max comb", 1, G, 1408, 512)
layer("readout final", 1, G, 1536, 512)
layer("fc0 gnn block", 1, G, 512, 2048)
layer("fc2", 1, G, 2048, 2048, None)
tot = 0.0
print(f"{'GEMM':48s} {'x':>3s} {'ms':>8s} {'TF/s':>7s} {'HBM floor ms':>12s} {'ms/step':>8s}")
for name, count, t, tf, floor, total in rows:
    tot += total
    print(f"{name:48s} {count:3d} {t:8.3f} {tf:7.1f} {floor:12.3f} {total:8.2f}")
print(f"total per step {tot:.2f} ms   (ADKF_GEMM_DEBUG={os.environ.get('ADKF_GEMM_DEBUG', '0')})")
