"""-m gpu tests of the tcgen05 3xTF32 GEMM (gemm_tf32x3.cu) through the C-ABI: dense-layer forward and the per-task
Gram / squared-distance epilogue.  The bar is fp32-class accuracy (the reason for the 3-pass split): errors are
compared with what torch's own fp32 matmul makes against an fp64 reference."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("M,K,N,act,bias", [(1000, 1408, 768, "relu", True), (133, 32, 128, None, False),
                                            (4096, 3072, 128, None, True), (257, 100, 12, "leaky_relu", True),
                                            (64, 96, 10, None, True), (300, 2048, 2048, "relu", True)])
def test_linear_forward_fp32_class_accuracy(cuda_device, M, K, N, act, bias):
    from adkf_ift_b200 import ops
    g = torch.Generator().manual_seed(M + K + N)
    x = torch.randn(M, K, generator=g)
    w = torch.randn(N, K, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g) if bias else None
    ref = x.double() @ w.double().t()
    if b is not None:
        ref = ref + b.double()
    if act == "relu":
        ref = ref.clamp_min(0)
    elif act == "leaky_relu":
        ref = torch.where(ref > 0, ref, 0.01 * ref)
    xd, wd = x.to(cuda_device), w.to(cuda_device)
    bd = b.to(cuda_device) if b is not None else None
    y = ops.linear_tf32x3_forward(xd, wd, bd, act)
    torch.cuda.synchronize()
    y32 = torch.nn.functional.linear(xd, wd, bd)
    if act == "relu":
        y32 = y32.clamp_min(0)
    elif act == "leaky_relu":
        y32 = torch.nn.functional.leaky_relu(y32, 0.01)
    e, e32 = rel(y, ref), rel(y32, ref)
    print(f"M={M} K={K} N={N}: 3xTF32 rel err {e:.3e}, cuBLAS fp32 {e32:.3e}")
    assert e < max(2e-6, 4 * e32)


def test_gram_tensor_core_matches_simt_and_oracle(cuda_device):
    from adkf_ift_b200 import ops
    from oracle import gp_closed_form as cf
    rng = np.random.default_rng(0)
    sizes = [(64, 256), (16, 32), (128, 192), (200, 40), (24, 7)]
    d = 2048
    Zs = []
    for (n, m) in sizes:
        c = rng.normal(size=(3, d)) * 0.3
        Zs.append((c[rng.integers(0, 3, size=n + m)] + 0.2 * rng.normal(size=(n + m, d)) + 1.5).astype(np.float32))
    layout = ops.TaskLayout.build([s[0] for s in sizes], [s[1] for s in sizes], cuda_device)
    Z = torch.from_numpy(np.concatenate(Zs)).to(cuda_device)
    for qq in (True, False):
        tc = ops.gp_sqdist(Z, layout, qq_block=qq, impl="tcgen05")
        simt = ops.gp_sqdist(Z, layout, qq_block=qq, impl="simt")
        torch.cuda.synchronize()
        off = 0
        for t, (n, m) in enumerate(sizes):
            N = n + m
            a = tc.D2[off:off + N * N].view(N, N).cpu().numpy()
            b = simt.D2[off:off + N * N].view(N, N).cpu().numpy()
            ref = cf.sq_dist_centered(Zs[t].astype(np.float64), n)
            if not qq:
                a, b, ref = a[:, :n], b[:, :n], ref[:, :n]      # only blocks touching the support set are defined
            e_tc = np.abs(a - ref).max() / ref.max()
            e_simt = np.abs(b - ref).max() / ref.max()
            print(f"task {t} N={N} qq={qq}: tcgen05 {e_tc:.2e} simt {e_simt:.2e}")
            assert e_tc < max(2e-6, 3 * e_simt)
            if qq:
                assert np.all(np.diag(a) == 0.0) and np.array_equal(a, a.T)
            off += N * N
