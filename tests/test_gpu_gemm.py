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


@pytest.mark.parametrize("M,N,K", [(1000, 3072, 128), (257, 1408, 768), (128, 128, 32), (5000, 132, 1024)])
def test_input_gradient_form_kt(cuda_device, M, N, K):
    """D[M,N] = A[M,K] . B  with B stored [K,N] (dX = dY W): the transposing splitter path for operand B."""
    from adkf_ift_b200 import ops
    g = torch.Generator().manual_seed(M * 7 + N)
    a = torch.randn(M, K, generator=g)
    b = torch.randn(K, N, generator=g) / K ** 0.5
    ref = a.double() @ b.double()
    ad, bd = a.to(cuda_device), b.to(cuda_device)
    out = ops.gemm_tf32x3(ad, False, bd, True)
    torch.cuda.synchronize()
    e, e32 = rel(out, ref), rel(ad @ bd, ref)
    print(f"KT M={M} N={N} K={K}: 3xTF32 {e:.3e} cuBLAS fp32 {e32:.3e}")
    assert e < max(2e-6, 4 * e32)


@pytest.mark.parametrize("M,N,K", [(128, 3072, 20000), (1024, 128, 33333), (768, 1408, 9000), (12 * 4, 768, 4100),
                                   (2048, 512, 320)])
def test_weight_gradient_form_tt_splitk(cuda_device, M, N, K):
    """D[M,N] = A^T B with A stored [K,M], B stored [K,N] (dW = dY^T X): both operands transposed, split-K over the
    (node) contraction with the deterministic second pass."""
    from adkf_ift_b200 import ops
    g = torch.Generator().manual_seed(M + N + K)
    a = torch.randn(K, M, generator=g)
    b = torch.randn(K, N, generator=g)
    ref = a.double().t() @ b.double()
    ad, bd = a.to(cuda_device), b.to(cuda_device)
    out = ops.gemm_tf32x3(ad, True, bd, True)
    out2 = ops.gemm_tf32x3(ad, True, bd, True)
    torch.cuda.synchronize()
    e, e32 = rel(out, ref), rel(ad.t() @ bd, ref)
    print(f"TT M={M} N={N} K={K}: 3xTF32 {e:.3e} cuBLAS fp32 {e32:.3e}")
    assert e < max(2e-6, 4 * e32)
    assert torch.equal(out, out2), "split-K reduction must be deterministic"


@pytest.mark.parametrize("M,K,N", [(1187, 768, 12), (3000, 32, 128), (257, 12, 768), (100, 8, 4)])
def test_linear_tc_tiny_layers(cuda_device, M, K, N):
    """The 12-wide attention-score layer (graph_readout.py:224-259), the 32 -> 128 input projection
    (graph_feature_extractor.py:76-98) and other narrow shapes on the tensor-core GEMM: forward, dX (K = N_layer), dW, db
    against fp64 autograd -- these used to fall back to cuBLAS SIMT."""
    from adkf_ift_b200 import ops
    g = torch.Generator().manual_seed(M + N)
    x, w, b, go = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / K ** 0.5, torch.randn(N, generator=g), torch.randn(M, N, generator=g)
    assert ops._tc_eligible(x.to(cuda_device), w.to(cuda_device))
    outs = []
    for dev, dt, fn in (("cpu", torch.float64, lambda a, ww, bb: torch.relu(torch.nn.functional.linear(a, ww, bb))),
                        (cuda_device, torch.float32, lambda a, ww, bb: ops.linear_tc(a, ww, bb, "relu"))):
        xx, ww, bb = (t.to(dev, dt).clone().requires_grad_(True) for t in (x, w, b))
        y = fn(xx, ww, bb)
        y.backward(go.to(dev, dt))
        outs.append((y.detach(), xx.grad, ww.grad, bb.grad))
    for name, a, r in zip(("y", "dx", "dw", "db"), outs[1], outs[0]):
        e = rel(a, r)
        print(f"tiny layer M={M} K={K} N={N} {name}: rel err {e:.3e}")
        assert e < 5e-6, name


@pytest.mark.parametrize("act", [None, "relu", "leaky_relu"])
def test_linear_tc_autograd(cuda_device, act):
    """linear_tc (forward + both gradient GEMMs + bias gradient) against fp64 autograd, also on a strided weight view
    (the fc[0] column block of adaptive_dkt.py:145-160)."""
    from adkf_ift_b200 import ops
    g = torch.Generator().manual_seed(5)
    M, K, N = 700, 512, 2048
    x = torch.randn(M, K, generator=g)
    wfull = torch.randn(N, K + 2048, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g)
    go = torch.randn(M, N, generator=g)

    def run(x, wfull, b, go, fn):
        x = x.clone().requires_grad_(True)
        wfull = wfull.clone().requires_grad_(True)
        b = b.clone().requires_grad_(True)
        y = fn(x, wfull[:, :K], b)
        y.backward(go)
        return y.detach(), x.grad, wfull.grad, b.grad

    def ref_fn(x, w, b):
        y = torch.nn.functional.linear(x, w, b)
        return y if act is None else (torch.relu(y) if act == "relu" else torch.nn.functional.leaky_relu(y, 0.01))

    ref = run(x.double(), wfull.double(), b.double(), go.double(), ref_fn)
    got = run(x.to(cuda_device), wfull.to(cuda_device), b.to(cuda_device), go.to(cuda_device),
              lambda x, w, b: ops.linear_tc(x, w, b, act))
    torch.cuda.synchronize()
    for name, a, r in zip(("y", "dx", "dw", "db"), got, ref):
        e = rel(a, r)
        print(f"linear_tc act={act} {name}: rel err {e:.3e}")
        assert e < 5e-6, name


def test_batched_forms_of_the_pna_message_mlp(cuda_device):
    """The three batched products the PNA message passing issues per edge type (4 towers per launch): message MLP
    (column blocks of the gathered inputs x row blocks of the stacked weights, bias), its input gradient, and its weight
    gradient with batched split-K."""
    from adkf_ift_b200 import ops
    g = torch.Generator().manual_seed(11)
    E, T = 5000, 4
    XG = torch.randn(E, T * 64, generator=g).to(cuda_device)
    W = (torch.randn(T, 3, 192, 64, generator=g) / 8).to(cuda_device)
    b = torch.randn(T, 3, 192, generator=g).to(cuda_device)
    tau = 1
    W2 = W.view(T * 3 * 192, 64)
    # forward: pre[:, k, :] = XG[:, 64k:64k+64] @ W[k][tau]^T + b[k][tau]
    pre = torch.zeros(E, T * 192, device=cuda_device)
    ops.gemm_tf32x3_batched(XG, False, 0, 64, W2[tau * 192:], False, 3 * 192, 0, E, 192, 64, T, pre, T * 192, 192,
                            bias=b.view(-1)[tau * 192:], bias_stride=3 * 192)
    ref = torch.stack([XG[:, 64 * k:64 * k + 64].double() @ W[k, tau].double().t() + b[k, tau].double() for k in range(T)], 1)
    assert rel(pre.view(E, T, 192), ref) < 2e-6
    # input gradient: dXG[:, 64k:64k+64] = dpre[:, k, :] @ W[k][tau]
    dpre = torch.randn(E, T * 192, generator=g).to(cuda_device)
    dXG = torch.zeros(E, T * 64, device=cuda_device)
    ops.gemm_tf32x3_batched(dpre, False, 0, 192, W2[tau * 192:], True, 0, 3 * 192, E, 64, 192, T, dXG, T * 64, 64)
    ref = torch.cat([dpre[:, 192 * k:192 * k + 192].double() @ W[k, tau].double() for k in range(T)], 1)
    assert rel(dXG, ref) < 2e-6
    # weight gradient: dW[k][tau] = dpre[:, k, :]^T @ XG[:, 64k:64k+64]   (split-K over the edges, batched)
    dW = torch.zeros(T, 3, 192, 64, device=cuda_device)
    out = dW.view(-1)[tau * 192 * 64:]
    ops.gemm_tf32x3_batched(dpre, True, 192, 0, XG, True, 64, 0, 192, 64, E, T, out, 64, 3 * 192 * 64, allow_splitk=True)
    torch.cuda.synchronize()
    ref = torch.stack([dpre[:, 192 * k:192 * k + 192].double().t() @ XG[:, 64 * k:64 * k + 64].double() for k in range(T)])
    assert rel(dW[:, tau], ref) < 2e-6
    assert float(dW[:, 0].abs().max()) == 0.0 and float(dW[:, 2].abs().max()) == 0.0     # other types untouched


@pytest.mark.parametrize("M,N,act", [(1000, 128, 0), (5000, 1024, 2), (133, 768, 1), (70000, 128, 2), (64, 12 * 4, 1)])
def test_act_bwd_colsum(cuda_device, M, N, act):
    """dZ = dY * act'(Y) is bit-exact against torch; the fused bias gradient matches an fp64 column sum and is
    deterministic (two-pass reduction)."""
    from adkf_ift_b200 import ops
    g = torch.Generator().manual_seed(M + N + act)
    dy = torch.randn(M, N, generator=g).to(cuda_device)
    y = torch.randn(M, N, generator=g).to(cuda_device)
    if act == 1:
        y = y.clamp_min(0)
    dz, db = ops.act_bwd_colsum(dy, y if act else None, act, True)
    dz2, db2 = ops.act_bwd_colsum(dy, y if act else None, act, True)
    want = dy if act == 0 else dy * (y > 0) if act == 1 else dy * torch.where(y > 0, 1.0, 0.01)
    assert torch.equal(dz, want)
    ref = want.double().sum(0)
    assert float((db.double() - ref).abs().max()) <= 2e-6 * float(want.double().abs().sum(0).max())
    assert torch.equal(db, db2)


@pytest.mark.parametrize("act,M,K,H,N", [("leaky_relu", 1000, 128, 1024, 128), ("relu", 333, 1408, 768, 768),
                                          ("relu", 130, 64, 36, 44)])
def test_mlp2_tc_fused_backward(cuda_device, act, M, K, H, N):
    """Linear -> act -> Linear as one autograd node (BOOM layer / hidden-layer MLPs): the input-gradient GEMM of the second
    layer applies act'(hidden) in its epilogue.  All six gradients against fp64 autograd, and equal (to rounding) to the
    unfused linear_tc chain; ragged tile edges (M, H, N not multiples of 128) included."""
    from adkf_ift_b200 import ops
    g = torch.Generator().manual_seed(9)
    x = torch.randn(M, K, generator=g)
    w1, b1 = torch.randn(H, K, generator=g) / K ** 0.5, torch.randn(H, generator=g) * 0.1
    w2, b2 = torch.randn(N, H, generator=g) / H ** 0.5, torch.randn(N, generator=g) * 0.1
    go = torch.randn(M, N, generator=g)

    def run(ts, fn):
        ts = [t.clone().requires_grad_(True) for t in ts]
        y = fn(*ts)
        y.backward(go.to(y.device, y.dtype))
        return [y.detach()] + [t.grad for t in ts]

    def ref_fn(x, w1, b1, w2, b2):
        h = torch.nn.functional.linear(x, w1, b1)
        h = torch.relu(h) if act == "relu" else torch.nn.functional.leaky_relu(h, 0.01)
        return torch.nn.functional.linear(h, w2, b2)

    ref = run([t.double() for t in (x, w1, b1, w2, b2)], ref_fn)
    dev = [t.to(cuda_device) for t in (x, w1, b1, w2, b2)]
    got = run(dev, lambda x, w1, b1, w2, b2: ops.mlp2_tc(x, w1, b1, act, w2, b2))
    unf = run(dev, lambda x, w1, b1, w2, b2: ops.linear_tc(ops.linear_tc(x, w1, b1, act), w2, b2))
    for name, a, u, r in zip(("y", "dx", "dw1", "db1", "dw2", "db2"), got, unf, ref):
        e, eu = rel(a, r), rel(u, r)
        print(f"mlp2_tc {act} {name}: rel err fused {e:.3e} unfused {eu:.3e}")
        assert e < 5e-6, name


@pytest.mark.parametrize("M,N,K", [(5000, 768, 1408), (4100, 200, 256), (2304, 1024, 384)])
def test_presplit_weight_images_equal_the_in_kernel_split(cuda_device, M, N, K):
    """adkf_gemm_presplit + adkf_gemm_tf32x3_pre (B operand = pre-split images of the weight, TMA-loaded straight into the MMA's B
    stages) against the kernel that splits B itself: same operand bits, same MMA order -> bit-identical outputs for the forward
    form (scaled-fp16 corrections), the input-gradient form (images of W^T, bf16 corrections) and the masked input gradient; the
    image cache follows the weight's version counter."""
    from adkf_ift_b200 import ops
    g = torch.Generator().manual_seed(M + N + K)
    x = torch.randn(M, K, generator=g).to(cuda_device)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(cuda_device)
    b = torch.randn(N, generator=g).to(cuda_device)
    # forward: y = relu(x w^T + b)
    plain = ops.gemm_tf32x3(x, False, w, False, b, "relu")
    pre = ops.gemm_weight_pre(x, w, False, b, "relu")
    assert torch.equal(pre, plain)
    ref = (x.double() @ w.double().t() + b.double()).clamp_min(0)
    assert rel(pre, ref) < 2e-6
    # input gradient: dx = dy w  (contraction over the weight's rows: images of w^T), plain and masked
    if N % 32 == 0 and N >= 256:
        dy = torch.randn(M, N, generator=g).to(cuda_device)
        plain = ops.gemm_tf32x3(dy, False, w, True)                  # (may run split-K at small M: same products, other summation order)
        pre = ops.gemm_weight_pre(dy, w, True)
        assert rel(pre, plain) < 1e-6
        assert rel(pre, dy.double() @ w.double()) < 4e-6
        mask = torch.randn(M, K, generator=g).to(cuda_device)
        plain = ops.gemm_tf32x3_masked(dy, w, mask, 2)
        pre = ops.gemm_weight_pre(dy, w, True, mask=mask, mask_act=2)
        assert torch.equal(pre, plain)
    # the cache is keyed by the version counter: an in-place update of the weight rebuilds the images
    hi0 = ops.weight_images(w, False)[0]
    assert ops.weight_images(w, False)[0] is hi0
    w.mul_(2.0)
    pre2 = ops.gemm_weight_pre(x, w, False, b, "relu")
    assert torch.equal(pre2, ops.gemm_tf32x3(x, False, w, False, b, "relu"))


@pytest.mark.parametrize("M", [129, 257, 1000, 4099])
@pytest.mark.parametrize("N,K", [(32, 32), (64, 100), (96, 192), (160, 64), (384, 36)])
def test_tma_store_epilogue_clipping_and_strided_outputs(cuda_device, M, N, K):
    """The TMA-store epilogue of the pair kernel (N % 32 == 0): partial row tiles are clipped by the tensor map, 32-column passes
    past N are skipped, and an output that is a column block of a wider buffer (ldd > N) leaves its neighbours untouched; K tails
    (K % 32 != 0) are zero-filled by the loads.  Checked against fp64 for plain, bias + relu and the masked form."""
    from adkf_ift_b200 import ops
    g = torch.Generator().manual_seed(1000 * M + 10 * N + K)
    a = torch.randn(M, K, generator=g).to(cuda_device)
    w = (torch.randn(N, K, generator=g) / max(K, 1) ** 0.5).to(cuda_device)
    b = torch.randn(N, generator=g).to(cuda_device)
    ref = a.double() @ w.double().t()
    y = ops.gemm_tf32x3(a, False, w, False)
    assert y.shape == (M, N) and rel(y, ref) < 2e-6
    y = ops.gemm_tf32x3(a, False, w, False, b, "relu")
    assert rel(y, (ref + b.double()).clamp_min(0)) < 2e-6
    # output = columns [32, 32 + N) of a wider buffer filled with a sentinel
    wide = torch.full((M + 3, N + 96), 7.0, device=cuda_device)
    view = wide[:M, 32:32 + N]
    ops.gemm_tf32x3(a, False, w, False, b, None, out=view)
    assert rel(view, ref + b.double()) < 2e-6
    wide[:M, 32:32 + N] = 7.0
    assert float(wide.min()) == 7.0 and float(wide.max()) == 7.0, "the epilogue wrote outside its column block / past row M"
    # masked input-gradient form (K,T): (dy @ w2) * leaky_relu'(mask), N_out = K here
    if K % 4 == 0:
        dy = torch.randn(M, N, generator=g).to(cuda_device)
        mask = torch.randn(M, K, generator=g).to(cuda_device)
        got = ops.gemm_tf32x3_masked(dy, w, mask, 2)
        want = (dy.double() @ w.double()) * torch.where(mask > 0, 1.0, 0.01).double()
        assert rel(got, want) < 4e-6
