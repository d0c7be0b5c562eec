"""-m gpu: the torch.library registration (adkf_ift_b200/torch_ops.py): torch.ops.adkf.* equal the ctypes path, `linear` is
differentiable through the dispatcher, CPU tensors reach no kernel, and the ops capture into a CUDA graph (no allocation outside
torch's caching allocator, no synchronisation)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / max(float(b.double().abs().max()), 1e-30))


def test_registered_ops_match_the_ctypes_path(cuda_device):
    import adkf_ift_b200.torch_ops as tops
    from adkf_ift_b200 import ops
    assert {"gemm_tf32x3", "gemm_tf32x3_masked", "linear", "act_bwd_colsum", "layernorm128", "pna_fused_fwd"} <= set(tops.registered_ops())
    g = torch.Generator().manual_seed(3)
    x = torch.randn(3000, 256, generator=g).to(cuda_device)
    w = (torch.randn(384, 256, generator=g) / 16).to(cuda_device)
    b = torch.randn(384, generator=g).to(cuda_device)
    y = torch.ops.adkf.gemm_tf32x3(x, False, w, False, b, 1)
    assert torch.equal(y, ops.gemm_tf32x3(x, False, w, False, b, "relu"))
    # differentiable linear through the dispatcher vs the autograd.Function of the product path vs fp64
    xs = [x.clone().requires_grad_(True) for _ in range(2)]
    ws = [w.clone().requires_grad_(True) for _ in range(2)]
    bs = [b.clone().requires_grad_(True) for _ in range(2)]
    gout = torch.randn(3000, 384, generator=g).to(cuda_device)
    torch.ops.adkf.linear(xs[0], ws[0], bs[0], 2).backward(gout)
    ops.linear_tc(xs[1], ws[1], bs[1], "leaky_relu").backward(gout)
    for a, c in zip((xs[0].grad, ws[0].grad, bs[0].grad), (xs[1].grad, ws[1].grad, bs[1].grad)):
        assert rel(a, c) < 1e-6
    xd, wd, bd = (t.detach().double().requires_grad_(True) for t in (x, w, b))
    torch.nn.functional.leaky_relu(xd @ wd.t() + bd, 0.01).backward(gout.double())
    assert rel(xs[0].grad, xd.grad) < 5e-6 and rel(ws[0].grad, wd.grad) < 5e-6 and rel(bs[0].grad, bd.grad) < 5e-6
    # CPU tensors reach no kernel
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops.adkf.gemm_tf32x3(x.cpu(), False, w.cpu(), False, None, 0)


def test_ops_capture_into_a_cuda_graph(cuda_device):
    import adkf_ift_b200.torch_ops  # noqa: F401
    g = torch.Generator().manual_seed(4)
    x = torch.randn(4096, 512, generator=g).to(cuda_device)
    w = (torch.randn(256, 512, generator=g) / 22).to(cuda_device)
    b = torch.randn(256, generator=g).to(cuda_device)
    want = torch.ops.adkf.gemm_tf32x3(x, False, w, False, b, 1)          # warm-up: function attributes, tensor-map entry point
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        torch.ops.adkf.gemm_tf32x3(x, False, w, False, b, 1)
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        y = torch.ops.adkf.gemm_tf32x3(x, False, w, False, b, 1)
        dz, db = torch.ops.adkf.act_bwd_colsum(y, y, 1, True)
    x.mul_(2.0)                                                            # new input values, same buffers
    graph.replay()
    torch.cuda.synchronize()
    want2 = torch.ops.adkf.gemm_tf32x3(x, False, w, False, b, 1)
    assert torch.equal(y, want2) and not torch.equal(want2, want)
    assert rel(db, (want2 * (want2 > 0)).double().sum(0)) < 1e-5
