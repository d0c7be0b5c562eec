"""Known-answer tests of the IFT hypergradient algebra -- the reference's only tests of the hot path
(/root/reference/test_hypergrad.ipynb cells 3-9, 13-25, 28-30), restated as pytest on our host-side routine."""
import numpy as np
import torch

from adkf_ift_b200.hypergradient import cauchy_hypergradient, cauchy_hypergradient_jvp

CPU = torch.device("cpu")


def test_sum_of_squares_cells_3_9():
    def f(po, pi):
        return sum((p ** 2).sum() for tup in (po, pi) for p in tup)

    torch.manual_seed(0)
    a = torch.randn(3, 4, requires_grad=True)
    b = torch.randn(5, requires_grad=True)
    for fn in (cauchy_hypergradient, cauchy_hypergradient_jvp):
        val = fn(f, f, params_outer=(a,), params_inner=(b,), device=CPU)
        assert torch.allclose(val, (a ** 2).sum() + (b ** 2).sum())
        assert torch.allclose(a.grad, 2 * a.detach())   # no coupling: correction term is zero
        assert torch.allclose(b.grad, 2 * b.detach())


def test_scalar_quadratic_cells_13_25():
    def f(po, pi):
        a, b, c = po
        (x,) = pi
        return a * (x ** 2) + b * x + c

    torch.manual_seed(1)
    for _ in range(20):
        a = torch.exp(torch.randn(1))[0].requires_grad_(True)
        b = torch.randn(1)[0].requires_grad_(True)
        c = torch.randn(1)[0].requires_grad_(True)
        x = torch.randn(1)[0].requires_grad_(True)
        with torch.no_grad():
            x.fill_(-b / 2 / a)
        for fn in (cauchy_hypergradient, cauchy_hypergradient_jvp):
            fn(f, f, params_outer=(a, b, c), params_inner=(x,), device=CPU)
            want = (float(b ** 2 / 4 / a ** 2), float(-b / 2 / a), 1.0)
            got = (float(a.grad), float(b.grad), float(c.grad))
            assert np.allclose(got, want, rtol=1e-5, atol=1e-6)
            # f == F: with the direct gradient ignored the residual hypergradient is zero
            fn(f, f, params_outer=(a, b, c), params_inner=(x,), device=CPU, ignore_direct_grad=True)
            assert np.allclose([float(a.grad), float(b.grad), float(c.grad)], 0.0, atol=1e-5)


def test_vector_quadratic_loop_cells_28_30():
    def f_inner(po, pi):
        a, b, c = po
        x1, x2 = pi
        return (torch.sum(a * (x1 ** 2) + b * x1 + c) + torch.sum(a * (x2 ** 2) + b * x2 + c)) / 2

    def f_outer(po, pi):
        a, b, c = po
        x1, x2 = pi
        return torch.sum(a * (x1 + x2)) / 2

    torch.manual_seed(2)
    for fn in (cauchy_hypergradient, cauchy_hypergradient_jvp):
        for _ in range(100):
            D = 3
            a = torch.exp(torch.randn(D)).requires_grad_(True)
            b = torch.randn(D, requires_grad=True)
            c = torch.randn(D, requires_grad=True)
            x1 = torch.randn(D, requires_grad=True)
            x2 = torch.randn(D, requires_grad=True)
            with torch.no_grad():
                x1.copy_(-b / 2 / a)
                x2.copy_(-b / 2 / a)
            fn(f_outer=f_outer, f_inner=f_inner, params_outer=(a, b, c), params_inner=(x1, x2), device=CPU)
            assert np.allclose(a.grad.numpy(), 0.0, atol=1e-5)
            assert np.allclose(b.grad.numpy(), -0.5, atol=1e-5)
            assert np.allclose(c.grad.numpy(), 0.0)      # unused in F and in the mixed term -> zeros
            assert x1.grad is not None and np.allclose(x1.grad.numpy(), a.detach().numpy() / 2)


def test_ignore_grad_correction_is_plain_gradient():
    def f_inner(po, pi):
        return ((po[0] - pi[0]) ** 2).sum()

    def f_outer(po, pi):
        return (po[0] * pi[0]).sum() + (po[0] ** 3).sum()

    t = torch.randn(4, requires_grad=True)
    p = t.detach().clone().requires_grad_(True)
    cauchy_hypergradient(f_outer, f_inner, (t,), (p,), CPU, ignore_grad_correction=True)
    assert torch.allclose(t.grad, p.detach() + 3 * t.detach() ** 2)
    # full: phi*(theta) = theta  =>  d/dtheta [theta*phi + theta^3] = 2 theta + 3 theta^2
    cauchy_hypergradient(f_outer, f_inner, (t,), (p,), CPU)
    assert torch.allclose(t.grad, 2 * t.detach() + 3 * t.detach() ** 2, atol=1e-5)
