"""Independent pin of the GP arithmetic against scikit-learn's GaussianProcessRegressor.

tests/golden/gp_sklearn.npz (tests/golden/make_golden_sklearn.py) holds sklearn's fp64 data-term log marginal likelihood,
its gradient in log-parameter space and the noisy predictive mean / variance at fixed hyper-parameters, n in {16, 64, 128,
256}, Matern-5/2 and RBF.  Neither this repo's kernels nor oracle/ took part in producing it.

  CPU: oracle/gpytorch_lite (the restated gpytorch the reference's files run on) and oracle/gp_closed_form (the closed
       form the CUDA kernels implement), both fp64, equal sklearn to 1e-9 once the LogNormal noise prior and the 1/n of
       ExactMarginalLogLikelihood are added analytically HERE (numpy, written out below).
  GPU: adkf_gp_fit(max_iter=0) (f, grad at fixed phi) and adkf_gp_predict equal it to 1e-4 (north_star tolerance).
"""
import math
import os

import numpy as np
import pytest
import torch

TAU = 0.25            # LogNormalPrior scale (fs_mol/models/adaptive_dkt.py:111-119)
FAMILIES = ("matern", "rbf")


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def load_case(golden_dir, tag):
    g = np.load(os.path.join(golden_dir, "gp_sklearn.npz"))
    n = int(g[f"{tag}_n"])
    if f"{tag}_X" in g:
        X, y = g[f"{tag}_X"].astype(np.float64), g[f"{tag}_y"].astype(np.float64)
    else:
        src = np.load(os.path.join(golden_dir, str(g[f"{tag}_source"]) + ".npz"))
        X = np.concatenate([src["z_support"], src["z_query"]]).astype(np.float64)
        y = np.concatenate([src["y_support"], src["y_query"]]).astype(np.float64)
    return g, X, y, n


def expected_inner(g, tag, family, n, noise_loc):
    """f = -MLL and its gradient w.r.t. the RAW parameters (noise, outputscale, lengthscale), assembled from sklearn's
    data term: MLL = [lml + log LogNormal(noise; loc, TAU)] / n (gpytorch ExactMarginalLogLikelihood, no lengthscale prior)."""
    phi = g[f"{tag}_phi"]
    noise, s, ell = g[f"{tag}_{family}_hyper"]
    lml = float(g[f"{tag}_{family}_lml"])
    d_log = g[f"{tag}_{family}_dlml_dlog"]                           # d lml / d (log s, log l, log noise)
    ln = math.log(noise)
    logp = -ln - math.log(TAU) - 0.5 * math.log(2 * math.pi) - (ln - noise_loc) ** 2 / (2 * TAU * TAU)
    dlogp_dnoise = -(1.0 + (ln - noise_loc) / (TAU * TAU)) / noise
    f = -(lml + logp) / n
    sg = sigmoid(phi)                                                 # d softplus(raw) / d raw
    g_raw = np.array([
        -(d_log[2] / noise + dlogp_dnoise) * sg[0],
        -(d_log[0] / s) * sg[1],
        -(d_log[1] / ell) * sg[2],
    ]) / n
    return f, g_raw


CASES = ["n16", "n64", "n128", "n256"]


@pytest.mark.parametrize("tag", CASES)
def test_oracle_gp_equals_sklearn(golden_dir, tag):
    from oracle import adkf_oracle as orc
    from oracle import gp_closed_form as cf
    g, X, y, n = load_case(golden_dir, tag)
    noise_loc = math.log(0.1) + TAU ** 2
    # the prior is constructed in the default dtype and cast afterwards (adaptive_dkt.py:111-119 followed by .to(dtype)),
    # so the library side carries the fp32-rounded loc
    noise_loc32 = float(np.float32(noise_loc))
    for family in FAMILIES:
        f_want, g_want = expected_inner(g, tag, family, n, noise_loc32)
        # (i) gpytorch_lite driven exactly as the reference drives gpytorch (fs_mol/utils/gp_utils.py:7-49), fp64
        tail = orc.GPTail(orc.OracleConfig(gp_kernel=family, use_lengthscale_prior=False), torch.float64)
        params = tail.params()
        with torch.no_grad():
            for p, v in zip(params, g[f"{tag}_phi"]):
                p.fill_(float(v))
        zs, ys = torch.from_numpy(X[:n]), torch.from_numpy(y[:n])
        loss = orc.inner_loss(tail, params, zs, ys)
        grads = torch.autograd.grad(loss, params)
        got_g = np.array([float(x.reshape(-1)[0]) for x in grads])
        assert abs(float(loss) - f_want) < 1e-9 * abs(f_want), (tag, family, float(loss), f_want)
        assert rel(got_g, g_want) < 1e-7, (tag, family, got_g, g_want)
        out = orc.predict(tail, zs, ys, torch.from_numpy(X[n:]))
        assert rel(out.mean.numpy(), g[f"{tag}_{family}_mean"]) < 1e-8
        assert rel(out.variance.numpy(), g[f"{tag}_{family}_var"]) < 1e-8
        # (ii) the closed form the CUDA kernels implement
        f_want, g_want = expected_inner(g, tag, family, n, noise_loc)
        D2 = cf.sq_dist_centered(X, n)
        r = cf.inner_all(D2[:n, :n], y[:n], g[f"{tag}_phi"], family, noise_loc, None, want_hess=False)
        assert abs(r["f"] - f_want) < 1e-9 * abs(f_want)
        assert rel(r["g"], g_want) < 1e-7
        m, v = cf.predict(D2, n, y[:n], g[f"{tag}_phi"], family)
        assert rel(m, g[f"{tag}_{family}_mean"]) < 1e-8 and rel(v, g[f"{tag}_{family}_var"]) < 1e-8


@pytest.mark.gpu
@pytest.mark.parametrize("tag", CASES)
def test_cuda_gp_equals_sklearn(cuda_device, golden_dir, tag):
    """f, grad f at fixed phi (adkf_gp_fit with max_iter = 0) and the posterior (adkf_gp_predict) vs scikit-learn: 1e-4."""
    from adkf_ift_b200 import ops
    g, X, y, n = load_case(golden_dir, tag)
    m = len(X) - n
    noise_loc = math.log(0.1) + TAU ** 2
    layout = ops.TaskLayout.build([n], [m], cuda_device)
    Zd = torch.from_numpy(X.astype(np.float32)).to(cuda_device)
    yd = torch.from_numpy(y.astype(np.float32)).to(cuda_device)
    sq = ops.gp_sqdist(Zd, layout, qq_block=False)
    phi = torch.from_numpy(g[f"{tag}_phi"].astype(np.float32)).view(1, 3).to(cuda_device)
    for family in FAMILIES:
        f_want, g_want = expected_inner(g, tag, family, n, noise_loc)
        _, aux, st = ops.gp_fit(layout, sq.D2, yd, family, 0.1, noise_loc, ls_prior=0, init_mode=0, max_iter=0, phi=phi)
        a = aux.cpu().numpy()[0]
        assert int(st.item()) & ops.ST_NOT_PD == 0
        e_f = abs(a[ops.AUX_F] - f_want) / abs(f_want)
        e_g = rel(a[ops.AUX_G:ops.AUX_G + 3], g_want)
        mean, var, _ = ops.gp_predict(layout, sq.D2, yd, family, phi)
        e_m = rel(mean.cpu().numpy()[n:], g[f"{tag}_{family}_mean"])
        e_v = rel(var.cpu().numpy()[n:], g[f"{tag}_{family}_var"])
        print(f"sklearn pin {tag} {family}: f {e_f:.2e} grad {e_g:.2e} mean {e_m:.2e} var {e_v:.2e}")
        assert e_f < 1e-4 and e_g < 1e-4 and e_m < 1e-4 and e_v < 1e-4


@pytest.mark.gpu
def test_report_third_party_presence(cuda_device):
    """Records in the test log whether the real third-party libraries exist on the GPU box (nothing reported
    which implementation ran); when gpytorch is present the restated library is cross-checked against it."""
    import importlib
    found = {}
    for name in ("gpytorch", "botorch", "torch_scatter", "torch_geometric", "sklearn", "scipy"):
        try:
            found[name] = getattr(importlib.import_module(name), "__version__", "present")
        except Exception:
            found[name] = "absent"
    print("third-party on this box:", found)
    if found["gpytorch"] != "absent":
        import gpytorch
        from oracle.gpytorch_lite import kernels as lk
        x = torch.randn(12, 5, dtype=torch.float64)
        real = gpytorch.kernels.ScaleKernel(gpytorch.kernels.MaternKernel(nu=2.5)).double()
        lite = lk.ScaleKernel(lk.MaternKernel(nu=2.5)).double()
        a = real(x).to_dense() if hasattr(real(x), "to_dense") else real(x).evaluate()
        b = lite(x)
        b = b.to_dense() if hasattr(b, "to_dense") else (b.evaluate() if hasattr(b, "evaluate") else b)
        assert torch.allclose(a, b, atol=1e-10)
