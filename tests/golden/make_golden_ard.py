"""Generates tests/golden/gp_ard.npz: --use-ard through the reference's OWN GP layer.

`ExactGPLayer(..., ard_num_dims=d)` (fs_mol/utils/gp_utils.py:14,27,30, unmodified, through oracle/shims.py) with the GP set up as
`ADKTModel.reinit_gp_params` / `__create_tail_GP` do for use_ard=True (fs_mol/models/adaptive_dkt.py:88-126: lengthscale [1,d]
initialised to the median heuristic, LogNormal prior summed over all d lengthscales, noise prior), evaluated by autograd in fp64
at FIXED raw parameters phi = (raw_noise, raw_outputscale, raw_lengthscale[0..d-1]):

    f = -MLL, grad f, H u for probe directions u, the DENSE inner Hessian H (cauchy_hypergradient.py:41-64),
    F = joint predictive NLL of the query set, dF/dphi, v = torch.linalg.solve(H, dF/dphi) (:136),
    dL/dZ = dF/dZ - d/dZ (v^T grad_phi f)   (SURVEY App. B),   predictive mean / variance,
    and phi*, f* of fit_gpytorch_scipy from the reference initialisation.

Cases: d = 64 (n 24, m 40) and the model's real d = 2048 (n 32, m 48: the dense Hessian is 2050 x 2050).

    python tests/golden/make_golden_ard.py            (authoring container only)
"""
import math
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import shims  # noqa: E402

REPORT = shims.install()
torch.set_default_dtype(torch.float64)

import gpytorch  # noqa: E402
from botorch.optim.fit import fit_gpytorch_scipy  # noqa: E402
from fs_mol.utils.gp_utils import ExactGPLayer  # noqa: E402  (reference file, unmodified)


def features(seed, n, m, d):
    rng = np.random.default_rng(seed)
    c = rng.normal(size=(3, d)) * 0.6
    lab = rng.integers(0, 3, size=n + m)
    X = (c[lab] + 0.5 * rng.normal(size=(n + m, d)) * rng.uniform(0.3, 1.7, size=d)).astype(np.float32)
    y = np.sign(np.array([1.0, -1.0, 0.4])[lab] + 0.4 * rng.normal(size=n + m)).astype(np.float32)
    y[y == 0] = 1.0
    return X, y


def build(zs, ys, kernel, d):
    """adaptive_dkt.py:88-126 with use_ard=True, use_lengthscale_prior=True, classification."""
    scale = 0.25
    lik = gpytorch.likelihoods.GaussianLikelihood(noise_prior=gpytorch.priors.LogNormalPrior(loc=math.log(0.1) + scale ** 2, scale=scale))
    gp = ExactGPLayer(train_x=zs, train_y=ys, likelihood=lik, kernel=kernel, ard_num_dims=d)
    mll = gpytorch.mlls.ExactMarginalLogLikelihood(lik, gp)
    d2 = torch.triu(torch.cdist(zs, zs) ** 2, diagonal=1)
    l0 = torch.sqrt(0.5 * torch.median(d2[d2 > 0.0]))
    loc = torch.log(l0).item() + scale ** 2
    bk = gp.covar_module.base_kernel
    bk.register_prior("lengthscale_prior", gpytorch.priors.LogNormalPrior(loc=loc, scale=scale), lambda m: m.lengthscale,
                      lambda m, v: m._set_lengthscale(v))
    bk.lengthscale = torch.ones_like(bk.lengthscale) * l0
    return lik, gp, mll, float(l0), loc


def params_of(mll):
    return [p for _, p in mll.named_parameters()]          # raw_noise [1], raw_outputscale [], raw_lengthscale [1,d]


def flat(ts):
    return torch.cat([t.reshape(-1) for t in ts])


def one_case(tag, seed, n, m, d, kernel, out):
    X, y = features(seed, n, m, d)
    Z, Y = torch.from_numpy(X).double(), torch.from_numpy(y).double()
    lik, gp, mll, l0, ls_loc = build(Z[:n], Y[:n], kernel, d)
    params = params_of(mll)
    phi_init = flat([p.detach() for p in params]).clone()
    # ---- the reference's inner fit from its own initialisation ----
    gp.train(); lik.train()
    fit_gpytorch_scipy(mll)
    phi_fit = flat([p.detach() for p in params]).clone()
    f_fit = float(-mll(gp(Z[:n]), Y[:n]))
    # ---- fixed phi: a perturbation of the fitted point (asymmetric lengthscales, nothing special about it) ----
    rng = np.random.default_rng(seed + 1)
    phi = phi_fit.clone()
    phi[0] += 0.05; phi[1] -= 0.1
    phi[2:] += torch.from_numpy(0.05 * rng.normal(size=d))
    with torch.no_grad():
        off = 0
        for p in params:
            p.copy_(phi[off:off + p.numel()].view(p.shape)); off += p.numel()
    zs = Z[:n].clone().requires_grad_(True)
    zq = Z[n:].clone().requires_grad_(True)

    def inner():
        gp.train(); lik.train()
        gp.set_train_data(inputs=zs, targets=Y[:n], strict=False)
        return -mll(gp(zs), gp.train_targets)

    def outer():
        gp.eval(); lik.eval()
        with gpytorch.settings.detach_test_caches(False):
            gp.set_train_data(inputs=zs, targets=Y[:n], strict=False)
            val = -lik(gp(zq)).log_prob(Y[n:])
        gp.train(); lik.train()
        return val

    f = inner()
    g = torch.autograd.grad(f, params, create_graph=True)
    gflat = flat(g)
    H = torch.stack([flat(torch.autograd.grad(gflat[i], params, retain_graph=True)) for i in range(gflat.numel())])
    F = outer()
    dF = torch.autograd.grad(F, params + [zs, zq])
    dF_phi = flat(dF[:len(params)])
    v = torch.linalg.solve(H, dF_phi)
    corr = torch.autograd.grad((gflat * v.detach()).sum(), [zs])[0]
    probes = torch.from_numpy(rng.normal(size=(3, d + 2)))
    gp.eval(); lik.eval()
    with torch.no_grad():
        gp.set_train_data(inputs=Z[:n], targets=Y[:n], strict=False)
        pred = lik(gp(Z[n:]))
        mean, var = pred.mean.clone(), pred.variance.clone()
    eig = torch.linalg.eigvalsh(0.5 * (H + H.T))
    print(tag, "d", d, "f", float(f), "F", float(F), "|g|inf", float(gflat.abs().max()), "eig(H) min/max", float(eig[0]), float(eig[-1]),
          "|v|", float(v.norm()), "f_fit", f_fit)
    out.update({f"{tag}_X": X, f"{tag}_y": y, f"{tag}_meta": np.array([n, m, d]), f"{tag}_kernel": np.array(kernel),
                f"{tag}_l0": np.array(l0), f"{tag}_ls_loc": np.array(ls_loc), f"{tag}_phi_init": phi_init.numpy(),
                f"{tag}_phi_fit": phi_fit.numpy(), f"{tag}_f_fit": np.array(f_fit), f"{tag}_phi": phi.numpy(),
                f"{tag}_f": np.array(float(f)), f"{tag}_g": gflat.detach().numpy(), f"{tag}_probes": probes.numpy(),
                f"{tag}_Hu": (H @ probes.T).T.detach().numpy(), f"{tag}_F": np.array(float(F)), f"{tag}_dF_dphi": dF_phi.numpy(),
                f"{tag}_v": v.detach().numpy(), f"{tag}_dL_dzs": (dF[len(params)] - corr).numpy().astype(np.float32),
                f"{tag}_dL_dzq": dF[len(params) + 1].numpy().astype(np.float32), f"{tag}_dF_dzs": dF[len(params)].numpy().astype(np.float32),
                f"{tag}_mean": mean.numpy(), f"{tag}_var": var.numpy(), f"{tag}_eig_minmax": np.array([float(eig[0]), float(eig[-1])])})


if __name__ == "__main__":
    print(REPORT)
    out = {}
    one_case("d64", 21, 24, 40, 64, "matern", out)
    one_case("d64rbf", 22, 24, 40, 64, "rbf", out)
    one_case("d2048", 23, 32, 48, 2048, "matern", out)
    path = os.path.join(HERE, "gp_ard.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")
