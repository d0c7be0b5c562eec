"""Generates tests/golden/gp_sklearn.npz: an INDEPENDENT pin of the GP arithmetic.

scikit-learn's GaussianProcessRegressor (fp64, its own Cholesky / kernels / gradients; nothing of this repo or of
oracle/ is imported) evaluates, at FIXED hyper-parameters,

    * the data term of the log marginal likelihood      log N(y; 0, s k(X,X) + sigma^2 I)
    * its gradient w.r.t. (log s, log l, log sigma^2)
    * the noisy predictive mean and variance on query rows

for kernel = ConstantKernel(s) * {Matern(l, nu=2.5) | RBF(l)} + WhiteKernel(sigma^2) on
    n = 16  : the reference model's own features of fixture c1_classification (d = 2048)
    n = 64  : fixture s64_rbf_l2 (d = 2048)
    n = 128, 256 : seeded clustered features (d = 96)

The tests turn the raw (softplus) parameters phi into (s, l, sigma^2) as gpytorch does
(fs_mol/utils/gp_utils.py:16-33: noise = softplus(raw)+1e-4, outputscale / lengthscale = softplus(raw)), add the
LogNormal prior terms and the 1/n of ExactMarginalLogLikelihood analytically, and compare
    CPU : oracle/gpytorch_lite (fp64) to 1e-9,        GPU : adkf_gp_fit(max_iter=0) / adkf_gp_predict to 1e-4.

    python tests/golden/make_golden_sklearn.py
"""
import os

import numpy as np
from sklearn.gaussian_process import GaussianProcessRegressor
from sklearn.gaussian_process.kernels import RBF, ConstantKernel, Matern, WhiteKernel

HERE = os.path.dirname(os.path.abspath(__file__))
NOISE_FLOOR = 1e-4


def softplus(x):
    return np.log1p(np.exp(-abs(x))) + max(x, 0.0)


def inv_softplus(v):
    return v + np.log(-np.expm1(-v))


def median_lengthscale(X):
    d2 = ((X[:, None, :] - X[None, :, :]) ** 2).sum(-1)
    v = np.sort(d2[np.triu_indices(len(X), 1)])
    v = v[v > 0]
    return float(np.sqrt(0.5 * v[(len(v) - 1) // 2]))            # torch.median = lower median


def clustered(seed, n, m, d=96):
    rng = np.random.default_rng(seed)
    c = rng.normal(size=(3, d))
    lab = rng.integers(0, 3, size=n + m)
    X = (c[lab] + 0.7 * rng.normal(size=(n + m, d))).astype(np.float32).astype(np.float64)
    y = (np.array([1.0, -1.0, 0.2])[lab] + 0.3 * rng.normal(size=n + m)).astype(np.float32).astype(np.float64)
    return X, y


def evaluate(X, y, n, family, phi_raw):
    """-> dict of sklearn's answers at the hyper-parameters gpytorch derives from the raw phi."""
    noise = softplus(phi_raw[0]) + NOISE_FLOOR
    s = softplus(phi_raw[1])
    ell = softplus(phi_raw[2])
    base = Matern(length_scale=ell, nu=2.5) if family == "matern" else RBF(length_scale=ell)
    kernel = ConstantKernel(s) * base + WhiteKernel(noise)
    gpr = GaussianProcessRegressor(kernel=kernel, alpha=0.0, optimizer=None, normalize_y=False)
    gpr.fit(X[:n], y[:n])
    theta = gpr.kernel_.theta                                     # log(s), log(l), log(sigma^2)
    assert np.allclose(np.exp(theta), [s, ell, noise])
    lml, grad = gpr.log_marginal_likelihood(theta, eval_gradient=True)
    mean, std = gpr.predict(X[n:], return_std=True)               # WhiteKernel diag included: the NOISY predictive variance
    return dict(lml=np.float64(lml), dlml_dlog=np.asarray(grad, dtype=np.float64), mean=mean, var=std ** 2,
                hyper=np.array([noise, s, ell]))


def main():
    out = {}
    cases = []
    for name in ("c1_classification", "s64_rbf_l2"):
        g = np.load(os.path.join(HERE, name + ".npz"))
        X = np.concatenate([g["z_support"], g["z_query"]]).astype(np.float64)
        y = np.concatenate([g["y_support"], g["y_query"]]).astype(np.float64)
        n = len(g["z_support"])
        cases.append((f"n{n}", X, y, n, np.asarray(g["train_phi_star"], dtype=np.float64)))
        # the big feature matrices are NOT duplicated: the tests read them from the fixture named here
        out[f"n{n}_source"] = np.array(name)
    for n, m, seed in ((128, 96, 11), (256, 64, 12)):
        X, y = clustered(seed, n, m)
        phi = np.array([inv_softplus(0.1 - NOISE_FLOOR), inv_softplus(0.8), inv_softplus(median_lengthscale(X[:n]))])
        cases.append((f"n{n}", X, y, n, phi))
        out[f"n{n}_X"] = X.astype(np.float32)
        out[f"n{n}_y"] = y.astype(np.float32)
    for tag, X, y, n, phi in cases:
        out[f"{tag}_n"] = np.array(n)
        out[f"{tag}_phi"] = phi
        for family in ("matern", "rbf"):
            r = evaluate(X, y, n, family, phi)
            for k, v in r.items():
                out[f"{tag}_{family}_{k}"] = v
            print(tag, family, "lml", r["lml"], "grad", r["dlml_dlog"], "hyper", r["hyper"])
    out["cases"] = np.array([c[0] for c in cases])
    import sklearn
    out["sklearn_version"] = np.array(sklearn.__version__)
    path = os.path.join(HERE, "gp_sklearn.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
