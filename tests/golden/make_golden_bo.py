"""Generates tests/golden/c5_bayes_opt.npz: the reference's OWN bayes_opt functions -- create_gp, CustomKernelGP,
compute_median_lengthscale_init (bayes_opt/bo_utils.py:400-461), extracted from the unmodified source file by name and
executed against the restated gpytorch / botorch (oracle/shims.py) -- on a dockstring-shaped pool (config 5):
2312 molecules, fit on 64, posterior / expected improvement over the pool, joint predictive NLL on the 2248 others.
The whole bo_utils module cannot be imported here (rdkit, matplotlib, torchmetrics, ... are absent); only these
definitions are needed.  Authoring container only:

    python tests/golden/make_golden_bo.py

The pool features are NOT stored: tests regenerate them from the seed with adkf_ift_b200.synthetic.bo_pool.
"""
import ast
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import shims  # noqa: E402

REPORT = shims.install()

import botorch  # noqa: E402  (restated)
import gpytorch  # noqa: E402  (restated)
from botorch.optim.fit import fit_gpytorch_scipy  # noqa: E402
from fs_mol.utils.gp_utils import TanimotoKernel  # noqa: E402  (reference file, unmodified)

from adkf_ift_b200 import synthetic  # noqa: E402


def reference_defs(names):
    """exec the named top-level definitions of bayes_opt/bo_utils.py, source untouched."""
    path = os.path.join(shims.REFERENCE_ROOT, "bayes_opt", "bo_utils.py")
    src = open(path).read()
    tree = ast.parse(src)
    ns = {"torch": torch, "np": np, "gpytorch": gpytorch, "botorch": botorch, "TanimotoKernel": TanimotoKernel,
          "fit_gpytorch_scipy": fit_gpytorch_scipy}
    for node in tree.body:
        if isinstance(node, (ast.FunctionDef, ast.ClassDef)) and node.name in names:
            exec(compile(ast.Module([node], []), path, "exec"), ns)
    return ns


def main():
    ns = reference_defs({"create_gp", "CustomKernelGP", "compute_median_lengthscale_init"})
    create_gp = ns["create_gp"]
    out = {}
    for kernel_type, d in (("matern", 64), ("tanimoto", 256)):
        x_all, y_all = synthetic.bo_pool(seed=5, pool=2312, dim=d, binary=(kernel_type == "tanimoto"))
        rng = np.random.RandomState(0)
        train_idx = np.sort(rng.choice(np.arange(len(y_all)), size=64, replace=False))
        test_idx = np.array([i for i in range(len(y_all)) if i not in set(train_idx.tolist())])
        for dt, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
            X = torch.from_numpy(x_all).to(dt)
            Y = torch.from_numpy(y_all).to(dt)
            x_tr, y_tr, x_te, y_te = X[train_idx], Y[train_idx], X[test_idx], Y[test_idx]
            mu, sd = y_tr.mean(), y_tr.std()
            y_tr, y_te = (y_tr - mu) / sd, (y_te - mu) / sd
            torch.manual_seed(0)
            likelihood, model, mll = create_gp(x_tr, y_tr, kernel_type, torch.device("cpu"), 0.01, True)
            likelihood.to(dt); model.to(dt); mll.to(dt)
            model.train(); likelihood.train()
            phi0 = np.array([p.detach().double().reshape(-1)[0].item() for p in mll.parameters()])
            fit_gpytorch_scipy(mll)
            phi = np.array([p.detach().double().reshape(-1)[0].item() for p in mll.parameters()])
            model.eval(); likelihood.eval()
            with torch.no_grad():
                pred = likelihood(model(x_te))
                nll = (-pred.log_prob(y_te) / y_te.shape[0]).item()
                pool = likelihood(model(X))
                acq = botorch.acquisition.analytic.ExpectedImprovement(model, y_tr.min().item(), maximize=False)
                ei = torch.stack([acq(X[j:j + 1, :]).reshape(()) for j in range(0, X.shape[0], 1)])
            k = f"{kernel_type}_{tag}_"
            out.update({k + "phi0": phi0, k + "phi": phi, k + "nll_per_point": np.array(nll),
                        k + "test_mean": pred.mean.double().numpy().astype(np.float32),
                        k + "test_var": pred.variance.double().numpy().astype(np.float32),
                        k + "pool_mean": pool.mean.double().numpy().astype(np.float32),
                        k + "pool_var": pool.variance.double().numpy().astype(np.float32),
                        k + "ei": ei.double().numpy().astype(np.float32), k + "best_f": np.array(y_tr.min().item()),
                        k + "noise": np.array(likelihood.noise.item())})
            if kernel_type == "matern":
                out[k + "ls_prior_loc"] = np.array(model.covar_module.base_kernel.lengthscale_prior.loc.item())
            print(kernel_type, tag, "phi0", phi0, "phi*", phi, "nll/m", nll, "EI>0:", int((ei > 0).sum()))
        out[f"{kernel_type}_train_idx"] = train_idx
    path = os.path.join(HERE, "c5_bayes_opt.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    print(REPORT)
    main()
