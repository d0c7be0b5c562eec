"""bayes_opt on the B200 kernels -- surface of bayes_opt/bo_utils.py: create_gp (:423-455), CustomKernelGP (:400-420),
compute_median_lengthscale_init (:458-461), run_gp_ei_bo (:342-397), min_so_far (:464-472).

What changes is the inner loop of run_gp_ei_bo: the reference builds a botorch ExpectedImprovement and calls it once per
pool molecule (2312 single-point posteriors per BO iteration); here one gp_predict launch produces the posterior of the
whole pool and one adkf_ei_pool launch the acquisition values.  Index bookkeeping / random fall-backs are the reference's.
"""
from typing import List, Optional

import numpy as np
import torch

from . import gp as gpk
from . import ops
from .gp_utils import (CustomKernelGP, ExactMarginalLogLikelihood, GaussianLikelihood, LogNormalPrior,
                       fit_gpytorch_scipy)


def compute_median_lengthscale_init(gp_input: torch.Tensor) -> torch.Tensor:
    n = gp_input.shape[0]
    layout = ops.TaskLayout.build([n], [0], gp_input.device)
    g = gpk.BatchedExactGP(gp_input.float(), torch.zeros(n, device=gp_input.device), layout, "matern")
    return g.init().aux[0, ops.AUX_L0]


def create_gp(train_x, train_y, kernel_type, device, noise_init, noise_prior):
    """-> (likelihood, model, mll) exactly as bo_utils.py:423-455 configures them."""
    prior = LogNormalPrior(loc=float(np.log(noise_init)) + gpk.PRIOR_SCALE ** 2) if noise_prior else None
    likelihood = GaussianLikelihood(prior).to(device)
    likelihood.noise = noise_init
    if kernel_type not in ("tanimoto", "matern"):
        raise ValueError
    model = CustomKernelGP(train_x, train_y, kernel_type, likelihood).to(device)
    if kernel_type == "matern":
        l0 = compute_median_lengthscale_init(train_x)
        bk = model.covar_module.base_kernel
        bk.register_prior("lengthscale_prior", LogNormalPrior(loc=float(torch.log(l0).item()) + gpk.PRIOR_SCALE ** 2))
        bk.lengthscale = l0
    mll = ExactMarginalLogLikelihood(likelihood, model)
    return likelihood, model, mll


def expected_improvement_pool(model, likelihood, x_all: torch.Tensor, best_f: float,
                              queried_idx: Optional[List[int]] = None) -> torch.Tensor:
    """EI(minimise) of every row of x_all under the fitted model; queried rows get -inf (bo_utils.py:368-378)."""
    dist = likelihood(model(x_all))
    exclude = None
    if queried_idx:
        exclude = torch.zeros(x_all.shape[0], dtype=torch.uint8, device=x_all.device)
        exclude[torch.as_tensor(sorted(queried_idx), device=x_all.device)] = 1
    return ops.ei_pool(dist.mean, dist.variance, float(likelihood.noise.item()), best_f, exclude)


def _select_queries(acq: torch.Tensor, seen: List[int], pool_size: int, batch: int) -> List[int]:
    """Top-`batch` molecules by EI; positions whose EI is not positive are replaced by uniform draws from the unseen pool
    (the three cases of bo_utils.py:380-390, with the same np.random.choice calls so that seeded runs agree)."""
    positive = int((acq > 0).sum())
    if positive == 0:
        unseen = [i for i in range(pool_size) if i not in seen]
        return np.random.choice(unseen, size=batch, replace=False).tolist()
    top = torch.topk(acq, batch).indices
    if positive >= batch:
        return top.tolist()
    picked = top[:positive].tolist()
    unseen = [i for i in range(pool_size) if i not in seen + picked]
    return picked + np.random.choice(unseen, size=batch - positive, replace=False).tolist()


def run_gp_ei_bo(dataset, x_all, y_all, num_init_points, query_batch_size, num_bo_iters, kernel_type, device, init_from,
                 noise_init, noise_prior):
    """GP + expected-improvement loop that MINIMISES y over a fixed pool whose rows are sorted by y ascending, so a pool
    index is also a rank (bo_utils.py:342-397).  Returns the record of queried indices: best initial point first, then
    every batch in reverse acquisition order, as the reference logs them."""
    pool = len(dataset)
    y_std_all = (y_all - y_all.mean()) / y_all.std()          # torch unbiased std, as the reference
    seen: List[int] = np.random.choice(np.arange(init_from, pool), size=num_init_points, replace=False).tolist()
    record = [min(seen)]
    for _ in range(num_bo_iters):
        x_obs, y_obs = x_all[seen, :], y_std_all[seen]
        incumbent = float(y_obs.min().item())
        likelihood, model, mll = create_gp(x_obs, y_obs, kernel_type, device, noise_init, noise_prior)
        model.train(); likelihood.train()
        fit_gpytorch_scipy(mll)
        model.eval(); likelihood.eval()
        acq = expected_improvement_pool(model, likelihood, x_all, incumbent, seen).cpu()
        chosen = _select_queries(acq, seen, pool, query_batch_size)
        seen = list(set(seen + chosen))
        record.extend(reversed(chosen))
    return record


def min_so_far(x):
    """Running minimum (bo_utils.py:464-472)."""
    return np.minimum.accumulate(np.asarray(x)).tolist()
