"""Batched exact-GP tail over the sm_100a kernels: T tasks per launch.

Replaces, per task of the reference (all third-party maths restated in SURVEY.md App. A/B):
  reinit_gp_params + median lengthscale init        fs_mol/models/adaptive_dkt.py:88-131
  ExactGPLayer + ExactMarginalLogLikelihood          fs_mol/utils/gp_utils.py:7-49; adaptive_dkt.py:121-126
  botorch fit_gpytorch_scipy                         fs_mol/utils/adaptive_dkt_utils.py:91
  eval-mode posterior + GaussianLikelihood           adaptive_dkt.py:198-203
  the GP side of cauchy_hypergradient                fs_mol/utils/cauchy_hypergradient.py:41-161

phi = (raw_noise, raw_outputscale, raw_lengthscale) per task, in the order of ADKTModel.gp_params().
"""
import math
from dataclasses import dataclass
from typing import Optional

import torch

from . import ops

NOISE_FLOOR = 1e-4            # gpytorch GreaterThan(1e-4) on the likelihood noise
PRIOR_SCALE = 0.25            # LogNormal scale used for both priors (adaptive_dkt.py:96,111-119)
SUPPORTED_KERNELS = ("matern", "rbf", "RBF", "linear", "poli1", "poli2", "cossim", "tanimoto")


def noise_init(use_numeric_labels: bool) -> float:
    """gp_utils.py:16-17."""
    return 0.01 if use_numeric_labels else 0.1


def noise_prior_loc(use_numeric_labels: bool) -> float:
    """LogNormal loc such that the prior mode is the initial noise (adaptive_dkt.py:111-119)."""
    return math.log(noise_init(use_numeric_labels)) + PRIOR_SCALE ** 2


def check_kernel(kernel: str) -> str:
    if kernel not in SUPPORTED_KERNELS:   # gp_utils.py:44-45 raises ValueError with this message
        raise ValueError(f"[ERROR] the kernel '{kernel}' is not supported! (sm_100a kernels: {SUPPORTED_KERNELS}; there is "
                         "no gpytorch / CPU fallback)")
    return kernel


@dataclass
class FitResult:
    phi: torch.Tensor      # [T,3] raw parameters at the optimum
    aux: torch.Tensor      # [T,32] see include/adkf_b200.h ADKF_AUX_*
    status: torch.Tensor   # [T] int32 ADKF_ST_* flags


class BatchedExactGP:
    """Binds T tasks' features/labels (device) and exposes fit / predict / hypergradient, each ONE launch
    of a one-CTA-per-task kernel.  `Z` is read as-is (detach before passing if it carries a graph)."""

    def __init__(self, Z: torch.Tensor, y: torch.Tensor, layout: ops.TaskLayout, kernel: str = "matern",
                 use_numeric_labels: bool = False, use_lengthscale_prior: bool = True, qq_block: bool = False,
                 freeze: int = 0):
        self.kernel = check_kernel(kernel)
        self.stat = ops.KERNEL_STATS[kernel]
        self.karg = ops.kernel_arg(kernel, freeze)      # family id | ADKF_KERNEL_FREEZE_* bits
        self.Z, self.y, self.layout = Z.detach().contiguous(), y.contiguous(), layout
        self.noise_init = noise_init(use_numeric_labels)
        self.noise_loc = noise_prior_loc(use_numeric_labels)
        self.ls_prior = 1 if (use_lengthscale_prior and self.stat == ops.STAT_SQDIST) else 0
        self.qq_block = qq_block
        self.sq = ops.gp_sqdist(self.Z, layout, qq_block=qq_block, stat=self.stat)

    # reinit_gp_params: phi0 = (inv_softplus(noise_init - 1e-4), 0, inv_softplus(l0)), prior loc from l0
    def init(self) -> FitResult:
        return FitResult(*ops.gp_fit(self.layout, self.sq.D2, self.y, self.karg, self.noise_init, self.noise_loc,
                                     ls_prior=self.ls_prior, init_mode=1, max_iter=0))

    def fit(self, start: Optional[FitResult] = None, max_iter: int = 50, gtol: float = 2e-6) -> FitResult:
        """Inner fit of phi on the support sets.  start=None: reference initialisation."""
        if start is None:
            return FitResult(*ops.gp_fit(self.layout, self.sq.D2, self.y, self.karg, self.noise_init,
                                         self.noise_loc, ls_prior=self.ls_prior, init_mode=1, max_iter=max_iter,
                                         gtol=gtol))
        return FitResult(*ops.gp_fit(self.layout, self.sq.D2, self.y, self.karg, self.noise_init, self.noise_loc,
                                     ls_prior=2 if self.ls_prior else 0, init_mode=0, max_iter=max_iter, gtol=gtol,
                                     phi=start.phi, aux=start.aux))

    def evaluate(self, phi: torch.Tensor, aux: Optional[torch.Tensor] = None) -> FitResult:
        """f = -MLL, its gradient and Hessian at fixed phi (aux F / G / H)."""
        return FitResult(*ops.gp_fit(self.layout, self.sq.D2, self.y, self.karg, self.noise_init, self.noise_loc,
                                     ls_prior=(2 if aux is not None else 1) if self.ls_prior else 0, init_mode=0,
                                     max_iter=0, phi=phi, aux=aux))

    def predict(self, phi: torch.Tensor):
        """(mean, variance, status): rows aligned with Z; only query rows are written (likelihood noise included)."""
        return ops.gp_predict(self.layout, self.sq.D2, self.y, self.karg, phi)

    def hypergradient(self, fit: FitResult, mode: int = 0):
        """IFT hypergradient w.r.t. the features.  mode 0: dL/dZ = dF/dZ - d/deps df/dZ(phi + eps v)
        (SURVEY App. B); 1: ignore_grad_correction (dF/dZ only); 2: df/dZ of the inner loss only.
        Returns (dZ [rows,d], aux with F / dF/dphi / H / v, status)."""
        if not self.qq_block:
            raise RuntimeError("hypergradient needs the query x query block: construct with qq_block=True")
        Wm, aux, status = ops.gp_ift(self.layout, self.sq.D2, self.y, self.karg, self.noise_loc, self.ls_prior,
                                     fit.phi, fit.aux, mode=mode)
        if self.stat == ops.STAT_TANIMOTO:
            raise NotImplementedError("the Tanimoto kernel acts on fixed fingerprints: there is no feature gradient")
        dZ = ops.gp_sqdist_bwd(self.Z, self.layout, Wm, self.sq.mu, 1.0, stat=self.stat)
        return dZ, aux, status


    def inner_mixed_term(self, fit: FitResult, u: torch.Tensor):
        """d/deps [df/dZ](phi + eps u) for the inner loss f = -MLL: the product of the mixed second derivative
        d2f/dZ dphi with a raw-space direction u [T,3] (adkf_gp_ift mode 3).  Returns (dZ_mixed [rows,d], aux with
        f / grad f / H).  This is what a double backward through the inner loss needs (cauchy_hypergradient.py:120-161
        differentiates grad_phi f once more w.r.t. theta)."""
        if not self.qq_block:
            raise RuntimeError("inner_mixed_term uses the IFT workspace layout: construct with qq_block=True")
        aux = fit.aux.clone()
        aux[:, ops.AUX_V:ops.AUX_V + 3] = u.to(aux.dtype).view(-1, 3)
        Wm, aux, status = ops.gp_ift(self.layout, self.sq.D2, self.y, self.karg, self.noise_loc, self.ls_prior,
                                     fit.phi, aux, mode=3)
        return ops.gp_sqdist_bwd(self.Z, self.layout, Wm, self.sq.mu, 1.0, stat=self.stat), aux, status


class PredictiveDistribution:
    """Duck-type of the gpytorch MultivariateNormal the reference gets from `gp_likelihood(gp_model(x))`
    (adaptive_dkt.py:203): what run_on_batches and bayes_opt read."""

    def __init__(self, mean: torch.Tensor, variance: torch.Tensor, neg_log_prob_fn=None):
        self.mean, self.variance = mean, variance
        self._nlp = neg_log_prob_fn

    @property
    def loc(self):
        return self.mean

    @property
    def stddev(self):
        return self.variance.sqrt()

    def log_prob(self, y: torch.Tensor) -> torch.Tensor:
        """Joint (full-covariance) log density of the query labels, as MultivariateNormal.log_prob."""
        if self._nlp is None:
            raise RuntimeError("log_prob needs the joint query block; not available for this distribution")
        return -self._nlp(y)

    @property
    def covariance_matrix(self):
        raise NotImplementedError("the full predictive covariance is never materialised on the sm_100a path; "
                                  "use .variance or .log_prob")
