"""Stand-alone exact GPs on the sm_100a kernels -- surface of fs_mol/utils/gp_utils.py (ExactGPLayer :7-49,
batch_tanimoto_sim :106-117, TanimotoKernel :120-134, ExactTanimotoGP :137-153) and of the gpytorch objects the
reference's GP-ST / bayes_opt code touches (fs_mol/gpst_test.py:71-110; bayes_opt/bo_utils.py:400-455;
bayes_opt/predictive_performance/adkt.py:86-96):

    likelihood = GaussianLikelihood(noise_prior=...)            parameter holder; likelihood(model(x)) adds the noise
    model      = ExactTanimotoGP(X, y, likelihood, use_numeric_labels) | ExactGP(X, y, likelihood, kernel)
    mll        = ExactMarginalLogLikelihood(likelihood, model)
    model.train(); likelihood.train(); fit_gpytorch_scipy(mll)   -> on-device Newton fit from the current parameters
    model.eval();  likelihood.eval();  dist = likelihood(model(X_test)); dist.mean / .variance / .log_prob(y)

One data set = the T = 1 case of the batched kernels (ops.gp_fit / gp_predict / gp_ift); above 512 joint rows the joint
log-probability goes through the blocked Cholesky (ops.gp_joint_nll_blocked).  No gpytorch, no CPU path.
"""
import math
from typing import Optional

import torch
import torch.nn as nn

from . import gp as gpk
from . import ops
from .adaptive_dkt import GaussianLikelihoodParams, _ScaleKernel

BLOCKED_ABOVE = 512      # joint rows above which log_prob uses the blocked (multi-CTA, tensor-core) factorisation


class LogNormalPrior:
    """Only what the reference configures: loc / scale = 0.25 (adaptive_dkt.py:94-119; bo_utils.py:425-443)."""

    def __init__(self, loc: float, scale: float = gpk.PRIOR_SCALE):
        if abs(scale - gpk.PRIOR_SCALE) > 1e-12:
            raise NotImplementedError("the sm_100a fit kernel has the reference's LogNormal scale 0.25 built in")
        self.loc, self.scale = float(loc), float(scale)


class GaussianLikelihood(GaussianLikelihoodParams):
    """gpytorch.likelihoods.GaussianLikelihood(noise_prior): noise = softplus(raw_noise) + 1e-4."""

    def __init__(self, noise_prior: Optional[LogNormalPrior] = None):
        super().__init__()
        self.noise_prior = noise_prior

    def forward(self, latent: "LatentPosterior"):
        return latent.with_noise()


class _BaseKernelHolder(nn.Module):
    has_lengthscale = True

    def __init__(self):
        super().__init__()
        self.raw_lengthscale = nn.Parameter(torch.zeros(1, 1))
        self.lengthscale_prior: Optional[LogNormalPrior] = None

    @property
    def lengthscale(self):
        return torch.nn.functional.softplus(self.raw_lengthscale)

    @lengthscale.setter
    def lengthscale(self, value):
        v = torch.as_tensor(value, dtype=self.raw_lengthscale.dtype, device=self.raw_lengthscale.device)
        with torch.no_grad():
            self.raw_lengthscale.copy_((v + torch.log(-torch.expm1(-v))).expand_as(self.raw_lengthscale))

    def register_prior(self, name, prior, getter=None, setter=None):
        if name != "lengthscale_prior":
            raise NotImplementedError(name)
        self.lengthscale_prior = prior


def batch_tanimoto_sim(x1: torch.Tensor, x2: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    """(x1.x2 + eps) / (eps + |x1|^2 + |x2|^2 - x1.x2) for 2-D CUDA inputs -> [len(x1), len(x2)] (gp_utils.py:106-117).
    eps is the reference's default, compiled into the kernel."""
    if abs(eps - 1e-6) > 1e-18:
        raise NotImplementedError("eps is fixed to the reference default 1e-6 in the sm_100a kernel")
    if x1.dim() != 2 or x2.dim() != 2:
        raise NotImplementedError("only 2-D inputs (one task) are on the sm_100a path")
    n1, n2 = x1.shape[0], x2.shape[0]
    layout = ops.TaskLayout.build([n1], [n2], x1.device)
    sq = ops.gp_sqdist(torch.cat([x1, x2]).float(), layout, qq_block=False, stat=ops.STAT_TANIMOTO)
    return sq.D2.view(n1 + n2, n1 + n2)[n1:, :n1].t().contiguous()


class TanimotoKernel(nn.Module):
    """Marker for CustomKernelGP / ScaleKernel(TanimotoKernel()) (gp_utils.py:120-134); forward = batch_tanimoto_sim."""
    is_stationary = False
    has_lengthscale = False

    def forward(self, x1, x2, diag=False, **params):
        if diag:
            return torch.ones(x1.shape[0], dtype=x1.dtype, device=x1.device)
        return batch_tanimoto_sim(x1, x2)


class LatentPosterior:
    """What `model(x)` returns in eval mode: the posterior of f at x, evaluated lazily by the kernels."""

    def __init__(self, model: "ExactGP", x: torch.Tensor, noisy: bool = False):
        self.model, self.x, self.noisy = model, x, noisy
        self._mv = None

    def with_noise(self):
        return LatentPosterior(self.model, self.x, noisy=True)

    def _compute(self):
        if self._mv is None:
            m = self.model
            g, n = m._bind(self.x, qq_block=False)
            mean, var, status = g.predict(m._phi())
            self._mv = (mean[n:], var[n:], status)
        return self._mv

    @property
    def mean(self):
        return self._compute()[0]

    loc = mean

    @property
    def variance(self):
        mean, var, _ = self._compute()
        return var if self.noisy else var - self.model.likelihood.noise.detach()

    @property
    def stddev(self):
        return self.variance.sqrt()

    def log_prob(self, y: torch.Tensor) -> torch.Tensor:
        """Joint (full-covariance) log density of y under the noisy predictive, as MultivariateNormal.log_prob."""
        if not self.noisy:
            raise NotImplementedError("log_prob is provided for likelihood(model(x)), as the reference uses it")
        m = self.model
        g, n = m._bind(self.x, qq_block=True, y_query=y)
        N = n + self.x.shape[0]
        if N > BLOCKED_ABOVE:
            out, status = ops.gp_joint_nll_blocked(g.sq.D2, N, n, g.y, m.kernel, m._phi())
            return -out[2]
        aux0 = torch.zeros(1, ops.AUX, dtype=torch.float32, device=self.x.device)
        _, aux, _ = ops.gp_ift(g.layout, g.sq.D2, g.y, g.karg, g.noise_loc, 0, m._phi(), aux0, mode=1)
        return -aux[0, ops.AUX_FOUTER]


class ExactGP(nn.Module):
    """Exact GP on one data set: zero mean, ScaleKernel(base kernel) (gp_utils.py:7-49; bo_utils.py:400-420)."""

    def __init__(self, train_x, train_y, likelihood: GaussianLikelihood, kernel: str = "matern",
                 freeze_noise: bool = False):
        super().__init__()
        self.kernel = gpk.check_kernel(kernel)
        self.likelihood = likelihood
        self.covar_module = _ScaleKernel()
        self.covar_module.base_kernel = _BaseKernelHolder()
        self.freeze = ops.FREEZE_NOISE if freeze_noise else 0
        self.train_inputs = (train_x,)
        self.train_targets = train_y
        self.last_fit: Optional[gpk.FitResult] = None

    def set_train_data(self, inputs=None, targets=None, strict=False):
        if inputs is not None:
            self.train_inputs = (inputs,)
        if targets is not None:
            self.train_targets = targets

    # ---- plumbing ------------------------------------------------------------------------------------
    def _phi(self) -> torch.Tensor:
        bk = self.covar_module.base_kernel
        return torch.stack([self.likelihood.noise_covar.raw_noise.detach().reshape(()),
                            self.covar_module.raw_outputscale.detach().reshape(()),
                            bk.raw_lengthscale.detach().reshape(())]).view(1, 3).float().contiguous()

    def _write_phi(self, phi3):
        with torch.no_grad():
            self.likelihood.noise_covar.raw_noise.copy_(phi3[0:1])
            self.covar_module.raw_outputscale.copy_(phi3[1])
            self.covar_module.base_kernel.raw_lengthscale.copy_(phi3[2].view(1, 1))

    def _noise_loc(self) -> float:
        p = self.likelihood.noise_prior
        return p.loc if p is not None else float("nan")

    def _bind(self, x_query: Optional[torch.Tensor], qq_block: bool, y_query: Optional[torch.Tensor] = None):
        xs = self.train_inputs[0]
        if not xs.is_cuda:
            raise RuntimeError("adkf_ift_b200 kernels need CUDA tensors (no CPU fallback)")
        n = xs.shape[0]
        m = 0 if x_query is None else x_query.shape[0]
        Z = xs if x_query is None else torch.cat([xs, x_query.to(xs.dtype)])
        ys = self.train_targets.float()
        yq = torch.zeros(m, device=xs.device) if y_query is None else y_query.float().to(xs.device)
        layout = ops.TaskLayout.build([n], [m], xs.device)
        g = gpk.BatchedExactGP(Z.float(), torch.cat([ys, yq]), layout, self.kernel, qq_block=qq_block, freeze=self.freeze,
                               use_lengthscale_prior=self.covar_module.base_kernel.lengthscale_prior is not None)
        g.noise_loc = self._noise_loc()
        return g, n

    def fit(self, max_iter: int = 100, gtol: float = 2e-6) -> gpk.FitResult:
        """fit_gpytorch_scipy(mll): minimise -MLL (with the registered priors) from the CURRENT parameter values."""
        g, _ = self._bind(None, qq_block=False)
        aux = torch.zeros(1, ops.AUX, dtype=torch.float32, device=g.Z.device)
        lsp = self.covar_module.base_kernel.lengthscale_prior
        if lsp is not None:
            aux[0, ops.AUX_LS_LOC] = lsp.loc
        res = gpk.FitResult(*ops.gp_fit(g.layout, g.sq.D2, g.y, g.karg, float(self.likelihood.noise.item()), g.noise_loc,
                                        ls_prior=2 if (lsp is not None and g.ls_prior) else 0, init_mode=0,
                                        max_iter=max_iter, gtol=gtol, phi=self._phi(), aux=aux))
        self._write_phi(res.phi[0])
        self.last_fit = res
        return res

    def median_lengthscale_init(self) -> torch.Tensor:
        """sqrt(0.5 * lower median of the positive pairwise squared distances) (bo_utils.py:458-461)."""
        g, _ = self._bind(None, qq_block=False)
        return g.init().aux[0, ops.AUX_L0]

    def forward(self, x):
        return LatentPosterior(self, x)

    def __call__(self, x):
        if self.training:
            raise NotImplementedError("train-mode outputs are consumed by mll.fit(); call model.eval() to predict")
        return LatentPosterior(self, x)


class ExactTanimotoGP(ExactGP):
    """gp_utils.py:137-153: Tanimoto kernel, noise 0.01 (trainable) for regression, 0.1 (fixed) for classification."""

    def __init__(self, train_x, train_y, likelihood: GaussianLikelihood, use_numeric_labels: bool = False):
        super().__init__(train_x, train_y, likelihood, "tanimoto", freeze_noise=not use_numeric_labels)
        likelihood.to(train_x.device)
        likelihood.noise = gpk.noise_init(use_numeric_labels)
        likelihood.noise_covar.raw_noise.requires_grad = use_numeric_labels


class CustomKernelGP(ExactGP):
    """bo_utils.py:400-420: ExactGP with a caller-built kernel (ScaleKernel(Matern) or ScaleKernel(Tanimoto))."""

    def __init__(self, train_x, train_y, kernel: str, likelihood: GaussianLikelihood):
        super().__init__(train_x, train_y, likelihood, kernel)


class ExactMarginalLogLikelihood:
    def __init__(self, likelihood: GaussianLikelihood, model: ExactGP):
        self.likelihood, self.model = likelihood, model

    def to(self, *a, **k):
        return self

    def fit(self, **kw):
        return self.model.fit(**kw)


def fit_gpytorch_scipy(mll: ExactMarginalLogLikelihood, **_ignored):
    """botorch.optim.fit.fit_gpytorch_scipy call-site compatible: returns (mll, info)."""
    res = mll.fit()
    return mll, {"fopt": res.aux[0, ops.AUX_F], "iterations": res.aux[0, ops.AUX_ITERS], "status": res.status}
