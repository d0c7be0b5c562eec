"""gpytorch.kernels subset, dense.  Call sites: fs_mol/utils/gp_utils.py:22-43 (ExactGPLayer kernel
zoo), :120-134 (TanimotoKernel subclasses Kernel), bayes_opt/bo_utils.py:435-451.

Restated from gpytorch ~1.6:
  * squared distance through the centred quadratic expansion  |a|^2 - 2 a.b + |b|^2, clamp_min(0);
    when x1 is x2 and nothing needs grad the diagonal is forced to exactly 0;
  * Matern: inputs centred by mean_rows(x1) and divided by the lengthscale, r = sqrt(clamp_min(d2, 1e-30));
  * RBF: exp(-d2 / 2) on lengthscale-scaled inputs;
  * Linear: variance * x1 x2^T; Polynomial: (x1 x2^T + offset)^power;
  * ScaleKernel multiplies by softplus(raw_outputscale).
"""
import math

import torch

from .constraints import Positive
from .module import Module


def sq_dist(x1, x2, x1_eq_x2=False):
    adjustment = x1.mean(-2, keepdim=True)
    x1 = x1 - adjustment
    x2 = x2 - adjustment
    x1_norm = x1.pow(2).sum(dim=-1, keepdim=True)
    x1_pad = torch.ones_like(x1_norm)
    no_grad = not x1.requires_grad and not x2.requires_grad
    if x1_eq_x2 and no_grad:
        x2_norm, x2_pad = x1_norm, x1_pad
    else:
        x2_norm = x2.pow(2).sum(dim=-1, keepdim=True)
        x2_pad = torch.ones_like(x2_norm)
    x1_ = torch.cat([-2.0 * x1, x1_norm, x1_pad], dim=-1)
    x2_ = torch.cat([x2, x2_pad, x2_norm], dim=-1)
    res = x1_.matmul(x2_.transpose(-2, -1))
    if x1_eq_x2 and no_grad:
        res.diagonal(dim1=-2, dim2=-1).fill_(0)
    return res.clamp_min(0)


def dist(x1, x2, x1_eq_x2=False):
    return sq_dist(x1, x2, x1_eq_x2).clamp_min(1e-30).sqrt()


class Kernel(Module):
    has_lengthscale = False
    is_stationary = False

    def __init__(self, ard_num_dims=None, batch_shape=torch.Size([]), active_dims=None,
                 lengthscale_prior=None, lengthscale_constraint=None, eps=1e-6, **kwargs):
        super().__init__()
        self.ard_num_dims = ard_num_dims
        self.eps = eps
        if self.has_lengthscale:
            lengthscale_num_dims = 1 if ard_num_dims is None else ard_num_dims
            self.register_parameter(
                "raw_lengthscale", torch.nn.Parameter(torch.zeros(1, lengthscale_num_dims)))
            self.register_constraint("raw_lengthscale", lengthscale_constraint or Positive())
            if lengthscale_prior is not None:
                self.register_prior("lengthscale_prior", lengthscale_prior,
                                    lambda m: m.lengthscale, lambda m, v: m._set_lengthscale(v))

    @property
    def lengthscale(self):
        if self.has_lengthscale:
            return self.raw_lengthscale_constraint.transform(self.raw_lengthscale)
        return None

    @lengthscale.setter
    def lengthscale(self, value):
        self._set_lengthscale(value)

    def _set_lengthscale(self, value):
        if not self.has_lengthscale:
            raise RuntimeError("Kernel has no lengthscale.")
        if not torch.is_tensor(value):
            value = torch.as_tensor(value)
        value = value.to(self.raw_lengthscale)
        self.initialize(raw_lengthscale=self.raw_lengthscale_constraint.inverse_transform(value))

    def forward(self, x1, x2, diag=False, **params):
        raise NotImplementedError

    def covar_dist(self, x1, x2, diag=False, square_dist=False, **params):
        x1_eq_x2 = torch.equal(x1, x2)
        if diag:
            if x1_eq_x2:
                return torch.zeros(*x1.shape[:-2], x1.shape[-2], dtype=x1.dtype, device=x1.device)
            res = (x1 - x2).pow(2).sum(-1)
            return res if square_dist else res.clamp_min(1e-30).sqrt()
        return sq_dist(x1, x2, x1_eq_x2) if square_dist else dist(x1, x2, x1_eq_x2)

    def __call__(self, x1, x2=None, diag=False, **params):
        if x1.ndimension() == 1:
            x1 = x1.unsqueeze(1)
        if x2 is None:
            x2 = x1
        elif x2.ndimension() == 1:
            x2 = x2.unsqueeze(1)
        return self.forward(x1, x2, diag=diag, **params)


class RBFKernel(Kernel):
    has_lengthscale = True
    is_stationary = True

    def forward(self, x1, x2, diag=False, **params):
        x1_ = x1.div(self.lengthscale)
        x2_ = x2.div(self.lengthscale)
        d2 = self.covar_dist(x1_, x2_, square_dist=True, diag=diag)
        return d2.div(-2).exp()


class MaternKernel(Kernel):
    has_lengthscale = True
    is_stationary = True

    def __init__(self, nu=2.5, **kwargs):
        if nu not in {0.5, 1.5, 2.5}:
            raise RuntimeError("nu expected to be 0.5, 1.5, or 2.5")
        super().__init__(**kwargs)
        self.nu = nu

    def forward(self, x1, x2, diag=False, **params):
        mean = x1.reshape(-1, x1.size(-1)).mean(0)[(None,) * (x1.dim() - 1)]
        x1_ = (x1 - mean).div(self.lengthscale)
        x2_ = (x2 - mean).div(self.lengthscale)
        distance = self.covar_dist(x1_, x2_, diag=diag)
        exp_component = torch.exp(-math.sqrt(self.nu * 2) * distance)
        if self.nu == 0.5:
            constant_component = 1
        elif self.nu == 1.5:
            constant_component = (math.sqrt(3) * distance).add(1)
        else:
            constant_component = (math.sqrt(5) * distance).add(1).add(5.0 / 3.0 * distance ** 2)
        return constant_component * exp_component


class LinearKernel(Kernel):
    def __init__(self, num_dimensions=None, offset_prior=None, variance_prior=None,
                 variance_constraint=None, **kwargs):
        super().__init__(**kwargs)
        self.register_parameter("raw_variance", torch.nn.Parameter(torch.zeros(1, 1)))
        self.register_constraint("raw_variance", variance_constraint or Positive())

    @property
    def variance(self):
        return self.raw_variance_constraint.transform(self.raw_variance)

    @variance.setter
    def variance(self, value):
        if not torch.is_tensor(value):
            value = torch.as_tensor(value)
        value = value.to(self.raw_variance)
        self.initialize(raw_variance=self.raw_variance_constraint.inverse_transform(value))

    def forward(self, x1, x2, diag=False, **params):
        x1_ = x1 * self.variance.sqrt()
        x2_ = x2 * self.variance.sqrt()
        if diag:
            return (x1_ * x2_).sum(-1)
        return x1_.matmul(x2_.transpose(-2, -1))


class PolynomialKernel(Kernel):
    def __init__(self, power, offset_prior=None, offset_constraint=None, **kwargs):
        super().__init__(**kwargs)
        self.power = power
        self.register_parameter("raw_offset", torch.nn.Parameter(torch.zeros(1)))
        self.register_constraint("raw_offset", offset_constraint or Positive())

    @property
    def offset(self):
        return self.raw_offset_constraint.transform(self.raw_offset)

    def forward(self, x1, x2, diag=False, **params):
        offset = self.offset.view(1, 1)
        if diag:
            return ((x1 * x2).sum(dim=-1) + self.offset).pow(self.power)
        return (x1.matmul(x2.transpose(-2, -1)) + offset).pow(self.power)


class ScaleKernel(Kernel):
    def __init__(self, base_kernel, outputscale_prior=None, outputscale_constraint=None, **kwargs):
        super().__init__(**kwargs)
        self.base_kernel = base_kernel
        self.register_parameter("raw_outputscale", torch.nn.Parameter(torch.zeros(())))
        self.register_constraint("raw_outputscale", outputscale_constraint or Positive())

    @property
    def is_stationary(self):
        return self.base_kernel.is_stationary

    @property
    def outputscale(self):
        return self.raw_outputscale_constraint.transform(self.raw_outputscale)

    @outputscale.setter
    def outputscale(self, value):
        if not torch.is_tensor(value):
            value = torch.as_tensor(value)
        value = value.to(self.raw_outputscale)
        self.initialize(raw_outputscale=self.raw_outputscale_constraint.inverse_transform(value))

    def forward(self, x1, x2, diag=False, **params):
        orig = self.base_kernel.forward(x1, x2, diag=diag, **params)
        return orig * self.outputscale
