"""gpytorch.likelihoods.GaussianLikelihood with HomoskedasticNoise: noise = softplus(raw_noise)+1e-4
(call sites: fs_mol/models/adaptive_dkt.py:121; fs_mol/utils/gp_utils.py:15-17; bayes_opt/bo_utils.py:432-433)."""
import torch

from .constraints import GreaterThan
from .distributions import MultivariateNormal
from .module import Module


class Likelihood(Module):
    pass


class HomoskedasticNoise(Module):
    def __init__(self, noise_prior=None, noise_constraint=None, batch_shape=torch.Size(), num_tasks=1):
        super().__init__()
        if noise_constraint is None:
            noise_constraint = GreaterThan(1e-4)
        self.register_parameter("raw_noise", torch.nn.Parameter(torch.zeros(*batch_shape, num_tasks)))
        if noise_prior is not None:
            self.register_prior("noise_prior", noise_prior, lambda m: m.noise, lambda m, v: m._set_noise(v))
        self.register_constraint("raw_noise", noise_constraint)

    @property
    def noise(self):
        return self.raw_noise_constraint.transform(self.raw_noise)

    @noise.setter
    def noise(self, value):
        self._set_noise(value)

    def _set_noise(self, value):
        if not torch.is_tensor(value):
            value = torch.as_tensor(value)
        value = value.to(self.raw_noise)
        self.initialize(raw_noise=self.raw_noise_constraint.inverse_transform(value))


class GaussianLikelihood(Likelihood):
    def __init__(self, noise_prior=None, noise_constraint=None, batch_shape=torch.Size(), **kwargs):
        super().__init__()
        self.noise_covar = HomoskedasticNoise(noise_prior=noise_prior, noise_constraint=noise_constraint,
                                              batch_shape=batch_shape)

    @property
    def noise(self):
        return self.noise_covar.noise

    @noise.setter
    def noise(self, value):
        self.noise_covar.noise = value

    @property
    def raw_noise(self):
        return self.noise_covar.raw_noise

    def marginal(self, function_dist):
        mean, covar = function_dist.mean, function_dist.covariance_matrix
        n = mean.shape[-1]
        eye = torch.eye(n, dtype=covar.dtype, device=covar.device)
        return MultivariateNormal(mean, covar + eye * self.noise)

    def forward(self, function_samples):
        raise NotImplementedError

    def __call__(self, input, *args, **kwargs):
        if isinstance(input, MultivariateNormal):
            return self.marginal(input)
        return super().__call__(input, *args, **kwargs)
