"""gpytorch.distributions.MultivariateNormal, dense (call sites: fs_mol/utils/gp_utils.py:49;
fs_mol/models/adaptive_dkt.py:189 `.log_prob`; fs_mol/utils/adaptive_dkt_utils.py:99-102 `.mean`)."""
import math

import torch

from .utils import psd_safe_cholesky


class MultivariateNormal:
    def __init__(self, mean, covariance_matrix, validate_args=False):
        self.loc = mean
        self._covar = covariance_matrix

    @property
    def mean(self):
        return self.loc

    @property
    def covariance_matrix(self):
        return self._covar

    @property
    def lazy_covariance_matrix(self):
        return self._covar

    @property
    def variance(self):
        # gpytorch clamps to a tiny positive minimum and warns; same here
        var = self._covar.diagonal(dim1=-1, dim2=-2)
        min_variance = 1e-6 if var.dtype == torch.float32 else 1e-10
        return var.clamp_min(min_variance) if (var < min_variance).any() else var

    @property
    def stddev(self):
        return self.variance.sqrt()

    @property
    def event_shape(self):
        return self.loc.shape[-1:]

    def confidence_region(self):
        std2 = self.stddev.mul(2)
        return self.mean.sub(std2), self.mean.add(std2)

    def log_prob(self, value):
        diff = value - self.loc
        L = psd_safe_cholesky(self._covar)
        sol = torch.cholesky_solve(diff.unsqueeze(-1), L).squeeze(-1)
        inv_quad = (diff * sol).sum(-1)
        logdet = 2.0 * torch.log(L.diagonal(dim1=-1, dim2=-2)).sum(-1)
        return -0.5 * (inv_quad + logdet + diff.size(-1) * math.log(2 * math.pi))
