"""gpytorch.priors subset: LogNormalPrior (used at fs_mol/models/adaptive_dkt.py:96,113,117;
bayes_opt/bo_utils.py:428,444)."""
import math

import torch
from torch import nn


class Prior(nn.Module):
    def log_prob(self, x):
        raise NotImplementedError


class LogNormalPrior(Prior):
    """log p(x) = -log x - log(scale) - 0.5 log(2 pi) - (log x - loc)^2 / (2 scale^2)."""

    def __init__(self, loc, scale, validate_args=None, transform=None):
        super().__init__()
        self.register_buffer("loc", torch.as_tensor(float(loc)))
        self.register_buffer("scale", torch.as_tensor(float(scale)))

    def log_prob(self, x):
        loc = self.loc.to(x)
        scale = self.scale.to(x)
        lx = torch.log(x)
        return -lx - torch.log(scale) - 0.5 * math.log(2 * math.pi) - (lx - loc) ** 2 / (2 * scale ** 2)

    @property
    def mean(self):
        """torch.distributions.LogNormal.mean = exp(loc + scale^2 / 2) (read by fs_mol/models/dkt.py:98)."""
        return torch.exp(self.loc + 0.5 * self.scale ** 2)


class NormalPrior(Prior):
    def __init__(self, loc, scale, validate_args=None, transform=None):
        super().__init__()
        self.register_buffer("loc", torch.as_tensor(float(loc)))
        self.register_buffer("scale", torch.as_tensor(float(scale)))

    def log_prob(self, x):
        loc = self.loc.to(x)
        scale = self.scale.to(x)
        return -torch.log(scale) - 0.5 * math.log(2 * math.pi) - (x - loc) ** 2 / (2 * scale ** 2)
