"""Parameter constraints (gpytorch.constraints subset): softplus transform + lower bound.
noise = softplus(raw)+1e-4 (GreaterThan(1e-4)); outputscale / lengthscale = softplus(raw) (Positive)."""
import math

import torch
import torch.nn.functional as F


def inv_softplus(x):
    # gpytorch.utils.transforms.inv_softplus
    return x + torch.log(-torch.expm1(-x))


class Interval(torch.nn.Module):
    enforced = True

    def __init__(self, lower_bound, upper_bound, transform=F.softplus, inv_transform=inv_softplus):
        super().__init__()
        self.lower_bound = torch.as_tensor(float(lower_bound))
        self.upper_bound = torch.as_tensor(float(upper_bound))
        self._transform = transform
        self._inv_transform = inv_transform

    def transform(self, tensor):
        return self._transform(tensor) + self.lower_bound.to(tensor)

    def inverse_transform(self, transformed):
        return self._inv_transform(transformed - self.lower_bound.to(transformed))


class GreaterThan(Interval):
    def __init__(self, lower_bound, transform=F.softplus, inv_transform=inv_softplus):
        super().__init__(lower_bound, math.inf, transform, inv_transform)


class Positive(GreaterThan):
    def __init__(self, transform=F.softplus, inv_transform=inv_softplus):
        super().__init__(0.0, transform, inv_transform)
