"""Analytic ExpectedImprovement (botorch ~0.6), used one pool point at a time by
bayes_opt/bo_utils.py:365-373 with maximize=False."""
import torch
from torch.distributions import Normal


class ExpectedImprovement:
    def __init__(self, model, best_f, maximize=True, **kwargs):
        self.model = model
        self.best_f = torch.as_tensor(best_f)
        self.maximize = maximize

    def __call__(self, X):
        with torch.no_grad():
            posterior = self.model.posterior(X)
        mean = posterior.mean
        sigma = posterior.variance.clamp_min(1e-9).sqrt().view(mean.shape)
        u = (mean - self.best_f.to(mean).expand_as(mean)) / sigma
        if not self.maximize:
            u = -u
        normal = Normal(torch.zeros_like(u), torch.ones_like(u))
        ei = sigma * (torch.exp(normal.log_prob(u)) + u * normal.cdf(u))
        return ei.squeeze(-1).squeeze(-1) if ei.dim() > 1 else ei
