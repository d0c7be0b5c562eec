"""gpytorch.mlls.ExactMarginalLogLikelihood:  [log N(y; 0, K + noise I) + sum_priors log p] / n
(call sites: fs_mol/models/adaptive_dkt.py:126,176,195; bayes_opt/bo_utils.py:453)."""
from .distributions import MultivariateNormal
from .module import Module


class MarginalLogLikelihood(Module):
    def __init__(self, likelihood, model):
        super().__init__()
        self.likelihood = likelihood
        self.model = model


class ExactMarginalLogLikelihood(MarginalLogLikelihood):
    def forward(self, function_dist, target, *params):
        if not isinstance(function_dist, MultivariateNormal):
            raise RuntimeError("ExactMarginalLogLikelihood can only operate on Gaussian random variables")
        output = self.likelihood(function_dist)
        res = output.log_prob(target)
        for _, module, prior, closure, _ in self.named_priors():
            res = res + prior.log_prob(closure(module)).sum()
        num_data = function_dist.event_shape.numel()
        return res / num_data
