"""botorch.models.gpytorch.GPyTorchModel subset (bayes_opt/bo_utils.py:400-411): posterior(X) =
eval-mode model(X) without observation noise."""
import torch


class _Posterior:
    def __init__(self, mvn):
        self.mvn = mvn

    @property
    def mean(self):
        return self.mvn.mean.unsqueeze(-1)

    @property
    def variance(self):
        return self.mvn.variance.unsqueeze(-1)


class GPyTorchModel:
    _num_outputs = 1

    def __init__(self):
        pass

    def posterior(self, X, observation_noise=False, **kwargs):
        self.eval()
        mvn = self(X)
        if observation_noise:
            mvn = self.likelihood(mvn)
        return _Posterior(mvn)
