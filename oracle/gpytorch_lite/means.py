"""gpytorch.means subset (fs_mol/utils/gp_utils.py:19)."""
import torch

from .module import Module


class ZeroMean(Module):
    def forward(self, x):
        return torch.zeros(x.shape[:-1], dtype=x.dtype, device=x.device)


class ConstantMean(Module):
    def __init__(self, prior=None, batch_shape=torch.Size(), **kwargs):
        super().__init__()
        self.register_parameter("constant", torch.nn.Parameter(torch.zeros(1)))

    def forward(self, x):
        return self.constant.expand(x.shape[:-1])
