from . import gpytorch
