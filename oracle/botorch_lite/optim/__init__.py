from . import fit
from .fit import fit_gpytorch_scipy
