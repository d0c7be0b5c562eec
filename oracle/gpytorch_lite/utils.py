"""psd_safe_cholesky (gpytorch.utils.cholesky): plain attempt, then diagonal jitter
1e-6 (fp32) / 1e-8 (fp64) times 10^i for i = 0, 1, 2; NotPSDError afterwards."""
import torch

from . import settings


class NotPSDError(RuntimeError):
    pass


def psd_safe_cholesky(A, upper=False, jitter=None, max_tries=3):
    L, info = torch.linalg.cholesky_ex(A)
    if not torch.any(info):
        return L.transpose(-1, -2) if upper else L
    if torch.isnan(A).any():
        raise NotPSDError(f"cholesky_cpu: {torch.isnan(A).sum().item()} of {A.numel()} elements are NaN.")
    if jitter is None:
        jitter = settings.cholesky_jitter.value(A.dtype)
    Aprime = A.clone()
    jitter_prev = 0
    for i in range(max_tries):
        jitter_new = jitter * (10 ** i)
        Aprime.diagonal(dim1=-1, dim2=-2).add_(jitter_new - jitter_prev)
        jitter_prev = jitter_new
        L, info = torch.linalg.cholesky_ex(Aprime)
        if not torch.any(info):
            return L.transpose(-1, -2) if upper else L
    raise NotPSDError(f"Matrix not positive definite after repeatedly adding jitter up to {jitter_new:.1e}.")
