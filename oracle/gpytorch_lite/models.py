"""gpytorch.models.ExactGP call semantics, dense (call sites: fs_mol/utils/gp_utils.py:7-49;
fs_mol/models/adaptive_dkt.py:174-203; bayes_opt/bo_utils.py:400-420).

train mode : model(train_x) -> prior MVN(mean(x), K(x, x))         (inputs must be the train inputs)
eval mode  : model(x_q)     -> posterior  mean = K_qs A^-1 y ,  cov = K_qq - K_qs A^-1 K_sq ,
             A = K_ss + noise I, evaluated from ONE joint kernel on cat(train, test) and sliced
             (DefaultPredictionStrategy with fast_pred_var off, Cholesky solves).
detach_test_caches(True) (the default) detaches the train-side caches (A^-1 y and the Cholesky of A);
the reference switches it off (adaptive_dkt.py:186) so that gradients reach Z_s for the hypergradient.
"""
import torch

from . import settings
from .distributions import MultivariateNormal
from .module import Module
from .utils import psd_safe_cholesky


class GP(Module):
    pass


class ExactGP(GP):
    def __init__(self, train_inputs, train_targets, likelihood):
        if train_inputs is not None and torch.is_tensor(train_inputs):
            train_inputs = (train_inputs,)
        super().__init__()
        if train_inputs is not None:
            self.train_inputs = tuple(t.unsqueeze(-1) if t.ndimension() == 1 else t for t in train_inputs)
            self.train_targets = train_targets
        else:
            self.train_inputs = None
            self.train_targets = None
        self.likelihood = likelihood

    def set_train_data(self, inputs=None, targets=None, strict=True):
        if inputs is not None:
            if torch.is_tensor(inputs):
                inputs = (inputs,)
            inputs = tuple(t.unsqueeze(-1) if t.ndimension() == 1 else t for t in inputs)
            if strict:
                for inp, old in zip(inputs, self.train_inputs or (None,)):
                    for attr in ("shape", "dtype", "device"):
                        if old is not None and getattr(inp, attr) != getattr(old, attr):
                            raise RuntimeError(f"Cannot modify {attr} of inputs")
            self.train_inputs = inputs
        if targets is not None:
            if strict and self.train_targets is not None:
                for attr in ("shape", "dtype", "device"):
                    if getattr(targets, attr) != getattr(self.train_targets, attr):
                        raise RuntimeError(f"Cannot modify {attr} of targets")
            self.train_targets = targets

    def __call__(self, *args, **kwargs):
        train_inputs = list(self.train_inputs) if self.train_inputs is not None else []
        inputs = [i.unsqueeze(-1) if i.ndimension() == 1 else i for i in args]

        if self.training:
            if self.train_inputs is None:
                raise RuntimeError("train_inputs, train_targets cannot be None in training mode.")
            res = Module.__call__(self, *inputs, **kwargs)
            return res

        # posterior mode
        train_x, test_x = train_inputs[0], inputs[0]
        n = train_x.shape[-2]
        full_inputs = torch.cat([train_x, test_x], dim=-2)
        full_output = Module.__call__(self, full_inputs, **kwargs)
        full_mean, full_covar = full_output.loc, full_output.covariance_matrix

        train_mean = full_mean[..., :n]
        test_mean = full_mean[..., n:]
        K_ss = full_covar[..., :n, :n]
        K_qs = full_covar[..., n:, :n]
        K_qq = full_covar[..., n:, n:]

        noise = self.likelihood.noise
        A = K_ss + torch.eye(n, dtype=K_ss.dtype, device=K_ss.device) * noise
        L = psd_safe_cholesky(A)
        resid = (self.train_targets - train_mean).unsqueeze(-1)
        mean_cache = torch.cholesky_solve(resid, L).squeeze(-1)
        if settings.detach_test_caches.on():
            mean_cache = mean_cache.detach()
            L_cov = L.detach()
        else:
            L_cov = L
        predictive_mean = test_mean + (K_qs @ mean_cache.unsqueeze(-1)).squeeze(-1)
        K_sq = K_qs.transpose(-1, -2)
        covar_correction_rhs = torch.cholesky_solve(K_sq, L_cov)
        predictive_covar = K_qq + K_qs @ covar_correction_rhs.mul(-1)
        return MultivariateNormal(predictive_mean, predictive_covar)


class ApproximateGP(GP):
    def __init__(self, variational_strategy):
        raise NotImplementedError("variational GP is out of scope (SURVEY.md section 2 row 2)")
