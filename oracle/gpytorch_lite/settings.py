"""Settings flags (restated from gpytorch.settings).  Only the ones the reference touches:
detach_test_caches (fs_mol/models/adaptive_dkt.py:186) and fast_computations
(botorch fit_gpytorch_scipy).  Dense maths here is always exact Cholesky, so the flags only
control whether prediction caches are detached from the autograd graph."""


class _Flag:
    _state = None
    _default = True

    def __init__(self, state=True, **kwargs):
        self.state = state
        self.prev = None

    @classmethod
    def on(cls):
        return cls._default if cls._state is None else cls._state

    @classmethod
    def off(cls):
        return not cls.on()

    def __enter__(self):
        self.prev = type(self)._state
        type(self)._state = self.state
        return self

    def __exit__(self, *args):
        type(self)._state = self.prev
        return False


class detach_test_caches(_Flag):
    _state = None
    _default = True


class fast_computations(_Flag):
    _state = None
    _default = True

    def __init__(self, covar_root_decomposition=True, log_prob=True, solves=True):
        super().__init__(log_prob)


class fast_pred_var(_Flag):
    _state = None
    _default = False


class max_cholesky_size:
    """gpytorch default 800: above it gpytorch switches to CG/Lanczos (stochastic).  The oracle
    always uses the exact Cholesky value ('Long-context')."""
    _value = 800

    @classmethod
    def value(cls):
        return cls._value


class cholesky_jitter:
    @staticmethod
    def value(dtype):
        import torch
        return 1e-6 if dtype == torch.float32 else 1e-8
