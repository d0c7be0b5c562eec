"""gpytorch.Module subset: priors registry, constraints registry, initialize()."""
from collections import OrderedDict

import torch
from torch import nn


class Module(nn.Module):
    def __init__(self):
        super().__init__()
        self._priors = OrderedDict()
        self._constraints = OrderedDict()

    # --- constraints ------------------------------------------------------------------
    def register_constraint(self, param_name, constraint):
        self._constraints[param_name + "_constraint"] = constraint
        # expose as attribute like gpytorch does (raw_x_constraint)
        object.__setattr__(self, param_name + "_constraint", constraint)

    # --- priors -----------------------------------------------------------------------
    def register_prior(self, name, prior, param_or_closure, setting_closure=None):
        if isinstance(param_or_closure, str):
            pname = param_or_closure
            closure = lambda m, _p=pname: getattr(m, _p)  # noqa: E731
        else:
            closure = param_or_closure
        self.add_module(name, prior)
        self._priors[name] = (prior, closure, setting_closure)

    def named_priors(self, memo=None, prefix=""):
        """Yields (name, module, prior, closure, setting_closure), de-duplicating shared
        sub-modules (the likelihood is reachable from mll.likelihood and mll.model.likelihood)."""
        if memo is None:
            memo = set()
        if hasattr(self, "_priors"):
            for name, (prior, closure, inv_closure) in self._priors.items():
                if prior is not None and prior not in memo:
                    memo.add(prior)
                    full_name = ".".join([prefix, name]) if prefix else name
                    yield full_name, self, prior, closure, inv_closure
        for mname, module_ in self.named_children():
            submodule_prefix = prefix + ("." if prefix else "") + mname
            if hasattr(module_, "named_priors"):
                for item in module_.named_priors(memo, submodule_prefix):
                    yield item

    # --- initialise parameters by name ------------------------------------------------
    def initialize(self, **kwargs):
        for name, val in kwargs.items():
            if "." in name:
                mod_name, rest = name.split(".", 1)
                getattr(self, mod_name).initialize(**{rest: val})
                continue
            if name in self._parameters:
                p = self._parameters[name]
                if not torch.is_tensor(val):
                    val = torch.as_tensor(val)
                with torch.no_grad():
                    p.data.copy_(val.to(p).expand_as(p) if val.numel() == 1 else val.to(p).view_as(p))
            else:
                setattr(self, name, val)
        return self
