"""fit_gpytorch_scipy restated (botorch ~0.6): minimise -mll(model(*train_inputs), train_targets).sum()
over all requires_grad parameters of `mll` (flattened, float64 x0, no bounds because every gpytorch
constraint is enforced through its softplus transform) with SciPy L-BFGS-B, jac=True, SciPy default
options (maxcor 10, ftol 2.22e-9, gtol 1e-5, maxfun/maxiter 15000, maxls 20).  Result written back in
the parameters' own dtype.  NaN objective -> (nan, nan-grad) like botorch's RuntimeError handling."""
from collections import OrderedDict

import numpy as np
import torch
from scipy.optimize import minimize


def _module_to_array(module):
    x, prop = [], OrderedDict()
    for name, p in module.named_parameters():
        if p.requires_grad:
            prop[name] = (p.shape, p.dtype, p.device)
            x.append(p.detach().reshape(-1).cpu().double().clone().numpy())
    return np.concatenate(x), prop


def _set_params_with_array(module, x, prop):
    params = OrderedDict(module.named_parameters())
    start = 0
    for name, (shape, dtype, device) in prop.items():
        n = int(np.prod(shape)) if len(shape) else 1
        vals = x[start:start + n]
        start += n
        new = torch.as_tensor(vals, dtype=dtype, device=device).reshape(tuple(shape))
        params[name].requires_grad_(True)
        params[name].data.copy_(new)
    return module


def _objective_and_grad(x, mll, prop):
    _set_params_with_array(mll, x, prop)
    train_inputs, train_targets = mll.model.train_inputs, mll.model.train_targets
    mll.zero_grad()
    try:
        output = mll.model(*train_inputs)
        loss = -mll(output, train_targets).sum()
    except RuntimeError as e:
        if "singular" in str(e) or "positive definite" in str(e) or "NaN" in str(e):
            return float("nan"), np.full_like(x, float("nan"))
        raise
    loss.backward()
    grads = []
    params = OrderedDict(mll.named_parameters())
    for name in prop:
        g = params[name].grad
        grads.append(np.zeros(params[name].numel()) if g is None
                     else g.detach().reshape(-1).cpu().double().clone().numpy())
    mll.zero_grad()
    return loss.item(), np.concatenate(grads)


def fit_gpytorch_scipy(mll, bounds=None, method="L-BFGS-B", options=None, track_iterations=False,
                       approx_mll=False, scipy_objective=None, module_to_array_func=None,
                       module_from_array_func=None):
    options = dict(options or {})
    x0, prop = _module_to_array(mll)
    x0 = x0.astype(np.float64)
    res = minimize(_objective_and_grad, x0, args=(mll, prop), bounds=None, method=method, jac=True,
                   options={k: v for k, v in options.items() if k not in ("method", "callback")})
    info = {"fopt": float(res.fun), "wall_time": None, "iterations": int(res.nit),
            "nfev": int(res.nfev), "message": str(res.message), "OptimizeResult": res}
    mll = _set_params_with_array(mll, res.x, prop)
    return mll, info
