"""gpytorch_lite -- ORACLE / TEST INFRASTRUCTURE ONLY (never imported by the product path).

A dense-torch restatement of exactly the slice of third-party `gpytorch` (~1.6.x,
un-pinned in the reference's environment.yml:25) that ADKF-IFT's hot path touches:

  reference call sites      fs_mol/utils/gp_utils.py:7-49,106-153
                            fs_mol/models/adaptive_dkt.py:88-131,170-203
                            bayes_opt/bo_utils.py:400-461

gpytorch itself is NOT vendored under /root/reference and is not installable here
(no network), so this file restates its *published* algorithm:
softplus/GreaterThan constraints, centred squared-distance expansion, Matern-5/2,
RBF, Linear, Polynomial, ScaleKernel, ZeroMean, GaussianLikelihood(noise >= 1e-4),
LogNormalPrior, ExactMarginalLogLikelihood (priors added, divided by n), ExactGP
train/eval call semantics with the exact (Cholesky) prediction strategy,
psd_safe_cholesky jitter policy and MultivariateNormal.log_prob.

PARITY STATUS: "parity unpinned" at the gpytorch boundary -- the reference holds no
golden vectors for GP numerics.  The reference's OWN files
(adaptive_dkt.py, gp_utils.py, cauchy_hypergradient.py, ...) are executed unmodified
on top of this package by oracle/ref_runner.py to generate tests/golden/*.npz.

The same package is inserted into sys.modules as `gpytorch` by oracle/shims.py.
Everything is dtype-generic (fp32 = reference dtype, fp64 = truth for tolerance budgets).
"""
from . import settings, constraints, priors, means, kernels, likelihoods, distributions
from . import models, mlls, variational, utils
from .module import Module

__version__ = "1.6.0+lite"
