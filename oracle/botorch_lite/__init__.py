"""botorch_lite -- ORACLE / TEST INFRASTRUCTURE ONLY.

Restatement of the slice of third-party `botorch` (~0.6.x, un-pinned, environment.yml:24) on the
ADKF-IFT hot path: `botorch.optim.fit.fit_gpytorch_scipy` (call sites fs_mol/utils/adaptive_dkt_utils.py:91,
MoleculeNet/chem_lib/models/adkfift_trainer.py:175,241,276, bayes_opt/bo_utils.py:362,
bayes_opt/predictive_performance/adkt.py:88), `botorch.models.gpytorch.GPyTorchModel` (bo_utils.py:400)
and `botorch.acquisition.analytic.ExpectedImprovement` (bo_utils.py:365).  The optimiser itself is the
REAL scipy.optimize.minimize(method="L-BFGS-B").  "parity unpinned" at this third-party boundary.
"""
from . import optim, models, acquisition
