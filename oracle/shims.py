"""oracle/shims.py -- ORACLE / TEST INFRASTRUCTURE ONLY.

Makes the reference's OWN Python files importable, unmodified, in the authoring container
(SURVEY.md section 8c, App. D recipe).  Only used by oracle/ref_runner.py and tests/golden/make_golden.py,
and only where /root/reference exists -- never on the GPU box, never by the product path.

  1. stub modules for I/O-only dependencies that are never *called* on the synthetic path
     (pyprojroot.here, dpu_utils RichPath/run_and_debug, more_itertools.partition, rdkit);
  2. `torch_scatter`  -> oracle/torch_scatter_lite.py
  3. `gpytorch`       -> oracle/gpytorch_lite        `botorch` -> oracle/botorch_lite
     (a real install is preferred when importable, and cross-checked by tests);
  4. GNNConfig.__hash__ workaround for the py>=3.11 dataclass mutable-default check at
     fs_mol/modules/graph_feature_extractor.py:20-24.
"""
import importlib
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
# /root/reference in the authoring container; on the GPU box the unmodified copy staged by oracle/build_ref.py
REFERENCE_ROOT = os.environ.get("ADKF_REFERENCE_ROOT") or (
    "/root/reference" if os.path.isdir("/root/reference/fs_mol") else os.path.join(_HERE, "_ref"))


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "fs_mol"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _have(name):
    try:
        importlib.import_module(name)
        return True
    except Exception:
        return False


def install(prefer_real=True):
    """Install the stand-ins and put the reference on sys.path.  Returns a dict describing which
    third-party packages are real and which are restated."""
    if not reference_available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}; the shims are authoring-container only")
    repo_root = os.path.dirname(_HERE)
    if repo_root not in sys.path:
        sys.path.insert(0, repo_root)
    report = {}

    # -- 1. I/O-only stubs -------------------------------------------------------------
    if "pyprojroot" not in sys.modules and not _have("pyprojroot"):
        _stub("pyprojroot", here=lambda *a, **k: REFERENCE_ROOT)
    if "dpu_utils" not in sys.modules and not _have("dpu_utils"):
        class RichPath:  # never instantiated on the synthetic path
            @staticmethod
            def create(path, *a, **k):
                raise NotImplementedError("RichPath stub (I/O is out of scope)")

        def run_and_debug(fn, debug):
            return fn()

        du = _stub("dpu_utils")
        duu = _stub("dpu_utils.utils", RichPath=RichPath, run_and_debug=run_and_debug)
        dur = _stub("dpu_utils.utils.richpath", RichPath=RichPath)
        du.utils = duu
        duu.richpath = dur
    if "more_itertools" not in sys.modules and not _have("more_itertools"):
        def partition(pred, iterable):
            items = list(iterable)
            return (x for x in items if not pred(x)), (x for x in items if pred(x))
        _stub("more_itertools", partition=partition)
    if "rdkit" not in sys.modules and not _have("rdkit"):
        rd = _stub("rdkit")
        chem = _stub("rdkit.Chem")
        ds = _stub("rdkit.DataStructs")
        rd.Chem, rd.DataStructs = chem, ds
        chem.rdFingerprintGenerator = _stub("rdkit.Chem.rdFingerprintGenerator")
        chem.Descriptors = _stub("rdkit.Chem.Descriptors")

    # -- 2. torch_scatter ----------------------------------------------------------------
    if prefer_real and _have("torch_scatter"):
        report["torch_scatter"] = "real"
    else:
        sys.modules["torch_scatter"] = importlib.import_module("oracle.torch_scatter_lite")
        report["torch_scatter"] = "restated (oracle/torch_scatter_lite.py)"

    # -- 3. gpytorch / botorch -----------------------------------------------------------
    if prefer_real and _have("gpytorch") and _have("botorch"):
        report["gpytorch"] = report["botorch"] = "real"
    else:
        gl = importlib.import_module("oracle.gpytorch_lite")
        sys.modules["gpytorch"] = gl
        for sub in ("settings", "constraints", "priors", "means", "kernels", "likelihoods",
                    "distributions", "models", "mlls", "variational", "utils"):
            sys.modules["gpytorch." + sub] = getattr(gl, sub)
        bl = importlib.import_module("oracle.botorch_lite")
        sys.modules["botorch"] = bl
        sys.modules["botorch.optim"] = bl.optim
        sys.modules["botorch.optim.fit"] = bl.optim.fit
        sys.modules["botorch.models"] = bl.models
        sys.modules["botorch.models.gpytorch"] = bl.models.gpytorch
        sys.modules["botorch.acquisition"] = bl.acquisition
        sys.modules["botorch.acquisition.analytic"] = bl.acquisition.analytic
        report["gpytorch"] = "restated (oracle/gpytorch_lite)"
        report["botorch"] = "restated (oracle/botorch_lite, real scipy L-BFGS-B)"

    # -- reference on the path -----------------------------------------------------------
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)

    # -- 4. GNNConfig hash workaround ------------------------------------------------------
    g = importlib.import_module("fs_mol.modules.gnn")
    g.GNNConfig.__hash__ = lambda s: id(s)
    r = importlib.import_module("fs_mol.modules.graph_readout")  # frozen dataclass, hashable already
    del r
    return report
