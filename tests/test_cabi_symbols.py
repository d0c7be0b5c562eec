"""The C-ABI shared library loads and exports every symbol include/adkf_b200.h declares (no compute calls: this
runs without a GPU), and the ctypes signature table covers the header one-to-one."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "adkf_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(adkf_[a-z0-9_]+)\s*\(", txt)))


def test_library_builds_and_exports_every_header_symbol():
    from adkf_ift_b200 import build
    lib_path = build.build()
    assert os.path.exists(lib_path)
    lib = ctypes.CDLL(lib_path)
    syms = header_symbols()
    assert len(syms) >= 17
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/adkf_b200.h but not exported"


def test_ctypes_table_matches_header():
    from adkf_ift_b200 import _cabi
    assert sorted(_cabi.SIGNATURES) == header_symbols()
    lib = _cabi.load()
    assert lib.adkf_version() == 100
    assert lib.adkf_launch_count() == 0 or lib.adkf_launch_count() > 0


def test_no_cpu_fallback():
    """Product ops refuse CPU tensors instead of silently computing something else."""
    import torch
    from adkf_ift_b200 import ops
    with pytest.raises((RuntimeError, AssertionError)):
        ops.gp_sqdist(torch.zeros(4, 8), ops.TaskLayout.build([2], [2], "cpu"), True)
    with pytest.raises(RuntimeError):
        ops.pna_message_passing(torch.zeros(3, 128), torch.zeros(4, 3, 192, 64), torch.zeros(4, 3, 192),
                                ops.GraphIndex(3, 0, *(torch.zeros(1, dtype=torch.int32),) * 7))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "adkf_ift_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), fn
            assert "/root/reference" not in src, fn


def test_torch_library_registration_has_no_cpu_kernels():
    """torch.ops.adkf.* (adkf_ift_b200/torch_ops.py) are registered for the CUDA key only: a CPU tensor reaches no kernel."""
    import torch
    import adkf_ift_b200.torch_ops as tops
    assert set(tops.registered_ops()) == set(tops.OPS)
    with pytest.raises(NotImplementedError):
        torch.ops.adkf.gemm_tf32x3(torch.zeros(4, 4), False, torch.zeros(4, 4), False, None, 0)
    with pytest.raises(NotImplementedError):
        torch.ops.adkf.linear(torch.zeros(4, 4), torch.zeros(4, 4), None, 0)
