"""-m gpu test of adkf_act_bwd_colsum_noalloc (include/adkf_b200.h): the allocation-free variant of adkf_act_bwd_colsum whose
column-sum partials live in a file-scope `static __device__` array.  No counterpart in the reference,
so the checker is the existing entry point: same inputs through both, through the C-ABI, on two calls in a row
(a large problem first, then a smaller one with other data: stale partials of the first call must not leak into the second).
Collected last on purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-6          # relative; the two entry points run the same kernels with the same block plan, so they are in fact bit-identical


def _p(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None else 0)


def _run_both(lib, dY, Y, act, stream):
    """(dZ, db) of the existing entry point (caller workspace) and of the no-alloc entry point; all buffers are made by the CALLER
    before the calls."""
    M, N = dY.shape
    nws = int(lib.adkf_act_bwd_colsum_ws_floats(M, N))
    ws = torch.empty(nws, dtype=torch.float32, device=dY.device)
    outs = []
    for _ in range(2):
        dZ = torch.full_like(dY, float("nan")) if act else None
        db = torch.full((N,), float("nan"), dtype=torch.float32, device=dY.device)
        outs.append((dZ, db))
    (dZ0, db0), (dZ1, db1) = outs
    rc = lib.adkf_act_bwd_colsum(_p(dY), dY.stride(0), _p(Y), Y.stride(0) if Y is not None else 0, M, N, act,
                                 _p(dZ0), N if act else 0, _p(db0), _p(ws), nws, stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dY.device)[0]
    rc = lib.adkf_act_bwd_colsum_noalloc(_p(dY), dY.stride(0), _p(Y), Y.stride(0) if Y is not None else 0, M, N, act,
                                         _p(dZ1), N if act else 0, _p(db1), stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info(dY.device)[0]
    return (dZ0, db0), (dZ1, db1), free_before - free_after


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / max(float(b.double().abs().max()), 1e-30))


def test_noalloc_matches_existing_entry_point_on_two_calls(cuda_device):
    from adkf_ift_b200 import _cabi
    lib = _cabi.load()
    stream = _cabi.stream_ptr()
    gen = torch.Generator(device=cuda_device).manual_seed(11)
    max_floats = int(lib.adkf_act_bwd_colsum_noalloc_max_floats())
    assert max_floats == 1184 * 1024
    # call 1: many row blocks, relu mask; call 2: fewer rows / other width, leaky-relu; call 3: bias gradient only (act 0)
    cases = [(70001, 1024, 1), (4099, 128, 2), (513, 3072, 0)]
    leaked = []
    for M, N, act in cases:
        assert int(lib.adkf_act_bwd_colsum_ws_floats(M, N)) <= max_floats
        dY = torch.randn(M, N, device=cuda_device, generator=gen)
        Y = torch.randn(M, N, device=cuda_device, generator=gen) if act else None
        (dZ0, db0), (dZ1, db1), grew = _run_both(lib, dY, Y, act, stream)
        leaked.append(grew)
        assert bool(torch.isfinite(db1).all())
        e = _rel(db1, db0)
        assert e <= TOL, (M, N, act, e)
        assert torch.equal(db1, db0)                     # same kernels, same plan: bit-identical
        if act:
            assert torch.equal(dZ1, dZ0)
        # and against plain torch (fp64 column sums): the tolerance the dense-layer tests use
        neg = {0: None, 1: 0.0, 2: 0.01}[act]
        dz_ref = dY.double() if act == 0 else dY.double() * torch.where(Y > 0, 1.0, neg).double()
        assert _rel(db1, dz_ref.sum(0)) < 2e-5
    # the calls after the first one (module loaded, symbol address cached) did not take device memory
    assert all(g <= 0 for g in leaked[1:]), leaked


def test_noalloc_rejects_sizes_beyond_the_static_scratch(cuda_device):
    from adkf_ift_b200 import _cabi
    lib = _cabi.load()
    M, N = 148 * 8 * 64, 2048                            # 1184 row blocks x 2048 columns = twice the static array
    assert int(lib.adkf_act_bwd_colsum_ws_floats(M, N)) > int(lib.adkf_act_bwd_colsum_noalloc_max_floats())
    dY = torch.zeros(8, N, device=cuda_device)           # never read: the call must refuse before launching anything
    db = torch.full((N,), 7.0, device=cuda_device)
    rc = lib.adkf_act_bwd_colsum_noalloc(_p(dY), N, _p(None), 0, M, N, 0, _p(None), 0, _p(db), _cabi.stream_ptr())
    assert rc == -2                                      # ADKF_ERR_TOOLARGE
    torch.cuda.synchronize()
    assert float(db.min()) == 7.0 and float(db.max()) == 7.0
