"""ctypes binding of libadkf_b200.so (the C-ABI declared in include/adkf_b200.h).

PyTorch is only used here for device memory and streams: tensors are passed as raw device pointers
(`data_ptr()`), work is enqueued on torch's current CUDA stream.  There is NO CPU fallback: every wrapper
raises if a tensor is not a contiguous CUDA tensor of the declared dtype, and importing the library
fails loudly if the shared object has not been built (run `python -m adkf_ift_b200.build`).
"""
import ctypes
import os
from ctypes import c_float, c_int32, c_int64, c_uint64, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libadkf_b200.so")

# every exported symbol of include/adkf_b200.h with its ctypes signature
_P = c_void_p
SIGNATURES = {
    "adkf_version": ([], ctypes.c_int),
    "adkf_launch_count": ([], c_int64),
    "adkf_graph_csr_build": ([_P, c_int64, _P, c_int64, _P, c_int64, c_int32, _P, _P, _P, _P, _P, _P, _P, _P, _P], ctypes.c_int),
    "adkf_mp_pna_fwd": ([_P, _P, _P, _P, _P, _P, c_int32, _P, _P, _P, _P], ctypes.c_int),
    "adkf_mp_pna_bwd": ([_P, _P, _P, _P, _P, _P, _P, _P, c_int32, c_int32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P,
                         c_int32, _P], ctypes.c_int),
    "adkf_pna_gather_inputs": ([_P, _P, _P, c_int64, _P, _P], ctypes.c_int),
    "adkf_pna_aggregate_fwd": ([_P, _P, c_int32, _P, _P, _P, _P], ctypes.c_int),
    "adkf_pna_dpre": ([_P, _P, c_int32, _P, _P, _P, _P, _P, _P], ctypes.c_int),
    "adkf_pna_dh_gather": ([_P, _P, _P, c_int32, _P, _P], ctypes.c_int),
    "adkf_pna_fused_num_tiles": ([c_int64], c_int32),
    "adkf_pna_fused_max_degree": ([], c_int32),
    "adkf_pna_fused_set_debug": ([_P], None),
    "adkf_pna_fused_fwd": ([_P, _P, _P, _P, _P, _P, _P, _P, c_int32, c_int64, _P, _P, _P, _P], ctypes.c_int),
    "adkf_pna_fused_dpre": ([_P, _P, _P, _P, _P, _P, _P, _P, _P, c_int32, c_int64, _P, _P, _P, _P, _P, _P], ctypes.c_int),
    "adkf_mha_gather_inputs": ([_P, _P, _P, c_int64, _P, _P], ctypes.c_int),
    "adkf_mha_dx_gather": ([_P, _P, _P, c_int32, _P, _P], ctypes.c_int),
    "adkf_mha_attention_fwd": ([_P, _P, _P, _P, c_int32, _P, _P, _P], ctypes.c_int),
    "adkf_mha_attention_bwd": ([_P, _P, _P, _P, _P, c_int32, _P, _P, _P, _P, _P, _P], ctypes.c_int),
    "adkf_mp_aggregate_fwd": ([_P, _P, c_int32, c_int32, _P, _P, _P, _P], ctypes.c_int),
    "adkf_mp_dpre": ([_P, _P, c_int32, c_int32, _P, _P, _P, _P, _P, _P], ctypes.c_int),
    "adkf_act_bwd_colsum_ws_floats": ([c_int64, c_int32], c_int64),
    "adkf_act_bwd_colsum": ([_P, c_int64, _P, c_int64, c_int64, c_int32, c_int32, _P, c_int64, _P, _P, c_int64, _P], ctypes.c_int),
    "adkf_act_bwd_colsum_noalloc_max_floats": ([], c_int64),
    "adkf_act_bwd_colsum_noalloc": ([_P, c_int64, _P, c_int64, c_int64, c_int32, c_int32, _P, c_int64, _P, _P], ctypes.c_int),
    "adkf_rezero_bwd_ws_doubles": ([], c_int64),
    "adkf_rezero_fwd": ([_P, _P, _P, c_int64, _P, _P], ctypes.c_int),
    "adkf_rezero_bwd": ([_P, _P, _P, c_int64, _P, _P, _P, _P], ctypes.c_int),
    "adkf_pna_degree_scalers": ([_P, c_int32, _P, _P, _P], ctypes.c_int),
    "adkf_pna_combine_fwd": ([_P, _P, _P, _P, c_int64, _P, _P], ctypes.c_int),
    "adkf_pna_combine_bwd_ws_floats": ([c_int64], c_int64),
    "adkf_pna_combine_bwd": ([_P, _P, _P, c_int64, _P, _P, _P, c_int64, _P], ctypes.c_int),
    "adkf_layernorm128_fwd": ([_P, c_int64, _P, _P, c_float, _P, _P, _P, _P], ctypes.c_int),
    "adkf_layernorm128_bwd_ws_floats": ([c_int64], c_int64),
    "adkf_layernorm128_bwd": ([_P, _P, _P, _P, _P, c_int64, _P, _P, _P, _P, c_int64, _P], ctypes.c_int),
    "adkf_readout_weighted_fwd": ([_P, _P, _P, c_int32, c_int32, c_int32, c_int32, _P, _P, _P], ctypes.c_int),
    "adkf_readout_weighted_bwd": ([_P, _P, _P, c_int32, c_int32, c_int32, c_int32, _P, _P, _P, _P], ctypes.c_int),
    "adkf_readout_max_fwd": ([_P, _P, c_int32, c_int32, _P, _P, _P], ctypes.c_int),
    "adkf_readout_max_bwd": ([_P, _P, _P, c_int32, c_int32, _P, _P], ctypes.c_int),
    "adkf_ecfp_linear_fwd": ([_P, c_int32, c_int32, _P, c_int32, _P, _P], ctypes.c_int),
    "adkf_ecfp_linear_bwd": ([_P, c_int32, c_int32, _P, c_int32, _P, _P], ctypes.c_int),
    "adkf_gram_sqdist": ([_P, c_int32, c_int32, _P, _P, _P, c_int32, c_int32, _P, _P, _P, _P], ctypes.c_int),
    "adkf_gram_sqdist_tc": ([_P, c_int32, c_int32, c_int64, _P, _P, _P, c_int32, c_int32, _P, _P, _P, _P], ctypes.c_int),
    "adkf_gram_pairstat": ([c_int32, _P, c_int32, c_int32, _P, _P, _P, c_int32, c_int32, _P, _P, _P, _P], ctypes.c_int),
    "adkf_gram_pairstat_tc": ([c_int32, _P, c_int32, c_int32, c_int64, _P, _P, _P, c_int32, c_int32, _P, _P, _P, _P],
                              ctypes.c_int),
    "adkf_gram_pairstat_bwd": ([c_int32, _P, c_int32, c_int32, _P, _P, c_int32, _P, _P, c_float, _P, _P], ctypes.c_int),
    "adkf_linear_tf32x3": ([_P, c_int64, c_int32, c_int64, _P, c_int32, c_int64, _P, c_int32, _P, c_int64, _P], ctypes.c_int),
    "adkf_gemm_tf32x3": ([_P, c_int32, c_int64, _P, c_int32, c_int64, c_int64, c_int32, c_int32, _P, c_int32, _P, c_int64,
                          _P, c_int64, _P], ctypes.c_int),
    "adkf_gemm_presplit": ([_P, c_int64, c_int32, c_int64, c_int32, _P, _P, _P], ctypes.c_int),
    "adkf_gemm_tf32x3_pre": ([_P, c_int64, _P, _P, c_int32, c_int64, c_int32, c_int32, _P, c_int32, _P, c_int64, c_int32, _P, c_int64, _P],
                             ctypes.c_int),
    "adkf_gemm_tf32x3_masked": ([_P, c_int32, c_int64, _P, c_int32, c_int64, c_int64, c_int32, c_int32, _P, c_int64, c_int32, _P,
                                 c_int64, _P], ctypes.c_int),
    "adkf_gemm_tf32x3_batched": ([_P, c_int32, c_int64, c_int64, c_int64, c_int32, c_int32, _P, c_int32, c_int64, c_int64, c_int64,
                                  c_int32, c_int32, c_int64, c_int32, c_int32, c_int32, _P, c_int64, c_int32, _P, c_int64,
                                  c_int64, _P, c_int64, _P], ctypes.c_int),
    "adkf_gemm_batched_splitk_ws_floats": ([c_int64, c_int32, c_int32, c_int32], c_int64),
    "adkf_gemm_splitk_ws_floats": ([c_int64, c_int32, c_int32], c_int64),
    "adkf_gemm_configure": ([c_int32, c_int32], ctypes.c_int),
    "adkf_ard_wsqdist": ([_P, c_int32, c_int32, _P, _P, _P, c_int32, _P, c_int32, _P, _P], ctypes.c_int),
    "adkf_ard_pairquad": ([_P, c_int32, c_int32, _P, _P, _P, _P, c_int32, _P, _P], ctypes.c_int),
    "adkf_gram_sqdist_bwd": ([_P, c_int32, c_int32, _P, _P, c_int32, _P, _P, c_float, _P, _P], ctypes.c_int),
    "adkf_gp_fit": ([c_int32, _P, _P, _P, _P, _P, c_int32, c_float, c_float, c_int32, c_int32, c_int32, c_float,
                     _P, _P, _P, _P, c_int64, _P], ctypes.c_int),
    "adkf_gp_predict": ([c_int32, _P, _P, _P, _P, _P, c_int32, _P, _P, _P, _P, _P, c_int64, _P], ctypes.c_int),
    "adkf_gp_ift": ([c_int32, _P, _P, _P, _P, _P, c_int32, c_float, c_int32, c_int32, _P, _P, _P, _P, _P, c_int64,
                     _P], ctypes.c_int),
    "adkf_chol_blocked_ws_floats": ([c_int32], c_int64),
    "adkf_gp_joint_nll_blocked": ([_P, c_int32, c_int32, _P, c_int32, _P, c_float, _P, c_int64, _P, _P, _P], ctypes.c_int),
    "adkf_gp_joint_nll_blocked_ladder": ([_P, c_int32, c_int32, _P, c_int32, _P, _P, c_int64, _P, _P, _P], ctypes.c_int),
    "adkf_ei_pool": ([_P, _P, c_int32, c_float, c_float, _P, _P, _P], ctypes.c_int),
    "adkf_csr_from_edge_index": ([_P, c_int64, c_int32, c_int32, _P, _P, _P, _P, _P], ctypes.c_int),
    "adkf_embed_sum2": ([_P, c_int64, c_int32, _P, _P, _P, _P], ctypes.c_int),
    "adkf_gin_gather": ([_P, c_int32, c_int32, _P, _P, _P, _P, _P, _P, _P, _P], ctypes.c_int),
    "adkf_seg_bn_fwd": ([_P, c_int32, _P, c_int32, _P, _P, _P, _P, c_float, c_int32, c_int32, c_float, c_uint64, _P, _P, _P,
                         _P, _P], ctypes.c_int),
    "adkf_seg_bn_bwd": ([_P, _P, c_int32, _P, c_int32, _P, _P, _P, _P, c_int32, c_float, c_uint64, _P, _P, _P, _P],
                        ctypes.c_int),
    "adkf_seg_mean_fwd": ([_P, c_int32, _P, c_int32, _P, _P], ctypes.c_int),
    "adkf_seg_mean_bwd": ([_P, c_int32, _P, c_int32, _P, _P], ctypes.c_int),
    "adkf_gp_ws_floats": ([c_int32, _P, _P, c_int32], c_int64),
}

_lib = None


def load():
    """dlopen the in-tree shared library (never silently falls back)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: the CUDA extension has not been built. Run `python -m adkf_ift_b200.build` "
            "(needs nvcc; cross-compiles for sm_100a without a GPU). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (argtypes, restype) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.argtypes = argtypes
        fn.restype = restype
    _lib = lib
    return lib


_RAW_STREAM = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_GET_DEVICE = getattr(torch._C, "_cuda_getDevice", None)


def stream_ptr():
    """cudaStream_t of torch's current stream on the current device.  One C call: torch.cuda.current_stream() builds a Stream object
    through several Python layers (is_available, device-index parsing, os.environ) -- 15 us, ~400 times per training step."""
    if _RAW_STREAM is not None and _GET_DEVICE is not None:
        return c_void_p(_RAW_STREAM(_GET_DEVICE()))
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def dptr(t, dtype=None, allow_none=False):
    """Raw device pointer of a contiguous CUDA tensor (raises instead of falling back)."""
    if t is None:
        if allow_none:
            return c_void_p(0)
        raise ValueError("required tensor is None")
    if not t.is_cuda:
        raise RuntimeError("adkf_ift_b200 kernels need CUDA tensors (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("adkf_ift_b200 kernels need contiguous tensors")
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f"expected {dtype}, got {t.dtype}")
    return c_void_p(t.data_ptr())


def check(rc, what):
    if rc != 0:
        if rc < 0:
            raise ValueError(f"{what}: argument error {rc}")
        raise RuntimeError(f"{what}: CUDA launch failed with cudaError {rc}")
