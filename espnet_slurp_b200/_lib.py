"""ctypes binding of libb200slu.so (the C ABI declared in include/b200slu.h).

The library is built in-tree by ``espnet_slurp_b200.build`` and loaded from this
directory.  There is no fallback: if the shared object is missing or fails to load the
import raises, and every op raises ``RuntimeError`` with the library's own message on a
non-zero status.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# B200SLU_LIB: another build of the same library (A/B measurements of compile-time tuning choices only)
LIB_PATH = os.environ.get("B200SLU_LIB") or os.path.join(HERE, "libb200slu.so")


class Epilogue(C.Structure):
    """b200slu_epilogue_t"""

    _fields_ = [
        ("bias", C.c_void_p),
        ("act", C.c_int),
        ("act_grad_out", C.c_void_p),
        ("alpha", C.c_float),
        ("dropout_p", C.c_float),
        ("dropout_seed", C.c_ulonglong),
        ("dropout_site", C.c_uint),
        ("act_grad_input", C.c_void_p),
        ("ld_act_grad_input", C.c_longlong),
        ("act_grad", C.c_int),
        ("residual", C.c_void_p),
        ("ld_residual", C.c_longlong),
        ("out_f32", C.c_int),
        ("accumulate", C.c_int),
        ("out_colsum_accum", C.c_void_p),
    ]


class GemmDesc(C.Structure):
    """b200slu_gemm_t"""

    _fields_ = [
        ("trans_a", C.c_int), ("trans_b", C.c_int),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("A", C.c_void_p), ("lda", C.c_longlong), ("a_stride_inner", C.c_longlong), ("a_stride_outer", C.c_longlong),
        ("B", C.c_void_p), ("ldb", C.c_longlong), ("b_stride_inner", C.c_longlong), ("b_stride_outer", C.c_longlong),
        ("C", C.c_void_p), ("ldc", C.c_longlong), ("c_stride_inner", C.c_longlong), ("c_stride_outer", C.c_longlong),
        ("batch_inner", C.c_int), ("batch_outer", C.c_int),
    ]


_P, _I, _L, _F, _U64, _U32 = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_ulonglong, C.c_uint

# argument types of every entry point of include/b200slu.h (return type int unless listed in _RESTYPES)
_SIGNATURES = {
    "b200slu_version": [],
    "b200slu_last_error": [],
    "b200slu_prof_reset": [],
    "b200slu_prof_enable_timing": [_I],
    "b200slu_prof_num_families": [],
    "b200slu_prof_family_name": [_I],
    "b200slu_prof_get": [_I, C.POINTER(C.c_double)],
    "b200slu_prof_total_launches": [],
    "b200slu_dropout_mask": [_F, _U64, _U32, _L, _P, _P],
    "b200slu_gemm_bf16": [_I, _I, _I, _I, _P, _L, _P, _L, _P, _L, _P, _P],
    "b200slu_gemm_bf16_ex": [_P, _P, _P],
    "b200slu_layernorm_fwd": [_P, _P, _P, _I, _I, _F, _P, _P, _P, _P],
    "b200slu_layernorm_bwd": [_P, _I, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P, _F, _P, _F, _U64, _U32, _P],
    "b200slu_cast_bf16": [_P, _P, _L, _F, _F, _U64, _U32, _P],
    "b200slu_cast_colsum_bf16": [_P, _P, _I, _I, _F, _F, _U64, _U32, _P, _P],
    "b200slu_colsum_bf16": [_P, _I, _I, _L, _P, _P],
    "b200slu_pos_bias_fwd": [_P, _L, _P, _P, _I, _I, _P, _P, _P],
    "b200slu_pos_bias_bwd": [_P, _P, _I, _I, _P, _L, _P, _P, _P],
    "b200slu_pos_bias_bwd_q": [_P, _P, _I, _I, _P, _L, _P, _P, _P, _P],
    "b200slu_embed_fwd": [_P, _P, _P, _I, _I, _I, _F, _P, _F, _U64, _U32, _P],
    "b200slu_embed_bwd": [_P, _P, _I, _I, _F, _P, _F, _U64, _U32, _P],
    "b200slu_softmax_fwd": [_P, _L, _P, _L, _P, _I, _I, _I, _I, _I, _I, _F, _P, _P, _L, _F, _U64, _U32, _P],
    "b200slu_softmax_bwd": [_P, _P, _L, _I, _I, _I, _I, _I, _F, _P, _P, _L, _F, _U64, _U32, _P],
    "b200slu_relattn_supported": [_I, _I, _I, _I],
    "b200slu_relattn_fwd": [_P, _P, _L, _P, _P, _L, _P, _L, _P, _I, _I, _I, _I, _I, _I, _F, _P, _P, _L, _P, _L, _F, _U64, _U32, _P],
    "b200slu_relattn_bwd_supported": [_I, _I, _I, _I],
    "b200slu_relattn_bwd": [_P, _L, _P, _L, _P, _P, _L, _P, _L, _P, _L, _I, _I, _I, _I, _I, _F, _P, _L, _P, _P, _L, _P, _F, _U64, _U32,
                            _P, _P, _L, _P, _P, _L, _P],
    "b200slu_relattn_chunk_supported": [_I, _I, _I, _I],
    "b200slu_relattn_fwd_chunk": [_P, _P, _L, _P, _P, _L, _P, _L, _P, _I, _I, _I, _I, _I, _I, _F, _P, _L, _P, _L, _F, _U64, _U32,
                                  _I, _I, _P, _P],
    "b200slu_relattn_combine": [_P, _L, _I, _I, _I, _P, _P, _L, _L, _P, _L, _I, _I, _I, _P],
    "b200slu_relattn_bwd_chunks": [_P, _L, _P, _L, _P, _P, _L, _P, _L, _P, _L, _I, _I, _I, _I, _I, _I, _F, _P, _L, _L, _P, _P, _L, _F,
                                   _U64, _U32, _P, _P, _L, _P, _P, _L, _I, _I, _P, _P, _P],
    "b200slu_sum_chunks_bf16": [_P, _L, _L, _I, _L, _I, _P, _L, _P],
    "b200slu_rel_unshift": [_P, _L, _I, _I, _I, _I, _P, _L, _P],
    "b200slu_rel_shift_index": [_I, _I, _I, _I],
    "b200slu_glu_dwconv_fwd": [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P],
    "b200slu_glu_dwconv_bwd": [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P, _P],
    "b200slu_glu_dwconv_bwd_b": [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P, _P, _P],
    "b200slu_bn_finalize": [_P, _I, _L, _F, _F, _P, _P, _P, _P, _P],
    "b200slu_bn_swish_fwd": [_P, _P, _P, _P, _L, _I, _P, _P],
    "b200slu_bn_swish_bwd": [_P, _P, _P, _P, _P, _L, _I, _P, _P, _P, _P, _P],
    "b200slu_conv1_fwd": [_P, _P, _P, _I, _I, _I, _I, _P, _P],
    "b200slu_conv1_bwd_workspace_size": [_I, _I, _I, _I],
    "b200slu_conv1_bwd": [_P, _P, _I, _I, _I, _I, _P, _P, _P, _L, _P],
    "b200slu_im2col3x3s2": [_P, _I, _I, _I, _I, _P, _P],
    "b200slu_col2im3x3s2_relu": [_P, _P, _I, _I, _I, _I, _P, _P],
    "b200slu_conv2_implicit_supported": [_I, _I, _I, _I],
    "b200slu_conv2_fwd": [_P, _I, _I, _I, _I, _P, _P, _P, _P],
    "b200slu_conv2_wgrad": [_P, _P, _I, _I, _I, _I, _P, _P],
    "b200slu_conv2_dgrad_relu": [_P, _P, _P, _I, _I, _I, _I, _P, _P],
    "b200slu_conv1_fwd_bits": [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P],
    "b200slu_relu_bits": [_P, _L, _I, _P, _P],
    "b200slu_ctc_workspace_size": [_I, _I, _I],
    "b200slu_ctc_lattice_states": [_P, _I, _P, _I, _I, _P, _P, _P],
    "b200slu_ctc_fwd": [_P, _L, _I, _I, _I, _P, _P, _I, _P, _I, _P, _P, _P, _P],
    "b200slu_ctc_bwd": [_P, _L, _I, _I, _I, _P, _P, _I, _P, _I, _P, _F, _P, _P, _L, _P],
    "b200slu_lsm_workspace_size": [_L],
    "b200slu_lsm_loss_fwd_bwd": [_P, _L, _L, _I, _P, _I, _F, _F, _P, _P, _P, _P, _L, _P],
    "b200slu_logmel_fwd": [_P, _L, _I, _I, _P, _P, _I, _I, _I, _P, _P, _P],
    "b200slu_set_step_params": [_P],
    "b200slu_adam_step": [_P, _P, _P, _P, _L, _F, _F, _F, _F, _F, _I, _F, _F, _P, _P],
    "b200slu_set_sm_budget": [_I],
    "b200slu_adam_step_sched": [_P, _P, _P, _P, _L, _F, _F, _F, _F, _F, _F, _F, _F, _P, _P],
    "b200slu_adam_step_sched_lp": [_P, _P, _P, _P, _L, _F, _F, _F, _F, _F, _F, _F, _F, _P, _P, _P],
    "b200slu_add_shards_f32": [_P, _P, _L, _I, _L, _P],
    "b200slu_lsm_static_max_rows": [],
    "b200slu_lsm_loss_fwd_bwd_static": [_P, _L, _L, _I, _P, _I, _F, _F, _P, _P, _P, _L, _P],
    "b200slu_vocab_softmax_fwd": [_P, _L, _L, _I, _P, _L, _P],
    "b200slu_vocab_log_softmax_fwd": [_P, _L, _L, _I, _P, _L, _P],
    "b200slu_vocab_softmax_bwd": [_P, _L, _P, _L, _L, _I, _P, _L, _P, _L, _P],
    "b200slu_global_mvn_fwd": [_P, _I, _I, _I, _P, _P, _P, _P, _P],
    "b200slu_specaug_mvn_fwd": [_P, _I, _I, _I, _P, _P, _P, _P, _I, _P, _I, _I, _P, _P],
}
_RESTYPES = {
    "b200slu_last_error": C.c_char_p,
    "b200slu_prof_family_name": C.c_char_p,
    "b200slu_prof_total_launches": C.c_longlong,
    "b200slu_prof_reset": None,
    "b200slu_prof_enable_timing": None,
    "b200slu_ctc_workspace_size": C.c_longlong,
    "b200slu_conv1_bwd_workspace_size": C.c_longlong,
    "b200slu_lsm_workspace_size": C.c_longlong,
    "b200slu_lsm_static_max_rows": C.c_longlong,
    "b200slu_rel_shift_index": C.c_longlong,
}
EXPORTS = tuple(_SIGNATURES)

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -m espnet_slurp_b200.build` "
                "(there is no CPU or PyTorch fallback for the B200 kernels)")
        _lib = C.CDLL(LIB_PATH)
        for name, args in _SIGNATURES.items():
            fn = getattr(_lib, name)  # AttributeError here = header / library mismatch
            fn.argtypes = args
            fn.restype = _RESTYPES.get(name, C.c_int)
    return _lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib().b200slu_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"libb200slu {what} failed ({rc}): {msg}")


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    return None if t is None else t.data_ptr()


def stream_ptr():
    import torch

    return torch.cuda.current_stream().cuda_stream
