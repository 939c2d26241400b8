"""Thin per-kernel Python wrappers over the C ABI (used by the drop-in modules and tests).

Every function takes/returns CUDA torch tensors; torch only owns memory and streams.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import Epilogue, check, ptr, stream_ptr

ACT = {None: 0, "none": 0, "swish": 1, "relu": 2}


def gemm(a: torch.Tensor, b: torch.Tensor, *, trans: bool = False, bias: Optional[torch.Tensor] = None,
         act: Optional[str] = None, alpha: float = 1.0, residual: Optional[torch.Tensor] = None,
         act_grad_input: Optional[torch.Tensor] = None, act_grad: Optional[str] = None,
         pre_act_out: bool = False, out_dtype=torch.bfloat16, out: Optional[torch.Tensor] = None,
         accumulate: bool = False, dropout_p: float = 0.0, seed: int = 0, site: int = 0):
    """trans=False: a[M,K] @ b[N,K]^T ; trans=True: a[K,M]^T @ b[K,N].  bf16 in, fused epilogue."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_cuda
    assert a.stride(-1) == 1 and b.stride(-1) == 1
    if not trans:
        M, K = a.shape
        N = b.shape[0]
        assert b.shape[1] == K
    else:
        K, M = a.shape
        N = b.shape[1]
        assert b.shape[0] == K
    if out is None:
        out = torch.empty(M, N, device=a.device, dtype=out_dtype)
    ep = Epilogue()
    ep.bias = bias.data_ptr() if bias is not None else None
    ep.act = ACT[act]
    pre = None
    if pre_act_out:
        pre = torch.empty(M, N, device=a.device, dtype=torch.bfloat16)
        ep.pre_act_out = pre.data_ptr()
    ep.alpha = alpha
    ep.dropout_p = dropout_p
    ep.dropout_seed = seed
    ep.dropout_site = site
    if act_grad_input is not None:
        ep.act_grad_input = act_grad_input.data_ptr()
        ep.ld_act_grad_input = act_grad_input.stride(0)
        ep.act_grad = ACT[act_grad]
    if residual is not None:
        assert residual.dtype == torch.float32
        ep.residual = residual.data_ptr()
        ep.ld_residual = residual.stride(0)
    ep.out_f32 = 1 if out.dtype == torch.float32 else 0
    ep.accumulate = 1 if accumulate else 0
    rc = _lib.lib().b200slu_gemm_bf16(C.c_int(1 if trans else 0), M, N, K, ptr(a), C.c_longlong(a.stride(0)), ptr(b),
                                      C.c_longlong(b.stride(0)), ptr(out), C.c_longlong(out.stride(0)), C.byref(ep),
                                      stream_ptr())
    check(rc, "gemm_bf16")
    return (out, pre) if pre_act_out else out
