"""Per-kernel Python wrappers over the C ABI of libb200slu (include/b200slu.h).

Every function takes / returns CUDA torch tensors; torch only owns memory and streams.  There is
no CPU or eager-PyTorch fallback: a missing library or a non-zero status raises.

Conventions shared with the step orchestration (``espnet_slurp_b200.step``):
  * activations fed to GEMMs are bf16, the residual stream / statistics / losses are fp32;
  * ``drop`` is ``None`` or ``(p, seed, site)``: counter-based Philox dropout, regenerated (never
    stored) in the backward from the same triple;
  * attention score tensors are head-major ``[H, B, Tq, ld]`` with ``ld`` = keys rounded up to 8.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import Epilogue, GemmDesc, check, ptr, stream_ptr

ACT = {None: 0, "none": 0, "swish": 1, "relu": 2, "mul": 3}
RELPOS = {None: 0, "none": 0, "latest": 1, "legacy": 2}
Drop = Optional[Tuple[float, int, int]]


def _drop(drop: Drop):
    if drop is None or drop[0] <= 0.0:
        return 0.0, 0, 0
    return float(drop[0]), int(drop[1]), int(drop[2])


def round8(n: int) -> int:
    return (n + 7) // 8 * 8


def _epilogue(bias, act, alpha, residual, act_grad_input, act_grad, pre_act_out, out, accumulate, drop, colsum_out=None):
    ep = Epilogue()
    ep.bias = ptr(bias)
    ep.act = ACT[act]
    ep.act_grad_out = ptr(pre_act_out)
    ep.alpha = alpha
    ep.dropout_p, ep.dropout_seed, ep.dropout_site = _drop(drop)
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
    if colsum_out is not None:
        assert colsum_out.dtype == torch.float32 and out.dtype == torch.bfloat16
        ep.out_colsum_accum = colsum_out.data_ptr()
    return ep


def gemm(a: torch.Tensor, b: torch.Tensor, *, trans: bool = False, trans_a: Optional[bool] = None,
         trans_b: Optional[bool] = None, bias: Optional[torch.Tensor] = None, act: Optional[str] = None,
         alpha: float = 1.0, residual: Optional[torch.Tensor] = None, act_grad_input: Optional[torch.Tensor] = None,
         act_grad: Optional[str] = None, act_grad_out=False, out_dtype=torch.bfloat16,
         out: Optional[torch.Tensor] = None, accumulate: bool = False, drop: Drop = None,
         dropout_p: float = 0.0, seed: int = 0, site: int = 0, colsum_out: Optional[torch.Tensor] = None):
    """C[M,N] = op(A) op(B)^T with the fused epilogue of b200slu_epilogue_t.

    trans_a=False: a is [M,K]; True: a is [K,M].  trans_b=False: b is [N,K]; True: b is [K,N].
    (``trans`` sets both.)  ``act_grad_out`` (True = allocate, or a bf16 [M,N] tensor) receives act'(y) times the
    element's dropout multiplier; feed it back as ``act_grad_input`` with ``act_grad="mul"`` in the dgrad GEMM.
    ``colsum_out`` (f32 [N], accumulated): column sums of the bf16 output, from the epilogue's staged tiles.
    """
    pre_act_out = act_grad_out
    ta = trans if trans_a is None else trans_a
    tb = trans if trans_b is None else trans_b
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_cuda
    assert a.stride(-1) == 1 and b.stride(-1) == 1
    (M, K) = (a.shape[1], a.shape[0]) if ta else (a.shape[0], a.shape[1])
    (N, Kb) = (b.shape[1], b.shape[0]) if tb else (b.shape[0], b.shape[1])
    assert K == Kb, (a.shape, b.shape, ta, tb)
    if out is None:
        out = torch.empty(M, N, device=a.device, dtype=out_dtype)
    pre = None
    if pre_act_out is True:
        pre = torch.empty(M, N, device=a.device, dtype=torch.bfloat16)
    elif pre_act_out is not False and pre_act_out is not None:
        pre = pre_act_out
    if pre is not None:
        assert pre.stride(0) == out.stride(0)
    if drop is None and dropout_p > 0.0:
        drop = (dropout_p, seed, site)
    ep = _epilogue(bias, act, alpha, residual, act_grad_input, act_grad, pre, out, accumulate, drop, colsum_out)
    g = GemmDesc()
    g.trans_a, g.trans_b = int(ta), int(tb)
    g.M, g.N, g.K = M, N, K
    g.A, g.lda = a.data_ptr(), a.stride(0)
    g.B, g.ldb = b.data_ptr(), b.stride(0)
    g.C, g.ldc = out.data_ptr(), out.stride(0)
    g.batch_inner = g.batch_outer = 1
    check(_lib.lib().b200slu_gemm_bf16_ex(C.addressof(g), C.addressof(ep), stream_ptr()), "gemm_bf16")
    return (out, pre) if pre_act_out is True else out


def _bgemm(ta, tb, M, N, K, A, lda, a_si, a_so, Bm, ldb, b_si, b_so, Cm, ldc, c_si, c_so, nbi, nbo, accumulate=False):
    g = GemmDesc()
    g.trans_a, g.trans_b = int(ta), int(tb)
    g.M, g.N, g.K = M, N, K
    g.A, g.lda, g.a_stride_inner, g.a_stride_outer = A.data_ptr(), lda, a_si, a_so
    g.B, g.ldb, g.b_stride_inner, g.b_stride_outer = Bm.data_ptr(), ldb, b_si, b_so
    g.C, g.ldc, g.c_stride_inner, g.c_stride_outer = Cm.data_ptr(), ldc, c_si, c_so
    g.batch_inner, g.batch_outer = nbi, nbo
    ep = Epilogue()
    ep.alpha = 1.0
    ep.out_f32 = 1 if Cm.dtype == torch.float32 else 0
    ep.accumulate = 1 if accumulate else 0
    check(_lib.lib().b200slu_gemm_bf16_ex(C.addressof(g), C.addressof(ep), stream_ptr()), "gemm_bf16_ex")


# ---- the four batched attention contractions (x, y are [rows, H*dk] bf16 views with row stride ld) ------------
def attn_f1(x, y, B: int, Tx: int, Ty: int, H: int, y_bcast: bool = False):
    """out[h,b,i,j] = sum_c x[b*Tx+i, h*dk+c] * y[b*Ty+j, h*dk+c]  -> f32 [H,B,Tx,round8(Ty)]
    (y_bcast: y is [Ty, H*dk], shared by all utterances)."""
    dk = x.shape[1] // H
    ld = round8(Ty)
    out = torch.empty(H, B, Tx, ld, device=x.device, dtype=torch.float32)
    _bgemm(0, 0, Tx, Ty, dk, x, x.stride(0), dk, Tx * x.stride(0), y, y.stride(0), dk, 0 if y_bcast else Ty * y.stride(0),
           out, ld, B * Tx * ld, Tx * ld, H, B)
    return out


def attn_f2(s, y, B: int, Tq: int, Tk: int, H: int, out=None, y_bcast: bool = False):
    """out[b*Tq+i, h*dk+c] = sum_j s[h,b,i,j] * y[b*Tk+j, h*dk+c]  -> bf16 [B*Tq, H*dk] (or into ``out`` view)."""
    dk = y.shape[1] // H
    ld = s.shape[-1]
    if out is None:
        out = torch.empty(B * Tq, H * dk, device=s.device, dtype=torch.bfloat16)
    _bgemm(0, 1, Tq, dk, Tk, s, ld, B * Tq * ld, Tq * ld, y, y.stride(0), dk, 0 if y_bcast else Tk * y.stride(0),
           out, out.stride(0), dk, Tq * out.stride(0), H, B)
    return out


def attn_f3(s, x, B: int, Tq: int, Tk: int, H: int, out=None):
    """out[b*Tk+j, h*dk+c] = sum_i s[h,b,i,j] * x[b*Tq+i, h*dk+c]  -> bf16 [B*Tk, H*dk] (or into ``out`` view)."""
    dk = x.shape[1] // H
    ld = s.shape[-1]
    if out is None:
        out = torch.empty(B * Tk, H * dk, device=s.device, dtype=torch.bfloat16)
    _bgemm(1, 1, Tk, dk, Tq, s, ld, B * Tq * ld, Tq * ld, x, x.stride(0), dk, Tq * x.stride(0),
           out, out.stride(0), dk, Tk * out.stride(0), H, B)
    return out


def attn_f4(s, x, B: int, T: int, P: int, H: int, out):
    """out[n, h*dk+c] += sum_{b,i} s[h,b,i,n] * x[b*T+i, h*dk+c]  (f32 [P, H*dk], accumulated; split-K atomics)."""
    dk = x.shape[1] // H
    ld = s.shape[-1]
    _bgemm(1, 1, P, dk, B * T, s, ld, B * T * ld, 0, x, x.stride(0), dk, 0, out, out.stride(0), dk, 0, H, 1,
           accumulate=True)
    return out


# ---- LayerNorm ------------------------------------------------------------------------------------------------
def layernorm_fwd(x, gamma, beta, eps: float = 1e-12, want_bf16: bool = True, want_f32: bool = False):
    M, D = x.shape
    y16 = torch.empty(M, D, device=x.device, dtype=torch.bfloat16) if want_bf16 else None
    y32 = torch.empty(M, D, device=x.device, dtype=torch.float32) if want_f32 else None
    stats = torch.empty(M, 2, device=x.device, dtype=torch.float32)
    check(_lib.lib().b200slu_layernorm_fwd(ptr(x), ptr(gamma), ptr(beta), M, D, eps, ptr(y16), ptr(y32), ptr(stats),
                                           stream_ptr()), "layernorm_fwd")
    return y16, y32, stats


def layernorm_bwd(dy, x, gamma, stats, dgamma, dbeta, dx_add=None, want_g: bool = False, g_alpha: float = 1.0,
                  g_drop: Drop = None, g_colsum_out=None):
    """dx (+ dx_add) and, with want_g, g16 = bf16(g_alpha * dropout(dx)) whose column sums accumulate into g_colsum_out."""
    M, D = x.shape
    dx = torch.empty(M, D, device=x.device, dtype=torch.float32)
    g16 = torch.empty(M, D, device=x.device, dtype=torch.bfloat16) if want_g else None
    assert g_colsum_out is None or want_g
    p, seed, site = _drop(g_drop)
    check(_lib.lib().b200slu_layernorm_bwd(ptr(dy), int(dy.dtype == torch.float32), ptr(x), ptr(gamma), ptr(stats),
                                           ptr(dx_add), M, D, ptr(dx), ptr(dgamma), ptr(dbeta), ptr(g16), g_alpha,
                                           ptr(g_colsum_out), p, seed, site, stream_ptr()), "layernorm_bwd")
    return dx, g16


# ---- small helpers ----------------------------------------------------------------------------------------------
def cast_bf16(x, alpha: float = 1.0, drop: Drop = None, out=None, colsum_out=None):
    """bf16(alpha * dropout(x)); with ``colsum_out`` (f32 [N], accumulated) also the column sums of a 2-D x."""
    assert x.dtype == torch.float32 and x.is_contiguous()
    if out is None:
        out = torch.empty(x.shape, device=x.device, dtype=torch.bfloat16)
    p, seed, site = _drop(drop)
    if colsum_out is not None:
        M, N = x.shape
        check(_lib.lib().b200slu_cast_colsum_bf16(ptr(x), ptr(out), M, N, alpha, p, seed, site, ptr(colsum_out), stream_ptr()),
              "cast_colsum")
        return out
    check(_lib.lib().b200slu_cast_bf16(ptr(x), ptr(out), x.numel(), alpha, p, seed, site, stream_ptr()), "cast_bf16")
    return out


def colsum(dy, out_accum, n: Optional[int] = None):
    M = dy.shape[0]
    N = dy.shape[1] if n is None else n
    check(_lib.lib().b200slu_colsum_bf16(ptr(dy), M, N, dy.stride(0), ptr(out_accum), stream_ptr()), "colsum")
    return out_accum


def pos_bias_fwd(q, u, v):
    M, d = q.shape
    qu = torch.empty(M, d, device=q.device, dtype=torch.bfloat16)
    qv = torch.empty(M, d, device=q.device, dtype=torch.bfloat16)
    check(_lib.lib().b200slu_pos_bias_fwd(ptr(q), q.stride(0), ptr(u), ptr(v), M, d, ptr(qu), ptr(qv), stream_ptr()),
          "pos_bias_fwd")
    return qu, qv


def pos_bias_bwd(dqu, dqv, dq_out, du, dv, dq_colsum=None):
    """dq = dqu + dqv; du += colsum(dqu); dv += colsum(dqv); dq_colsum (optional, f32 [d]) += colsum(dq)."""
    M, d = dqu.shape
    check(_lib.lib().b200slu_pos_bias_bwd_q(ptr(dqu), ptr(dqv), M, d, ptr(dq_out), dq_out.stride(0), ptr(du), ptr(dv),
                                            ptr(dq_colsum), stream_ptr()), "pos_bias_bwd")


def embed_fwd(emb, tokens, pe, L: int, scale: float, drop: Drop = None):
    R, d = tokens.numel(), emb.shape[1]
    x = torch.empty(R, d, device=emb.device, dtype=torch.float32)
    p, seed, site = _drop(drop)
    check(_lib.lib().b200slu_embed_fwd(ptr(emb), ptr(tokens), ptr(pe), R, L, d, scale, ptr(x), p, seed, site, stream_ptr()),
          "embed_fwd")
    return x


def embed_bwd(dx, tokens, scale: float, demb, drop: Drop = None):
    R, d = dx.shape
    p, seed, site = _drop(drop)
    check(_lib.lib().b200slu_embed_bwd(ptr(dx), ptr(tokens), R, d, scale, ptr(demb), p, seed, site, stream_ptr()),
          "embed_bwd")


# ---- softmax ------------------------------------------------------------------------------------------------------
def softmax_fwd(ac, bd, key_lens, Tk: int, mode: str, causal: bool, scale: float, drop: Drop = None):
    H, B, Tq, ld = ac.shape
    p = torch.empty(H, B, Tq, ld, device=ac.device, dtype=torch.bfloat16)
    dp, seed, site = _drop(drop)
    pd = torch.empty_like(p) if dp > 0.0 else p
    check(_lib.lib().b200slu_softmax_fwd(ptr(ac), ld, ptr(bd), bd.shape[-1] if bd is not None else 0, ptr(key_lens), H, B, Tq,
                                         Tk, RELPOS[mode], int(causal), scale, ptr(p), ptr(pd), ld, dp, seed, site,
                                         stream_ptr()), "softmax_fwd")
    return p, pd


def softmax_bwd(dpd, p, Tk: int, mode: str, scale: float, ld_bd: int = 0, drop: Drop = None):
    H, B, Tq, ld = p.shape
    dac = torch.empty(H, B, Tq, ld, device=p.device, dtype=torch.bfloat16)
    dbd = torch.empty(H, B, Tq, ld_bd, device=p.device, dtype=torch.bfloat16) if RELPOS[mode] else None
    dp, seed, site = _drop(drop)
    check(_lib.lib().b200slu_softmax_bwd(ptr(dpd), ptr(p), ld, H, B, Tq, Tk, RELPOS[mode], scale, ptr(dac), ptr(dbd), ld_bd,
                                         dp, seed, site, stream_ptr()), "softmax_bwd")
    return dac, dbd


# ---- fused attention (scores -> rel_shift -> softmax -> dropout -> context) --------------------------------------
CHUNK_KEYS = 192  # keys per launch of the fused attention kernels (multiple of 64: a chunk is whole 64-column p blocks)
TAIL_MIN_UNITS = 128  # (utterance, head) units from which a short last key chunk gets a backward launch of its own


def relattn_supported(Tq: int, Tk: int, dk: int, mode: str, causal: bool = False) -> bool:
    """The fused attention kernels take any number of keys: up to 192 in one launch, more as key chunks + a combine pass
    (causal masks only within one launch -- the decoder's token axis is short)."""
    L = _lib.lib()
    if Tk <= CHUNK_KEYS:
        return bool(L.b200slu_relattn_supported(Tq, Tk, dk, RELPOS[mode]))
    if causal or (RELPOS[mode] and Tq != Tk) or -(-Tk // CHUNK_KEYS) > 16:
        return False
    if os.environ.get("B200SLU_ATTN_CHUNKS", "1") == "0":  # A/B measurements against the unfused chain
        return False
    return bool(L.b200slu_relattn_chunk_supported(Tq, CHUNK_KEYS, dk, RELPOS[mode]))


def _key_chunks(Tk: int):
    return [(o, min(CHUNK_KEYS, Tk - o)) for o in range(0, Tk, CHUNK_KEYS)]


def relattn_fwd(qu, qv, k, v, pos, key_lens, B: int, Tq: int, Tk: int, H: int, mode: str, causal: bool, scale: float,
                drop: Drop = None, want_pd: bool = True):
    """One kernel for attention.py:281-308 / :98-114.  qu, qv [B*Tq, H*64] (qv, pos only for mode "latest"), k, v
    [B*Tk, H*64] (column slices of wider buffers are fine) -> (p, p_drop) bf16 [H,B,Tq,round8(Tk)], ctx bf16 [B*Tq, H*64].
    More than 192 keys: one launch per key chunk + the combine pass (p_drop is not produced then)."""
    d = k.shape[1]
    ld = round8(Tk)
    L = _lib.lib()
    p = torch.empty(H, B, Tq, ld, device=k.device, dtype=torch.bfloat16)
    dp, seed, site = _drop(drop)
    ctx = torch.empty(B * Tq, d, device=k.device, dtype=torch.bfloat16)
    assert qu.stride(0) == (qv.stride(0) if qv is not None else qu.stride(0)) and k.stride(0) == v.stride(0)
    ldpos = pos.stride(0) if pos is not None else 0
    if Tk <= CHUNK_KEYS:
        pd = (torch.empty_like(p) if want_pd else None) if dp > 0.0 else p  # the fused backward regenerates dropout(p) from p
        check(L.b200slu_relattn_fwd(ptr(qu), ptr(qv), qu.stride(0), ptr(k), ptr(v), k.stride(0), ptr(pos), ldpos, ptr(key_lens), B,
                                    H, Tq, Tk, RELPOS[mode], int(causal), scale, ptr(p), ptr(pd), ld, ptr(ctx), ctx.stride(0), dp,
                                    seed, site, stream_ptr()), "relattn_fwd")
        return p, pd, ctx
    assert not causal and not (want_pd and dp > 0.0), "key chunks: no causal mask, no stored dropout(p)"
    chunks = _key_chunks(Tk)
    ctxc = torch.empty(len(chunks), B * Tq, d, device=k.device, dtype=torch.bfloat16)
    lse = torch.empty(len(chunks), H * B * Tq, device=k.device, dtype=torch.float32)
    if ld > Tk:
        p[..., Tk:].zero_()
    for c, (off, n) in enumerate(chunks):
        check(L.b200slu_relattn_fwd_chunk(ptr(qu), ptr(qv), qu.stride(0), ptr(k[off:]), ptr(v[off:]), k.stride(0), ptr(pos), ldpos,
                                          ptr(key_lens), B, H, Tq, n, RELPOS[mode], 0, scale, ptr(p[..., off:]), ld, ptr(ctxc[c]),
                                          d, dp, seed, site, off, Tk, ptr(lse[c]), stream_ptr()), "relattn_fwd_chunk")
    check(L.b200slu_relattn_combine(ptr(p), ld, Tk, CHUNK_KEYS, len(chunks), ptr(lse), ptr(ctxc), B * Tq * d, d, ptr(ctx),
                                    ctx.stride(0), B, H, Tq, stream_ptr()), "relattn_combine")
    return p, (None if dp > 0.0 else p), ctx


def relattn_bwd_supported(Tq: int, Tk: int, dk: int, mode: str) -> bool:
    if Tk <= CHUNK_KEYS:
        return bool(_lib.lib().b200slu_relattn_bwd_supported(Tq, Tk, dk, RELPOS[mode]))
    return relattn_supported(Tq, Tk, dk, mode)


def relattn_bwd(dctx, qu, k, v, p, ctx, B: int, Tq: int, Tk: int, H: int, mode: str, scale: float, dq, dk, dv,
                drop: Drop = None, want_dac: bool = False, qv=None, pos=None, dqv=None, dpos=None, dk_colsum=None,
                dv_colsum=None):
    """One kernel for the backward of the attention core (autograd through attention.py:78-95, :297-302 / :111-113, rel_shift
    :240-263).  dctx, qu, ctx [B*Tq, H*64], k, v [B*Tk, H*64], p bf16 [H,B,Tq,round8(Tk)] (as saved by the forward); writes
    dq [B*Tq, H*64], dk, dv [B*Tk, H*64] (views with their own row strides are fine).  Mode "latest": qv, pos (projected
    positions [2T-1, H*64]) in, dqv (row stride of dq) out, dpos (f32 [2T-1, H*64]) ACCUMULATED.  want_dac: also return dS
    (bf16, shifted layout [H,B,Tq,round8(Tk)]).  More than 192 keys: one launch per key chunk, dq / dqv summed over them.
    dk_colsum / dv_colsum (f32 [H*64], accumulated): column sums of dk / dv = the K / V projections' bias gradients."""
    ld = p.shape[-1]
    L = _lib.lib()
    dp, seed, site = _drop(drop)
    assert k.stride(0) == v.stride(0) and dk.stride(0) == dv.stride(0)
    rel = RELPOS[mode]
    if rel:
        assert qv.stride(0) == qu.stride(0) and dqv.stride(0) == dq.stride(0) and dpos.dtype == torch.float32
    ldpos = pos.stride(0) if pos is not None else 0
    lddpos = dpos.stride(0) if dpos is not None else 0
    if Tk <= CHUNK_KEYS and dk_colsum is not None:
        assert not want_dac and dv_colsum is not None
        check(L.b200slu_relattn_bwd_chunks(ptr(dctx), dctx.stride(0), ptr(qu), qu.stride(0), ptr(k), ptr(v), k.stride(0), ptr(p), ld,
                                           ptr(ctx), ctx.stride(0), B, H, Tq, Tk, Tk, rel, scale, ptr(dq), dq.stride(0), 0,
                                           ptr(dk), ptr(dv), dk.stride(0), dp, seed, site, ptr(qv), ptr(pos), ldpos, ptr(dqv),
                                           ptr(dpos), lddpos, 0, Tk, ptr(dk_colsum), ptr(dv_colsum), stream_ptr()),
              "relattn_bwd_chunks")
        return None
    if Tk <= CHUNK_KEYS:
        dac = torch.empty_like(p) if want_dac else None
        check(L.b200slu_relattn_bwd(ptr(dctx), dctx.stride(0), ptr(qu), qu.stride(0), ptr(k), ptr(v), k.stride(0), ptr(p), ld,
                                    ptr(ctx), ctx.stride(0), B, H, Tq, Tk, rel, scale, ptr(dq), dq.stride(0), ptr(dk), ptr(dv),
                                    dk.stride(0), ptr(dac), dp, seed, site, ptr(qv), ptr(pos), ldpos, ptr(dqv), ptr(dpos), lddpos,
                                    stream_ptr()), "relattn_bwd")
        return dac
    assert not want_dac
    chunks = _key_chunks(Tk)
    d = H * 64
    dqc = torch.empty(2 if rel else 1, len(chunks), B * Tq, d, device=k.device, dtype=torch.bfloat16)
    # the full chunks in one launch; a short last chunk in its own, sized to its keys (whole 64-key blocks)
    nfull = Tk // CHUNK_KEYS
    tail = Tk - nfull * CHUNK_KEYS
    if -(-tail // 64) * 64 >= CHUNK_KEYS or B * H < TAIL_MIN_UNITS:
        tail = 0  # nothing to save, or too few (utterance, head) units to fill the machine with a launch of its own
    for begin, end, ck, c0 in ([(0, Tk, CHUNK_KEYS, 0)] if tail == 0 else
                               [(0, nfull * CHUNK_KEYS, CHUNK_KEYS, 0), (nfull * CHUNK_KEYS, Tk, -(-tail // 64) * 64, nfull)]):
        check(L.b200slu_relattn_bwd_chunks(ptr(dctx), dctx.stride(0), ptr(qu), qu.stride(0), ptr(k), ptr(v), k.stride(0), ptr(p), ld,
                                           ptr(ctx), ctx.stride(0), B, H, Tq, Tk, ck, rel, scale, ptr(dqc[0, c0]), d,
                                           B * Tq * d, ptr(dk), ptr(dv), dk.stride(0), dp, seed, site, ptr(qv), ptr(pos), ldpos,
                                           ptr(dqc[1, c0]) if rel else None, ptr(dpos), lddpos, begin, end, ptr(dk_colsum),
                                           ptr(dv_colsum), stream_ptr()),
              "relattn_bwd_chunks")
    check(L.b200slu_sum_chunks_bf16(ptr(dqc[0]), B * Tq * d, d, len(chunks), B * Tq, d, ptr(dq), dq.stride(0), stream_ptr()), "sum_chunks")
    if rel:
        check(L.b200slu_sum_chunks_bf16(ptr(dqc[1]), B * Tq * d, d, len(chunks), B * Tq, d, ptr(dqv), dqv.stride(0), stream_ptr()),
              "sum_chunks")
    return None


def rel_unshift(dac, T: int, mode: str, ld_bd: int):
    """d(matrix_bd) (un-shifted, bf16 [H,B,T,ld_bd]) from d(scores) (shifted layout)."""
    H, B, Tq, ld = dac.shape
    dbd = torch.empty(H, B, Tq, ld_bd, device=dac.device, dtype=torch.bfloat16)
    check(_lib.lib().b200slu_rel_unshift(ptr(dac), ld, H, B, T, RELPOS[mode], ptr(dbd), ld_bd, stream_ptr()), "rel_unshift")
    return dbd


# ---- conv module --------------------------------------------------------------------------------------------------
def glu_dwconv_fwd(pw1, w, bias, B: int, T: int, K: int, sums=None):
    """sums: zeroed f32 [2 C] the kernel adds the BatchNorm sums into (allocated here when None)."""
    Cc = pw1.shape[1] // 2
    out = torch.empty(B * T, Cc, device=pw1.device, dtype=torch.bfloat16)
    if sums is None:
        sums = torch.zeros(2 * Cc, device=pw1.device, dtype=torch.float32)
    check(_lib.lib().b200slu_glu_dwconv_fwd(ptr(pw1), ptr(w), ptr(bias), B, T, Cc, K, ptr(out), ptr(sums), stream_ptr()),
          "glu_dwconv_fwd")
    return out, sums


def glu_dwconv_bwd(d_out, pw1, w, B: int, T: int, K: int, dw, dbias, dpw1_colsum=None):
    """... dpw1_colsum (optional, f32 [2C], accumulated): column sums of d_pw1 = pointwise_conv1's bias gradient."""
    Cc = pw1.shape[1] // 2
    d_pw1 = torch.empty(B * T, 2 * Cc, device=pw1.device, dtype=torch.bfloat16)
    check(_lib.lib().b200slu_glu_dwconv_bwd_b(ptr(d_out), ptr(pw1), ptr(w), B, T, Cc, K, ptr(d_pw1), ptr(dw), ptr(dbias),
                                              ptr(dpw1_colsum), stream_ptr()), "glu_dwconv_bwd")
    return d_pw1


def bn_finalize(sums, n: int, eps: float, momentum: float, running_mean=None, running_var=None, num_batches=None):
    Cc = sums.numel() // 2
    mr = torch.empty(2 * Cc, device=sums.device, dtype=torch.float32)
    check(_lib.lib().b200slu_bn_finalize(ptr(sums), Cc, n, eps, momentum, ptr(mr), ptr(running_mean), ptr(running_var),
                                         ptr(num_batches), stream_ptr()), "bn_finalize")
    return mr


def bn_swish_fwd(x, mean_rstd, gamma, beta):
    M, Cc = x.shape
    y = torch.empty(M, Cc, device=x.device, dtype=torch.bfloat16)
    check(_lib.lib().b200slu_bn_swish_fwd(ptr(x), ptr(mean_rstd), ptr(gamma), ptr(beta), M, Cc, ptr(y), stream_ptr()),
          "bn_swish_fwd")
    return y


def bn_swish_bwd(dy, x, mean_rstd, gamma, beta, dgamma, dbeta, scratch=None):
    """scratch: zeroed f32 [2 C] for the two channel sums of the backward (allocated here when None)."""
    M, Cc = x.shape
    dx = torch.empty(M, Cc, device=x.device, dtype=torch.bfloat16)
    sums2 = scratch if scratch is not None else torch.zeros(2 * Cc, device=x.device, dtype=torch.float32)
    check(_lib.lib().b200slu_bn_swish_bwd(ptr(dy), ptr(x), ptr(mean_rstd), ptr(gamma), ptr(beta), M, Cc, ptr(sums2), ptr(dx),
                                          ptr(dgamma), ptr(dbeta), stream_ptr()), "bn_swish_bwd")
    return dx


# ---- subsampling --------------------------------------------------------------------------------------------------
def conv1_fwd(x, w, bias, want_bits: bool = False):
    """relu(conv.0(x) + bias) -> bf16 [B,T1,F1,C]; want_bits: also the (h1 > 0) bit rows int32 [B*T1*F1, C/32] that
    conv2_dgrad_relu reads instead of h1 (C % 256 == 0)."""
    B, Tin, F = x.shape
    Cc = w.shape[0]
    T1, F1 = (Tin - 3) // 2 + 1, (F - 3) // 2 + 1
    h1 = torch.empty(B, T1, F1, Cc, device=x.device, dtype=torch.bfloat16)
    if not want_bits:
        check(_lib.lib().b200slu_conv1_fwd(ptr(x), ptr(w), ptr(bias), B, Tin, F, Cc, ptr(h1), stream_ptr()), "conv1_fwd")
        return h1
    bits = torch.empty(B * T1 * F1, Cc // 32, device=x.device, dtype=torch.int32)
    check(_lib.lib().b200slu_conv1_fwd_bits(ptr(x), ptr(w), ptr(bias), B, Tin, F, Cc, ptr(h1), ptr(bits), stream_ptr()), "conv1_fwd")
    return h1, bits


def relu_bits(h1):
    Cc = h1.shape[-1]
    pixels = h1.numel() // Cc
    bits = torch.empty(pixels, Cc // 32, device=h1.device, dtype=torch.int32)
    check(_lib.lib().b200slu_relu_bits(ptr(h1), pixels, Cc, ptr(bits), stream_ptr()), "relu_bits")
    return bits


def conv1_bwd(dh1, x, dw, db):
    B, Tin, F = x.shape
    Cc = dh1.shape[-1]
    nbytes = int(_lib.lib().b200slu_conv1_bwd_workspace_size(B, Tin, F, Cc))
    ws = torch.empty(nbytes, device=x.device, dtype=torch.uint8)   # bf16 patches [pixels,16] + fp32 accumulator [C,16]
    check(_lib.lib().b200slu_conv1_bwd(ptr(dh1), ptr(x), B, Tin, F, Cc, ptr(dw), ptr(db), ptr(ws), nbytes, stream_ptr()),
          "conv1_bwd")


def im2col(h1):
    B, T1, F1, Cc = h1.shape
    T2, F2 = (T1 - 3) // 2 + 1, (F1 - 3) // 2 + 1
    col = torch.empty(B * T2 * F2, 9 * Cc, device=h1.device, dtype=torch.bfloat16)
    check(_lib.lib().b200slu_im2col3x3s2(ptr(h1), B, T1, F1, Cc, ptr(col), stream_ptr()), "im2col")
    return col


def col2im_relu(dcol, h1):
    B, T1, F1, Cc = h1.shape
    dh1 = torch.empty_like(h1)
    check(_lib.lib().b200slu_col2im3x3s2_relu(ptr(dcol), ptr(h1), B, T1, F1, Cc, ptr(dh1), stream_ptr()), "col2im")
    return dh1


def conv2_implicit_supported(B: int, T1: int, F1: int, C: int) -> bool:
    if os.environ.get("B200SLU_CONV2_IMPLICIT", "1") == "0":  # A/B measurements against the im2col + GEMM path
        return False
    return bool(_lib.lib().b200slu_conv2_implicit_supported(B, T1, F1, C))


def conv2_fwd(h1, w16, bias):
    """relu(conv.2(h1) + bias) as an implicit GEMM: h1 bf16 [B,T1,F1,C], w16 bf16 [C, 9C] -> bf16 [B*T2*F2, C]."""
    B, T1, F1, Cc = h1.shape
    T2, F2 = (T1 - 3) // 2 + 1, (F1 - 3) // 2 + 1
    h2 = torch.empty(B * T2 * F2, Cc, device=h1.device, dtype=torch.bfloat16)
    check(_lib.lib().b200slu_conv2_fwd(ptr(h1), B, T1, F1, Cc, ptr(w16), ptr(bias), ptr(h2), stream_ptr()), "conv2_fwd")
    return h2


def conv2_wgrad(dh2, h1, dw_accum):
    B, T1, F1, Cc = h1.shape
    check(_lib.lib().b200slu_conv2_wgrad(ptr(dh2), ptr(h1), B, T1, F1, Cc, ptr(dw_accum), stream_ptr()), "conv2_wgrad")


def conv2_dgrad_relu(dh2, w16, h1, bits=None):
    """dh1 = (h1 > 0) * conv_transpose(dh2, w): `bits` = the (h1 > 0) rows from conv1_fwd(want_bits=True), derived from h1 here
    when the caller has none."""
    B, T1, F1, Cc = h1.shape
    if bits is None:
        bits = relu_bits(h1)
    dh1 = torch.empty_like(h1)
    check(_lib.lib().b200slu_conv2_dgrad_relu(ptr(dh2), ptr(w16), ptr(bits), B, T1, F1, Cc, ptr(dh1), stream_ptr()), "conv2_dgrad")
    return dh1


# ---- losses -------------------------------------------------------------------------------------------------------
def ctc_fwd(logits, V: int, hlens, labels, label_lens, blank: int = 0):
    """logits f32 [B,T,ldv]; returns (loss_per_utt [B], loss_sum [1], workspace)."""
    B, T, ldv = logits.shape
    Lmax = labels.shape[1]
    ws = torch.empty(_lib.lib().b200slu_ctc_workspace_size(B, T, Lmax), device=logits.device, dtype=torch.uint8)
    loss = torch.empty(B, device=logits.device, dtype=torch.float32)
    loss_sum = torch.empty(1, device=logits.device, dtype=torch.float32)
    check(_lib.lib().b200slu_ctc_fwd(ptr(logits), ldv, B, T, V, ptr(hlens), ptr(labels), Lmax, ptr(label_lens), blank, ptr(ws),
                                     ptr(loss), ptr(loss_sum), stream_ptr()), "ctc_fwd")
    return loss, loss_sum, ws


def ctc_lattice_states(labels, label_lens, blank: int = 0):
    """(ext int32 [B, 2*Lmax+1], skip uint8): the lattice as the device code builds it (bit-exact test)."""
    B, Lmax = labels.shape
    ext = torch.empty(B, 2 * Lmax + 1, device=labels.device, dtype=torch.int32)
    skip = torch.empty(B, 2 * Lmax + 1, device=labels.device, dtype=torch.uint8)
    check(_lib.lib().b200slu_ctc_lattice_states(ptr(labels), Lmax, ptr(label_lens), B, blank, ptr(ext), ptr(skip), stream_ptr()),
          "ctc_lattice_states")
    return ext, skip


def ctc_bwd(logits, V: int, hlens, labels, label_lens, ws, scale: float, blank: int = 0, want_f32: bool = False,
            want_bf16: bool = True):
    B, T, ldv = logits.shape
    Lmax = labels.shape[1]
    ldg = round8(V)
    g32 = torch.empty(B, T, ldg, device=logits.device, dtype=torch.float32) if want_f32 else None
    g16 = torch.empty(B, T, ldg, device=logits.device, dtype=torch.bfloat16) if want_bf16 else None
    check(_lib.lib().b200slu_ctc_bwd(ptr(logits), ldv, B, T, V, ptr(hlens), ptr(labels), Lmax, ptr(label_lens), blank, ptr(ws),
                                     scale, ptr(g32), ptr(g16), ldg, stream_ptr()), "ctc_bwd")
    return g32, g16


def lsm_loss(logits, V: int, targets, smoothing: float, grad_scale: float, ignore: int = -1, want_f32: bool = False,
             want_bf16: bool = True):
    """logits f32 [R,ldv]; returns (stats3 = [sum loss, #correct, #rows], grad_f32, grad_bf16)."""
    R, ldv = logits.shape
    ldg = round8(V)
    ws = torch.empty(_lib.lib().b200slu_lsm_workspace_size(R), device=logits.device, dtype=torch.uint8)
    stats = torch.empty(3, device=logits.device, dtype=torch.float32)
    g32 = torch.empty(R, ldg, device=logits.device, dtype=torch.float32) if want_f32 else None
    g16 = torch.empty(R, ldg, device=logits.device, dtype=torch.bfloat16) if want_bf16 else None
    check(_lib.lib().b200slu_lsm_loss_fwd_bwd(ptr(logits), ldv, R, V, ptr(targets), ignore, smoothing, grad_scale, ptr(ws),
                                              ptr(stats), ptr(g32), ptr(g16), ldg, stream_ptr()), "lsm_loss")
    return stats, g32, g16


def vocab_softmax_fwd(logits, V: int):
    """softmax over the first V columns of f32 [R, ld]: bf16 [R, round8(V)], padding columns zero."""
    R, ldv = logits.shape
    out = torch.empty(R, round8(V), device=logits.device, dtype=torch.bfloat16)
    check(_lib.lib().b200slu_vocab_softmax_fwd(ptr(logits), ldv, R, V, ptr(out), out.shape[1], stream_ptr()), "vocab_softmax_fwd")
    return out


def vocab_log_softmax(logits, V: int):
    """log_softmax over the first V columns of f32 [R, ld] -> f32 [R, V]."""
    R, ldv = logits.shape
    out = torch.empty(R, V, device=logits.device, dtype=torch.float32)
    check(_lib.lib().b200slu_vocab_log_softmax_fwd(ptr(logits), ldv, R, V, ptr(out), V, stream_ptr()), "vocab_log_softmax")
    return out


def vocab_softmax_bwd(sm, dsm, V: int, add=None):
    """dlogits = sm * (dsm - <sm, dsm>) (+ add), bf16 [R, round8(V)]."""
    R = sm.shape[0]
    assert sm.dtype == dsm.dtype == torch.bfloat16 and sm.stride(1) == 1 and dsm.stride(1) == 1
    out = torch.empty(R, round8(V), device=sm.device, dtype=torch.bfloat16)
    check(_lib.lib().b200slu_vocab_softmax_bwd(ptr(sm), sm.stride(0), ptr(dsm), dsm.stride(0), R, V, ptr(add),
                                               add.stride(0) if add is not None else 0, ptr(out), out.shape[1], stream_ptr()),
          "vocab_softmax_bwd")
    return out


# ---- front-end ----------------------------------------------------------------------------------------------------
def logmel(wav, lens, melmat, n_fft: int = 512, hop: int = 128):
    B, N = wav.shape
    T = N // hop + 1
    n_mels = melmat.shape[1]
    melmat = melmat.contiguous()  # numpy hands over the (n_freq, n_mels) table as a transposed view
    assert wav.stride(1) == 1 and lens.dtype == torch.int32
    feats = torch.empty(B, T, n_mels, device=wav.device, dtype=torch.float32)
    flens = torch.empty(B, device=wav.device, dtype=torch.int32)
    check(_lib.lib().b200slu_logmel_fwd(ptr(wav), wav.stride(0), B, N, ptr(lens), ptr(melmat), n_fft, hop, n_mels, ptr(feats),
                                        ptr(flens), stream_ptr()), "logmel")
    return feats, flens


def specaug_mvn(x, lens, center=None, warped=None, fmask=None, tmask=None, do_mvn: bool = True):
    B, T, F = x.shape
    y = torch.empty_like(x)
    nf = fmask.shape[1] if fmask is not None else 0
    nt = tmask.shape[1] if tmask is not None else 0
    check(_lib.lib().b200slu_specaug_mvn_fwd(ptr(x), B, T, F, ptr(lens), ptr(center), ptr(warped), ptr(fmask), nf, ptr(tmask),
                                             nt, int(do_mvn), ptr(y), stream_ptr()), "specaug_mvn")
    return y


def global_mvn(x, lens, mean=None, inv_std=None):
    """GlobalMVN (global_mvn.py:66-94): (x - mean) * inv_std on the valid frames, 0 on the padding."""
    B, T, F = x.shape
    y = torch.empty_like(x)
    check(_lib.lib().b200slu_global_mvn_fwd(ptr(x), B, T, F, ptr(lens), ptr(mean), ptr(inv_std), ptr(y), stream_ptr()),
          "global_mvn")
    return y


def rel_shift_index(i: int, j: int, T: int, mode: str) -> int:
    return _lib.lib().b200slu_rel_shift_index(i, j, T, RELPOS[mode])


def dropout_mask(drop: Drop, n: int, device="cuda"):
    """uint8 [n]: 1 where flat element e of the dropout site ``drop`` = (p, seed, site) is kept (parity tests)."""
    p, seed, site = _drop(drop)
    keep = torch.empty(n, device=device, dtype=torch.uint8)
    check(_lib.lib().b200slu_dropout_mask(p, seed, site, n, ptr(keep), stream_ptr()), "dropout_mask")
    return keep


# ---- optimizer ----------------------------------------------------------------------------------------------------
def adam_step(params, grads, exp_avg, exp_avg_sq, lr: float, step: int, state3, betas=(0.9, 0.999), eps: float = 1e-8,
              weight_decay: float = 0.0, grad_inv_scale: float = 1.0, max_norm: float = 5.0):
    """In-place fused clip + Adam over flat fp32 buffers; state3 = float64[3] (sumsq, grad norm, applied flag)."""
    check(_lib.lib().b200slu_adam_step(ptr(params), ptr(grads), ptr(exp_avg), ptr(exp_avg_sq), params.numel(), lr, betas[0],
                                       betas[1], eps, weight_decay, step, grad_inv_scale, max_norm, ptr(state3), stream_ptr()),
          "adam_step")


def adam_step_sched(params, grads, exp_avg, exp_avg_sq, base_lr: float, warmup_steps: float, state4, betas=(0.9, 0.999),
                    eps: float = 1e-8, weight_decay: float = 0.0, grad_inv_scale: float = 1.0, max_norm: float = 5.0,
                    params_bf16=None):
    """clip + Adam + WarmupLR with the step count on the device: state4 = float64[4] (sumsq, grad norm, applied flag, number of
    applied updates so far -- persistent).  A non-finite gradient norm skips the update and does not advance the schedule.
    params_bf16: bf16 tensor of the same length that receives the updated parameters in the same pass."""
    if params_bf16 is not None:
        assert params_bf16.dtype == torch.bfloat16 and params_bf16.numel() == params.numel()
        check(_lib.lib().b200slu_adam_step_sched_lp(ptr(params), ptr(grads), ptr(exp_avg), ptr(exp_avg_sq), params.numel(), base_lr,
                                                    float(warmup_steps), betas[0], betas[1], eps, weight_decay, grad_inv_scale,
                                                    max_norm, ptr(state4), ptr(params_bf16), stream_ptr()), "adam_step_sched_lp")
        return
    check(_lib.lib().b200slu_adam_step_sched(ptr(params), ptr(grads), ptr(exp_avg), ptr(exp_avg_sq), params.numel(), base_lr,
                                             float(warmup_steps), betas[0], betas[1], eps, weight_decay, grad_inv_scale, max_norm,
                                             ptr(state4), stream_ptr()), "adam_step_sched")


def add_shards_f32(dst, src, shard_stride: int, nshards: int):
    """dst += sum_k src[k * shard_stride : k * shard_stride + dst.numel()] (fp32, in place, fixed order)."""
    check(_lib.lib().b200slu_add_shards_f32(ptr(dst), ptr(src), int(shard_stride), int(nshards), dst.numel(), stream_ptr()),
          "add_shards_f32")


def set_sm_budget(sms: int):
    """SMs the persistent kernels size their grids for on the current device (0 = all)."""
    check(_lib.lib().b200slu_set_sm_budget(int(sms)), "set_sm_budget")


def set_step_params(dev_tensor):
    """Register (or with None unregister) the 24-byte device buffer {u64 seed; f32 lr, bc1, bc2, pad}."""
    check(_lib.lib().b200slu_set_step_params(ptr(dev_tensor)), "set_step_params")
