"""Forward + hand-written backward of the Conformer CTC/attention train step over the libb200slu kernels.

This is the host-side orchestration only: every arithmetic step is one call into ``ops`` (= one or two
CUDA kernels behind the C ABI).  There is no autograd inside; the drop-in modules (``modules.py``) expose
each sub-network to PyTorch as ONE ``torch.autograd.Function`` whose backward calls the ``*_bwd`` functions
here.  No host<->device synchronisation happens anywhere in this file (lengths stay on the device), so
a whole step can be captured in a CUDA graph.

Reference being reproduced (paths relative to the reference root):
  EncoderLayer.forward            espnet/nets/pytorch_backend/conformer/encoder_layer.py:76-157
  RelPositionMultiHeadedAttention espnet/nets/pytorch_backend/transformer/attention.py:265-308 (legacy :167-209)
  ConvolutionModule.forward       espnet/nets/pytorch_backend/conformer/convolution.py:56-79
  PositionwiseFeedForward         espnet/nets/pytorch_backend/transformer/positionwise_feed_forward.py:30-32
  Conv2dSubsampling.forward       espnet/nets/pytorch_backend/transformer/subsampling.py:67-87
  ConformerEncoder.forward        espnet2/asr/encoder/conformer_encoder.py:292-368
  TransformerDecoder.forward      espnet2/asr/decoder/transformer_decoder.py:92-145, decoder_layer.py:63-134
  CTC.forward                     espnet2/asr/ctc.py:72-97
  LabelSmoothingLoss / accuracy   label_smoothing_loss.py:41-63, nets_utils.py:299-320
  ESPnetASRModel.forward          espnet2/asr/espnet_model.py:169-297

Tensors: ``P`` name -> fp32 parameter (reference state_dict names, plus fused views such as
``...self_attn.linear_qkv.weight``), ``W`` name -> bf16 copy in the layout the GEMM wants,
``G`` name -> fp32 gradient accumulator (same layout as ``W``).  See ``params.py``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

from . import ops as K  # tests substitute a torch emulation of the same functions for CPU checks of this file

F32 = torch.float32


@dataclass
class EncCfg:
    idim: int = 80
    d: int = 512
    heads: int = 8
    ffn: int = 2048
    blocks: int = 12
    kernel: int = 31
    rel_pos_type: str = "latest"
    dropout: float = 0.1
    pos_dropout: float = 0.1
    att_dropout: float = 0.1
    interctc_layer_idx: tuple = ()  # 1-based block indices whose (after_norm'ed) output also feeds the CTC (:336-347)
    interctc_use_conditioning: bool = False  # feed the CTC posterior of every tap back into the stream (:350-358)
    vocab: int = 0  # CTC output size (only the self-conditioning path needs it inside the encoder)


@dataclass
class DecCfg:
    vocab: int = 600
    d: int = 512
    heads: int = 8
    ffn: int = 2048
    blocks: int = 6
    dropout: float = 0.1
    pos_dropout: float = 0.1
    self_att_dropout: float = 0.0
    src_att_dropout: float = 0.0


class StepCtx:
    """Per-step dropout bookkeeping: one Philox seed per step, one site id per dropout call site."""

    def __init__(self, seed: int = 0, training: bool = True):
        self.seed = int(seed)
        self.training = training
        self._site = 0
        self.sites: Dict[str, tuple] = {}  # tag (the reference module that owns the Dropout) -> (p, site)

    def drop(self, p: float, tag: Optional[str] = None):
        """Allocate a site; returns the (p, seed, site) triple or None when dropout is inactive."""
        self._site += 1
        active = self.training and p > 0.0
        if tag is not None:
            self.sites[tag] = (float(p) if active else 0.0, self._site)
        if not active:
            return None
        return (float(p), self.seed, self._site)


# ------------------------------------------------------------------------------------------------------------------
# zeroed accumulators: the small f32 buffers kernels ADD into (BatchNorm sums, d(projected positions)) are cut out of one
# buffer that is cleared by one fill per step instead of three fills per encoder block
# ------------------------------------------------------------------------------------------------------------------
_ZEROS = [None, 0]  # (f32 buffer, floats handed out)


def _zeros_begin(device, floats: int):
    _ZEROS[0], _ZEROS[1] = torch.zeros(floats, device=device, dtype=F32), 0


def _zeros(n: int, device):
    """n zeroed floats (16-byte aligned piece of the step's buffer; a fresh tensor once that is used up)."""
    buf, used = _ZEROS
    if buf is None or buf.device != device or used + n > buf.numel():
        return torch.zeros(n, device=device, dtype=F32)
    _ZEROS[1] = used + (n + 3) // 4 * 4
    return buf[used:used + n]


# ------------------------------------------------------------------------------------------------------------------
# position tables (constants; built once per length with torch, like the reference's extend_pe)
# ------------------------------------------------------------------------------------------------------------------
_PE_CACHE: Dict[tuple, torch.Tensor] = {}
KEEP_IM2COL_BYTES = 8 << 30


def _sinusoid(pos: torch.Tensor, d: int) -> torch.Tensor:
    div = torch.exp(torch.arange(0, d, 2, dtype=F32) * -(math.log(10000.0) / d))
    pe = torch.zeros(pos.numel(), d, dtype=F32)
    pe[:, 0::2] = torch.sin(pos[:, None] * div)
    pe[:, 1::2] = torch.cos(pos[:, None] * div)
    return pe


def pos_table(kind: str, n: int, d: int, device) -> torch.Tensor:
    """kind = 'latest' (embedding.py:196-244: rows +(T-1)..-(T-1)), 'legacy' (embedding.py:137-170, reverse, maxlen 5000)
    or 'abs' (embedding.py:59-94)."""
    key = (kind, n, d, str(device))
    if key not in _PE_CACHE:
        if kind == "latest":
            pos = torch.arange(n - 1, -n, -1, dtype=F32)
        elif kind == "legacy":
            pos = torch.arange(4999, 4999 - n, -1, dtype=F32)
        else:
            pos = torch.arange(0, n, dtype=F32)
        _PE_CACHE[key] = _sinusoid(pos, d).to(device).contiguous()
    return _PE_CACHE[key]


def subsampled_len(t_in: int) -> int:
    """subsampling.py:57-59,87"""
    return ((t_in - 1) // 2 - 1) // 2


# ------------------------------------------------------------------------------------------------------------------
# position-wise feed-forward branch:  x + scale * dropout(w_2(dropout(act(w_1(LN(x))))))
# ------------------------------------------------------------------------------------------------------------------
def ffn_fwd(x, P, W, pre: str, norm: str, act: str, scale: float, p_drop: float, ctx: StepCtx):
    ln, _, st = K.layernorm_fwd(x, P[norm + ".weight"], P[norm + ".bias"])
    d1, d2 = ctx.drop(p_drop, pre + ".inner"), ctx.drop(p_drop, pre + ".outer")
    h, hpre = K.gemm(ln, W[pre + ".w_1.weight"], bias=P[pre + ".w_1.bias"], act=act, act_grad_out=True, drop=d1)
    y = K.gemm(h, W[pre + ".w_2.weight"], bias=P[pre + ".w_2.bias"], alpha=scale, residual=x, drop=d2, out_dtype=F32)
    return y, (x, ln, st, hpre, h, d1, d2, pre, norm, act, scale)


# Backward of "x + dropout(branch(LN(x)))": a branch starts from g = bf16(alpha * dropout(dy)) plus the bias gradient of its
# last Linear (column sums of g).  The LayerNorm backward that PRODUCES dy can emit both in the same pass (layernorm.cu),
# so each *_bwd takes an optional ready-made ``g`` and the ``nxt`` head (alpha, drop, bias-gradient tensor) of the branch
# that will consume its dx, and returns (dx, g for that branch or None).
def _head_g(dy, g, head):
    if g is not None:
        return g
    alpha, drop, bias_grad = head
    return K.cast_bf16(dy, alpha=alpha, drop=drop, colsum_out=bias_grad)


def _ln_bwd(dln, x, P, G, norm: str, st, dx_add, nxt):
    if nxt is None:
        return K.layernorm_bwd(dln, x, P[norm + ".weight"], st, G[norm + ".weight"], G[norm + ".bias"], dx_add=dx_add)
    alpha, drop, bias_grad = nxt
    return K.layernorm_bwd(dln, x, P[norm + ".weight"], st, G[norm + ".weight"], G[norm + ".bias"], dx_add=dx_add, want_g=True,
                           g_alpha=alpha, g_drop=drop, g_colsum_out=bias_grad)


def ffn_head(saved, G):
    pre, scale, d2 = saved[7], saved[10], saved[6]
    return scale, d2, G[pre + ".w_2.bias"]


def ffn_bwd(dy, saved, P, W, G, g=None, nxt=None):
    x, ln, st, hpre, h, d1, d2, pre, norm, act, scale = saved
    g = _head_g(dy, g, (scale, d2, G[pre + ".w_2.bias"]))
    K.gemm(g, h, trans=True, out=G[pre + ".w_2.weight"], accumulate=True)
    # hpre = act' * dropout mask; w_1's bias gradient = the column sums of dh, taken from the epilogue's staged tiles
    fuse = hpre.shape[1] % 64 == 0  # (whole 64-column boxes; the tiny fixtures take the separate pass)
    dh = K.gemm(g, W[pre + ".w_2.weight"], trans_b=True, act_grad_input=hpre, act_grad="mul",
                colsum_out=G[pre + ".w_1.bias"] if fuse else None)
    if not fuse:
        K.colsum(dh, G[pre + ".w_1.bias"])
    K.gemm(dh, ln, trans=True, out=G[pre + ".w_1.weight"], accumulate=True)
    dln = K.gemm(dh, W[pre + ".w_1.weight"], trans_b=True)
    return _ln_bwd(dln, x, P, G, norm, st, dy, nxt)


# ------------------------------------------------------------------------------------------------------------------
# multi-head attention core shared by the encoder (rel-pos) and the decoder (plain): scores -> softmax -> context
# ------------------------------------------------------------------------------------------------------------------
def relmha_fwd(x, pos16, klens, P, W, p: str, norm: str, cfg: EncCfg, B: int, T: int, ctx: StepCtx):
    d, H = cfg.d, cfg.heads
    a = p + ".self_attn"
    ln, _, st = K.layernorm_fwd(x, P[norm + ".weight"], P[norm + ".bias"])
    qkv = K.gemm(ln, W[a + ".linear_qkv.weight"], bias=P[a + ".linear_qkv.bias"])
    q, k, v = qkv[:, :d], qkv[:, d:2 * d], qkv[:, 2 * d:]
    qu, qv = K.pos_bias_fwd(q, P[a + ".pos_bias_u"].reshape(-1), P[a + ".pos_bias_v"].reshape(-1))
    pp = K.gemm(pos16, W[a + ".linear_pos.weight"])
    Pn = pos16.shape[0]
    da, do = ctx.drop(cfg.att_dropout, a + ".att"), ctx.drop(cfg.dropout, a + ".out")
    scale = 1.0 / math.sqrt(d // H)
    fb = False
    if K.relattn_supported(T, T, d // H, cfg.rel_pos_type):
        # short utterances: scores, rel_shift, softmax, dropout and context in one tcgen05 kernel (scores never leave TMEM);
        # the fused backward regenerates dropout(p) from p, so only p is saved
        fb = ctx.training and K.relattn_bwd_supported(T, T, d // H, cfg.rel_pos_type)
        pr, pd, c = K.relattn_fwd(qu, qv, k, v, pp, klens, B, T, T, H, cfg.rel_pos_type, False, scale, drop=da, want_pd=not fb)
    else:
        ac = K.attn_f1(qu, k, B, T, T, H)
        bd = K.attn_f1(qv, pp, B, T, Pn, H, y_bcast=True)
        pr, pd = K.softmax_fwd(ac, bd, klens, T, cfg.rel_pos_type, False, scale, drop=da)
        c = K.attn_f2(pd, v, B, T, T, H)
    y = K.gemm(c, W[a + ".linear_out.weight"], bias=P[a + ".linear_out.bias"], residual=x, drop=do, out_dtype=F32)
    return y, (x, ln, st, qkv, qu, qv, pp, pr, pd, c, da, do, scale, Pn, K.round8(Pn), fb)


def relmha_head(saved, G, p: str):
    return 1.0, saved[11], G[p + ".self_attn.linear_out.bias"]


def relmha_bwd(dy, saved, pos16, P, W, G, p: str, norm: str, cfg: EncCfg, B: int, T: int, g=None, nxt=None):
    x, ln, st, qkv, qu, qv, pp, pr, pd, c, da, do, scale, Pn, ld_bd, fb = saved
    d, H = cfg.d, cfg.heads
    a = p + ".self_attn"
    k, v = qkv[:, d:2 * d], qkv[:, 2 * d:]
    g = _head_g(dy, g, (1.0, do, G[a + ".linear_out.bias"]))
    K.gemm(g, c, trans=True, out=G[a + ".linear_out.weight"], accumulate=True)
    dc = K.gemm(g, W[a + ".linear_out.weight"], trans_b=True)
    dqkv = torch.empty_like(qkv)
    dpp = _zeros(Pn * d, x.device).view(Pn, d)
    if fb:
        # fused backward (keys <= 192): dP, dS, dropout(p) and d(matrix_bd) stay on the SM; dq_u, dq_v, dk, dv and the
        # gradient of the projected positions come out of ONE kernel
        dqu = torch.empty(B * T, d, device=x.device, dtype=dc.dtype)
        dqv = torch.empty_like(dqu)
        # ... and the K / V parts of linear_qkv's bias gradient (column sums of dk, dv) on the way out of tensor memory
        gb = G[a + ".linear_qkv.bias"]
        K.relattn_bwd(dc, qu, k, v, pr, c, B, T, T, H, cfg.rel_pos_type, scale, dqu, dqkv[:, d:2 * d], dqkv[:, 2 * d:], drop=da,
                      qv=qv, pos=pp, dqv=dqv, dpos=dpp, dk_colsum=gb[d:2 * d], dv_colsum=gb[2 * d:])
    else:
        dpd = K.attn_f1(dc, v, B, T, T, H)
        K.attn_f3(pd, dc, B, T, T, H, out=dqkv[:, 2 * d:])
        dac, dbd = K.softmax_bwd(dpd, pr, T, cfg.rel_pos_type, scale, ld_bd=ld_bd, drop=da)
        dqu = K.attn_f2(dac, k, B, T, T, H)
        K.attn_f3(dac, qu, B, T, T, H, out=dqkv[:, d:2 * d])
        dqv = K.attn_f2(dbd, pp, B, T, Pn, H, y_bcast=True)
        K.attn_f4(dbd, qv, B, T, Pn, H, dpp)
    K.gemm(K.cast_bf16(dpp), pos16, trans=True, out=G[a + ".linear_pos.weight"], accumulate=True)
    if fb:  # colsum(dq) = colsum(dq_u) + colsum(dq_v) comes with the position-bias gradients: no pass over dqkv
        K.pos_bias_bwd(dqu, dqv, dqkv[:, :d], G[a + ".pos_bias_u"].reshape(-1), G[a + ".pos_bias_v"].reshape(-1),
                       dq_colsum=G[a + ".linear_qkv.bias"][:d])
    else:
        K.pos_bias_bwd(dqu, dqv, dqkv[:, :d], G[a + ".pos_bias_u"].reshape(-1), G[a + ".pos_bias_v"].reshape(-1))
        K.colsum(dqkv, G[a + ".linear_qkv.bias"])
    K.gemm(dqkv, ln, trans=True, out=G[a + ".linear_qkv.weight"], accumulate=True)
    dln = K.gemm(dqkv, W[a + ".linear_qkv.weight"], trans_b=True)
    return _ln_bwd(dln, x, P, G, norm, st, dy, nxt)


# ------------------------------------------------------------------------------------------------------------------
# convolution module branch
# ------------------------------------------------------------------------------------------------------------------
def convmod_fwd(x, P, W, p: str, norm: str, cfg: EncCfg, B: int, T: int, ctx: StepCtx, bn_buffers=None):
    cm = p + ".conv_module"
    ln, _, st = K.layernorm_fwd(x, P[norm + ".weight"], P[norm + ".bias"])
    pw1 = K.gemm(ln, W[cm + ".pointwise_conv1.weight"], bias=P[cm + ".pointwise_conv1.bias"])
    wdw = P[cm + ".depthwise_conv.weight"].reshape(cfg.d, cfg.kernel)
    dw, sums = K.glu_dwconv_fwd(pw1, wdw, P[cm + ".depthwise_conv.bias"], B, T, cfg.kernel, sums=_zeros(2 * cfg.d, x.device))
    rm = rv = nb = None
    if bn_buffers is not None and ctx.training:
        rm, rv, nb = bn_buffers[cm + ".norm.running_mean"], bn_buffers[cm + ".norm.running_var"], \
            bn_buffers[cm + ".norm.num_batches_tracked"]
    if not ctx.training and bn_buffers is not None:
        # eval mode: running statistics (convolution.py:45 BatchNorm1d in eval)
        mr = torch.cat([bn_buffers[cm + ".norm.running_mean"].float(),
                        torch.rsqrt(bn_buffers[cm + ".norm.running_var"].float() + 1e-5)]).contiguous()
    else:
        mr = K.bn_finalize(sums, B * T, 1e-5, 0.1, rm, rv, nb)
    bs = K.bn_swish_fwd(dw, mr, P[cm + ".norm.weight"], P[cm + ".norm.bias"])
    do = ctx.drop(cfg.dropout, cm + ".out")
    y = K.gemm(bs, W[cm + ".pointwise_conv2.weight"], bias=P[cm + ".pointwise_conv2.bias"], residual=x, drop=do,
               out_dtype=F32)
    return y, (x, ln, st, pw1, dw, mr, bs, do)


def convmod_head(saved, G, p: str):
    return 1.0, saved[7], G[p + ".conv_module.pointwise_conv2.bias"]


def convmod_bwd(dy, saved, P, W, G, p: str, norm: str, cfg: EncCfg, B: int, T: int, g=None, nxt=None):
    x, ln, st, pw1, dw, mr, bs, do = saved
    cm = p + ".conv_module"
    g = _head_g(dy, g, (1.0, do, G[cm + ".pointwise_conv2.bias"]))
    K.gemm(g, bs, trans=True, out=G[cm + ".pointwise_conv2.weight"], accumulate=True)
    dbs = K.gemm(g, W[cm + ".pointwise_conv2.weight"], trans_b=True)
    ddw = K.bn_swish_bwd(dbs, dw, mr, P[cm + ".norm.weight"], P[cm + ".norm.bias"], G[cm + ".norm.weight"],
                         G[cm + ".norm.bias"], scratch=_zeros(2 * cfg.d, dy.device))
    wdw = P[cm + ".depthwise_conv.weight"].reshape(cfg.d, cfg.kernel)
    # ... which also accumulates pointwise_conv1's bias gradient (the column sums of d_pw1)
    dpw1 = K.glu_dwconv_bwd(ddw, pw1, wdw, B, T, cfg.kernel, G[cm + ".depthwise_conv.weight"].reshape(cfg.d, cfg.kernel),
                            G[cm + ".depthwise_conv.bias"], dpw1_colsum=G[cm + ".pointwise_conv1.bias"])
    K.gemm(dpw1, ln, trans=True, out=G[cm + ".pointwise_conv1.weight"], accumulate=True)
    dln = K.gemm(dpw1, W[cm + ".pointwise_conv1.weight"], trans_b=True)
    return _ln_bwd(dln, x, P, G, norm, st, dy, nxt)


# ------------------------------------------------------------------------------------------------------------------
# one Conformer block
# ------------------------------------------------------------------------------------------------------------------
def enc_layer_fwd(x, pos16, klens, P, W, p: str, cfg: EncCfg, B: int, T: int, ctx: StepCtx, bn_buffers=None):
    x, s1 = ffn_fwd(x, P, W, p + ".feed_forward_macaron", p + ".norm_ff_macaron", "swish", 0.5, cfg.dropout, ctx)
    x, s2 = relmha_fwd(x, pos16, klens, P, W, p, p + ".norm_mha", cfg, B, T, ctx)
    x, s3 = convmod_fwd(x, P, W, p, p + ".norm_conv", cfg, B, T, ctx, bn_buffers)
    x, s4 = ffn_fwd(x, P, W, p + ".feed_forward", p + ".norm_ff", "swish", 0.5, cfg.dropout, ctx)
    _, y, st = K.layernorm_fwd(x, P[p + ".norm_final.weight"], P[p + ".norm_final.bias"], want_bf16=False, want_f32=True)
    return y, (s1, s2, s3, s4, x, st)


def enc_layer_bwd(dy, saved, pos16, P, W, G, p: str, cfg: EncCfg, B: int, T: int, nxt=None):
    """Returns (dx, g): g is the ready-made branch gradient for the head ``nxt`` of whatever consumes dx (or None)."""
    s1, s2, s3, s4, x4, st = saved
    dx, g = _ln_bwd(dy, x4, P, G, p + ".norm_final", st, None, ffn_head(s4, G))
    dx, g = ffn_bwd(dx, s4, P, W, G, g, convmod_head(s3, G, p))
    dx, g = convmod_bwd(dx, s3, P, W, G, p, p + ".norm_conv", cfg, B, T, g, relmha_head(s2, G, p))
    dx, g = relmha_bwd(dx, s2, pos16, P, W, G, p, p + ".norm_mha", cfg, B, T, g, ffn_head(s1, G))
    return ffn_bwd(dx, s1, P, W, G, g, nxt)


# ------------------------------------------------------------------------------------------------------------------
# encoder: Conv2dSubsampling + rel-pos encoding + N blocks + after_norm
# ------------------------------------------------------------------------------------------------------------------
def encoder_fwd(feats, flens, P, W, cfg: EncCfg, ctx: StepCtx, p: str = "encoder", bn_buffers=None,
                want_inter_f32: bool = False):
    """feats f32 [B, T_in, idim], flens int32 [B] (device).  Returns (enc f32 [B*T, d], enc16 bf16, olens int32, saved).
    saved[11] lists the intermediate-CTC taps (block idx, block output, LN stats, bf16 tap, fp32 tap or None)."""
    B, Tin, _ = feats.shape
    e = p + ".embed"
    C_ = P[e + ".conv.0.weight"].shape[0]
    Tin_, F_ = feats.shape[1], feats.shape[2]
    implicit = K.conv2_implicit_supported(B, (Tin_ - 3) // 2 + 1, (F_ - 3) // 2 + 1, C_)
    h1 = K.conv1_fwd(feats, P[e + ".conv.0.weight"].reshape(C_, 9), P[e + ".conv.0.bias"], want_bits=implicit)
    h1, h1_bits = h1 if implicit else (h1, None)
    _, T1, F1, _ = h1.shape
    T, F2 = (T1 - 3) // 2 + 1, (F1 - 3) // 2 + 1
    if implicit:
        # implicit GEMM: the operand loads walk h1 through TMA im2col maps, no [pixels, 9 C] matrix anywhere; the backward
        # gets h1's ReLU mask as bit rows
        col = h1_bits
        h2 = K.conv2_fwd(h1, W[e + ".conv.2.weight"], P[e + ".conv.2.bias"])
    else:
        col = K.im2col(h1)
        h2 = K.gemm(col, W[e + ".conv.2.weight"], bias=P[e + ".conv.2.bias"], act="relu")
        # the im2col matrix (9x the conv.0 output: 27 MB per utterance of 5 s) is kept for the wgrad GEMM while it fits
        # comfortably (<= 8 GB of the 180 GB), else dropped here and rebuilt in the backward
        if col.numel() * 2 > KEEP_IM2COL_BYTES:
            col = None
    h2v = h2.view(B * T, F2 * C_)
    dpos = ctx.drop(cfg.pos_dropout, e + ".x")
    x = K.gemm(h2v, W[e + ".out.0.weight"], bias=P[e + ".out.0.bias"], alpha=math.sqrt(cfg.d), drop=dpos, out_dtype=F32)
    # RelPositionalEncoding: table rows + dropout (embedding.py:228-244 / 156-170)
    Pn = 2 * T - 1 if cfg.rel_pos_type == "latest" else T
    dpe = ctx.drop(cfg.pos_dropout, e + ".pos")
    pos16 = K.cast_bf16(pos_table(cfg.rel_pos_type, T if cfg.rel_pos_type == "latest" else Pn, cfg.d, feats.device), drop=dpe)
    olens = torch.clamp((flens + 3) // 4, max=T).to(torch.int32)  # mask[:, :, :-2:2][:, :, :-2:2] (subsampling.py:87)
    layers, inter = [], []
    _zeros_begin(feats.device, cfg.blocks * (4 * cfg.d + (Pn * cfg.d if ctx.training else 0)))
    for n in range(cfg.blocks):
        x, s = enc_layer_fwd(x, pos16, olens, P, W, f"{p}.encoders.{n}", cfg, B, T, ctx, bn_buffers)
        layers.append(s)
        if n + 1 in cfg.interctc_layer_idx:
            # intermediate CTC tap: the block output through after_norm (conformer_encoder.py:339-347)
            i16, i32, sti = K.layernorm_fwd(x, P[p + ".after_norm.weight"], P[p + ".after_norm.bias"], want_bf16=True,
                                            want_f32=want_inter_f32)
            cond = None
            if cfg.interctc_use_conditioning:
                # x += conditioning_layer(softmax(ctc_lo(tap))): the tap's CTC logits are computed here and handed to the
                # CTC head later; the posterior is kept (bf16) for the backward
                V = cfg.vocab
                lg = torch.empty(x.shape[0], K.round8(V), device=x.device, dtype=F32)
                K.gemm(i16, W["ctc.ctc_lo.weight"], bias=P["ctc.ctc_lo.bias"], out=lg)
                sm = K.vocab_softmax_fwd(lg, V)
                x_tap = x
                x = K.gemm(sm[:, :V], W[p + ".conditioning_layer.weight"], bias=P[p + ".conditioning_layer.bias"], residual=x,
                           out_dtype=F32)
                cond = (lg, sm)
                inter.append((n + 1, x_tap, sti, i16, i32, cond))
                continue
            inter.append((n + 1, x, sti, i16, i32, cond))
    enc16, enc, st = K.layernorm_fwd(x, P[p + ".after_norm.weight"], P[p + ".after_norm.bias"], want_bf16=True, want_f32=True)
    return enc, enc16, olens, (feats, h1, h2v, dpos, pos16, layers, x, st, col, B, T, inter)


def encoder_bwd(denc, saved, P, W, G, cfg: EncCfg, p: str = "encoder", on_ready=None, dinter=None):
    """denc f32 [B*T, d]: gradient w.r.t. the encoder output.  Accumulates every encoder parameter gradient into G.
    on_ready("enc<n>") is called when block n's parameter gradients are final (data-parallel bucket hook).
    dinter = {block idx: f32 [B*T, d] gradient w.r.t. that intermediate-CTC tap} (missing / None = no gradient)."""
    feats, h1, h2v, dpos, pos16, layers, xl, st, col, B, T, inter = saved
    taps = {t[0]: t for t in inter}
    e = p + ".embed"
    dx, _ = K.layernorm_bwd(denc, xl, P[p + ".after_norm.weight"], st, G[p + ".after_norm.weight"], G[p + ".after_norm.bias"])
    embed_head = (math.sqrt(cfg.d), dpos, G[e + ".out.0.bias"])
    g = None
    for n in reversed(range(cfg.blocks)):
        tap = taps.get(n + 1)
        dtap = dinter.get(n + 1) if dinter is not None else None
        if tap is not None and tap[5] is not None:
            # self-conditioning backward: dx is the gradient of x + cond(softmax(ctc_lo(tap))); the posterior's gradient
            # goes through the softmax and ctc_lo into the tap, next to the CTC loss's own gradient (dtap)
            V = cfg.vocab
            _, i16, (_, sm) = tap[1], tap[3], tap[5]
            c = p + ".conditioning_layer"
            g = K.cast_bf16(dx, colsum_out=G[c + ".bias"])
            K.gemm(g, sm[:, :V], trans=True, out=G[c + ".weight"], accumulate=True)
            dsm = K.gemm(g, W[c + ".weight"], trans_b=True)
            dlg = K.vocab_softmax_bwd(sm, dsm, V)
            gl = dlg[:, :V]
            K.colsum(dlg, G["ctc.ctc_lo.bias"], n=V)
            K.gemm(gl, i16, trans=True, out=G["ctc.ctc_lo.weight"], accumulate=True)
            dcond = K.gemm(gl, W["ctc.ctc_lo.weight"], trans_b=True, out_dtype=F32)
            dtap = dcond if dtap is None else dtap + dcond
        if tap is not None and dtap is not None:
            # the tap's after_norm backward joins the gradient of block n's output (one pass: dx = LN'(dtap) + dx)
            dx, _ = K.layernorm_bwd(dtap, tap[1], P[p + ".after_norm.weight"], tap[2], G[p + ".after_norm.weight"],
                                    G[p + ".after_norm.bias"], dx_add=dx)
        # block 0 hands its dx to the embedding's output Linear, the others to the previous block's norm_final (fp32 only)
        dx, g = enc_layer_bwd(dx, layers[n], pos16, P, W, G, f"{p}.encoders.{n}", cfg, B, T, embed_head if n == 0 else None)
        layers[n] = None  # free the block's activations as we go
        if on_ready is not None:
            on_ready("enc%d" % n)
    C_ = h1.shape[-1]
    g = _head_g(dx, g, embed_head)
    K.gemm(g, h2v, trans=True, out=G[e + ".out.0.weight"], accumulate=True)
    dh2 = K.gemm(g, W[e + ".out.0.weight"], trans_b=True, act_grad_input=h2v, act_grad="relu").view(-1, C_)
    K.colsum(dh2, G[e + ".conv.2.bias"])
    if col is not None and col.dtype == torch.int32:
        K.conv2_wgrad(dh2, h1, G[e + ".conv.2.weight"])
        dh1 = K.conv2_dgrad_relu(dh2, W[e + ".conv.2.weight"], h1, bits=col)
    else:
        if col is None:
            col = K.im2col(h1)
        K.gemm(dh2, col, trans=True, out=G[e + ".conv.2.weight"], accumulate=True)
        del col
        dcol = K.gemm(dh2, W[e + ".conv.2.weight"], trans_b=True)
        dh1 = K.col2im_relu(dcol, h1)
        del dcol
    K.conv1_bwd(dh1, feats, G[e + ".conv.0.weight"].reshape(C_, 9), G[e + ".conv.0.bias"])


# ------------------------------------------------------------------------------------------------------------------
# decoder
# ------------------------------------------------------------------------------------------------------------------
def _mha_fwd(xq16, kv_src16, q_w: str, kv_w: str, P, W, a: str, klens, B, Tq, Tk, H, d, causal, p_att, ctx, fused_qkv):
    """Plain MultiHeadedAttention (attention.py:98-114) up to (not including) linear_out."""
    if fused_qkv:
        qkv = K.gemm(xq16, W[a + ".linear_qkv.weight"], bias=P[a + ".linear_qkv.bias"])
        q, k, v = qkv[:, :d], qkv[:, d:2 * d], qkv[:, 2 * d:]
        proj = (qkv,)
    else:
        q = K.gemm(xq16, W[a + ".linear_q.weight"], bias=P[a + ".linear_q.bias"])
        kv = K.gemm(kv_src16, W[a + ".linear_kv.weight"], bias=P[a + ".linear_kv.bias"])
        k, v = kv[:, :d], kv[:, d:]
        proj = (q, kv)
    da = ctx.drop(p_att, a + ".att")
    scale = 1.0 / math.sqrt(d // H)
    fb = False
    if K.relattn_supported(Tq, Tk, d // H, "none", causal):
        fb = ctx.training and K.relattn_bwd_supported(Tq, Tk, d // H, "none")
        pr, pd, c = K.relattn_fwd(q, None, k, v, None, klens, B, Tq, Tk, H, "none", causal, scale, drop=da, want_pd=not fb)
    else:
        ac = K.attn_f1(q, k, B, Tq, Tk, H)
        pr, pd = K.softmax_fwd(ac, None, klens, Tk, "none", causal, scale, drop=da)
        c = K.attn_f2(pd, v, B, Tq, Tk, H)
    return c, (proj, pr, pd, da, scale, fb)


def dec_layer_fwd(x, mem16, ylens, mlens, P, W, p: str, cfg: DecCfg, B: int, L: int, T: int, ctx: StepCtx):
    d, H = cfg.d, cfg.heads
    # masked self-attention (decoder_layer.py:88-110)
    ln1, _, st1 = K.layernorm_fwd(x, P[p + ".norm1.weight"], P[p + ".norm1.bias"])
    c1, sa = _mha_fwd(ln1, None, "", "", P, W, p + ".self_attn", ylens, B, L, L, H, d, True, cfg.self_att_dropout, ctx, True)
    do1 = ctx.drop(cfg.dropout, p + ".self_attn.out")
    x1 = K.gemm(c1, W[p + ".self_attn.linear_out.weight"], bias=P[p + ".self_attn.linear_out.bias"], residual=x, drop=do1,
                out_dtype=F32)
    # source attention over the encoder memory (decoder_layer.py:112-123)
    ln2, _, st2 = K.layernorm_fwd(x1, P[p + ".norm2.weight"], P[p + ".norm2.bias"])
    c2, ca = _mha_fwd(ln2, mem16, "", "", P, W, p + ".src_attn", mlens, B, L, T, H, d, False, cfg.src_att_dropout, ctx, False)
    do2 = ctx.drop(cfg.dropout, p + ".src_attn.out")
    x2 = K.gemm(c2, W[p + ".src_attn.linear_out.weight"], bias=P[p + ".src_attn.linear_out.bias"], residual=x1, drop=do2,
                out_dtype=F32)
    # feed-forward, ReLU (decoder_layer.py:125-131)
    x3, sf = ffn_fwd(x2, P, W, p + ".feed_forward", p + ".norm3", "relu", 1.0, cfg.dropout, ctx)
    return x3, (x, ln1, st1, c1, sa, do1, x1, ln2, st2, c2, ca, do2, sf)


def dec_layer_bwd(dy, dmem, saved, mem16, P, W, G, p: str, cfg: DecCfg, B: int, L: int, T: int, g=None, nxt=None):
    """Returns (dx, g for the head ``nxt``); accumulates the memory gradient into dmem (f32 [B*T, d], in place via the GEMM
    residual input)."""
    x, ln1, st1, c1, sa, do1, x1, ln2, st2, c2, ca, do2, sf = saved
    d, H = cfg.d, cfg.heads
    a = p + ".src_attn"
    dx2, g = ffn_bwd(dy, sf, P, W, G, g, (1.0, do2, G[a + ".linear_out.bias"]))
    # ---- source attention
    (q, kv), pr, pd, da, scale, fb = ca
    k, v = kv[:, :d], kv[:, d:]
    K.gemm(g, c2, trans=True, out=G[a + ".linear_out.weight"], accumulate=True)
    dc = K.gemm(g, W[a + ".linear_out.weight"], trans_b=True)
    dkv = torch.empty_like(kv)
    if fb:
        dq = torch.empty_like(q)
        K.relattn_bwd(dc, q, k, v, pr, c2, B, L, T, H, "none", scale, dq, dkv[:, :d], dkv[:, d:], drop=da)
    else:
        dpd = K.attn_f1(dc, v, B, L, T, H)
        K.attn_f3(pd, dc, B, L, T, H, out=dkv[:, d:])
        dac, _ = K.softmax_bwd(dpd, pr, T, "none", scale, drop=da)
        dq = K.attn_f2(dac, k, B, L, T, H)
        K.attn_f3(dac, q, B, L, T, H, out=dkv[:, :d])
    K.colsum(dkv, G[a + ".linear_kv.bias"])
    K.gemm(dkv, mem16, trans=True, out=G[a + ".linear_kv.weight"], accumulate=True)
    K.gemm(dkv, W[a + ".linear_kv.weight"], trans_b=True, residual=dmem, out=dmem)
    K.colsum(dq, G[a + ".linear_q.bias"])
    K.gemm(dq, ln2, trans=True, out=G[a + ".linear_q.weight"], accumulate=True)
    dln2 = K.gemm(dq, W[a + ".linear_q.weight"], trans_b=True)
    a = p + ".self_attn"
    dx1, g = _ln_bwd(dln2, x1, P, G, p + ".norm2", st2, dx2, (1.0, do1, G[a + ".linear_out.bias"]))
    # ---- self attention
    (qkv,), pr, pd, da, scale, fb = sa
    q, k, v = qkv[:, :d], qkv[:, d:2 * d], qkv[:, 2 * d:]
    K.gemm(g, c1, trans=True, out=G[a + ".linear_out.weight"], accumulate=True)
    dc = K.gemm(g, W[a + ".linear_out.weight"], trans_b=True)
    dqkv = torch.empty_like(qkv)
    if fb:
        K.relattn_bwd(dc, q, k, v, pr, c1, B, L, L, H, "none", scale, dqkv[:, :d], dqkv[:, d:2 * d], dqkv[:, 2 * d:], drop=da)
    else:
        dpd = K.attn_f1(dc, v, B, L, L, H)
        K.attn_f3(pd, dc, B, L, L, H, out=dqkv[:, 2 * d:])
        dac, _ = K.softmax_bwd(dpd, pr, L, "none", scale, drop=da)
        K.attn_f2(dac, k, B, L, L, H, out=dqkv[:, :d])
        K.attn_f3(dac, q, B, L, L, H, out=dqkv[:, d:2 * d])
    K.colsum(dqkv, G[a + ".linear_qkv.bias"])
    K.gemm(dqkv, ln1, trans=True, out=G[a + ".linear_qkv.weight"], accumulate=True)
    dln1 = K.gemm(dqkv, W[a + ".linear_qkv.weight"], trans_b=True)
    return _ln_bwd(dln1, x, P, G, p + ".norm1", st1, dx1, nxt)


def decoder_fwd(mem16, mlens, ys_in, ys_in_lens, P, W, cfg: DecCfg, ctx: StepCtx, B: int, T: int, p: str = "decoder"):
    """mem16 bf16 [B*T, d]; ys_in int32 [B, L]; returns (logits f32 [B*L, round8(V)] (columns >= V unspecified), saved)."""
    L = ys_in.shape[1]
    tok = ys_in.reshape(-1)
    de = ctx.drop(cfg.pos_dropout, p + ".embed")
    x = K.embed_fwd(P[p + ".embed.0.weight"], tok, pos_table("abs", L, cfg.d, mem16.device), L, math.sqrt(cfg.d), drop=de)
    layers = []
    for n in range(cfg.blocks):
        x, s = dec_layer_fwd(x, mem16, ys_in_lens, mlens, P, W, f"{p}.decoders.{n}", cfg, B, L, T, ctx)
        layers.append(s)
    ln, _, st = K.layernorm_fwd(x, P[p + ".after_norm.weight"], P[p + ".after_norm.bias"])
    V = cfg.vocab
    logits = torch.empty(B * L, K.round8(V), device=mem16.device, dtype=F32)
    K.gemm(ln, W[p + ".output_layer.weight"], bias=P[p + ".output_layer.bias"], out=logits)
    return logits, (tok, de, layers, x, ln, st, L)


def decoder_bwd(dlogits16, dmem, saved, mem16, P, W, G, cfg: DecCfg, B: int, T: int, p: str = "decoder"):
    """dlogits16 bf16 [B*L, round8(V)]; dmem f32 [B*T, d] is accumulated in place."""
    tok, de, layers, xl, ln, st, L = saved
    V = cfg.vocab
    gl = dlogits16[:, :V]
    K.colsum(dlogits16, G[p + ".output_layer.bias"], n=V)
    K.gemm(gl, ln, trans=True, out=G[p + ".output_layer.weight"], accumulate=True)
    dln = K.gemm(gl, W[p + ".output_layer.weight"], trans_b=True)
    dx, g = _ln_bwd(dln, xl, P, G, p + ".after_norm", st, None, ffn_head(layers[-1][12], G))
    for n in reversed(range(cfg.blocks)):
        nxt = ffn_head(layers[n - 1][12], G) if n > 0 else None  # the block below starts its backward with its FFN branch
        dx, g = dec_layer_bwd(dx, dmem, layers[n], mem16, P, W, G, f"{p}.decoders.{n}", cfg, B, L, T, g, nxt)
        layers[n] = None
    K.embed_bwd(dx, tok, math.sqrt(cfg.d), G[p + ".embed.0.weight"], drop=de)


# ------------------------------------------------------------------------------------------------------------------
# heads: CTC and label-smoothing attention loss (each computes its own input gradient in the same pass)
# ------------------------------------------------------------------------------------------------------------------
def ctc_head(enc16, olens, labels, label_lens, P, W, G, V: int, B: int, T: int, weight: float, p: str = "ctc",
             want_grad: bool = True, logits=None):
    """CTC.forward (espnet2/asr/ctc.py:72-97) + its backward.  Returns (loss_ctc scalar tensor = sum/B, denc f32 or None).
    ``weight`` = d(total loss)/d(loss_ctc); the gradient written is already scaled by it.  ``logits``: ctc_lo(enc16) when
    the caller has it already (self-conditioned taps)."""
    if logits is None:
        logits = torch.empty(B * T, K.round8(V), device=enc16.device, dtype=F32)
        K.gemm(enc16, W[p + ".ctc_lo.weight"], bias=P[p + ".ctc_lo.bias"], out=logits)
    l3 = logits.view(B, T, -1)
    _, loss_sum, ws = K.ctc_fwd(l3, V, olens, labels, label_lens, blank=0)
    loss = loss_sum / B
    if not want_grad:
        return loss, None
    _, g16 = K.ctc_bwd(l3, V, olens, labels, label_lens, ws, weight / B, blank=0)
    g16 = g16.view(B * T, -1)
    gl = g16[:, :V]
    K.colsum(g16, G[p + ".ctc_lo.bias"], n=V)
    K.gemm(gl, enc16, trans=True, out=G[p + ".ctc_lo.weight"], accumulate=True)
    denc = K.gemm(gl, W[p + ".ctc_lo.weight"], trans_b=True, out_dtype=F32)
    return loss, denc


def att_head(logits, ys_out, V: int, B: int, smoothing: float, weight: float):
    """LabelSmoothingLoss + th_accuracy.  Returns (loss_att scalar tensor, acc scalar tensor, dlogits bf16)."""
    stats, _, g16 = K.lsm_loss(logits, V, ys_out.reshape(-1), smoothing, weight / B)
    return stats[0:1] / B, stats[1:2] / stats[2:3], g16


# ------------------------------------------------------------------------------------------------------------------
# whole step from encoder input on (espnet_model.py:202-297): loss + all parameter gradients
# ------------------------------------------------------------------------------------------------------------------
@dataclass
class ModelCfg:
    enc: EncCfg = field(default_factory=EncCfg)
    dec: DecCfg = field(default_factory=DecCfg)
    vocab: int = 600
    ctc_weight: float = 0.3
    lsm_weight: float = 0.1
    n_fft: int = 512
    hop: int = 128
    n_mels: int = 80
    interctc_weight: float = 0.0  # espnet_model.py:223-245; the taps are cfg.enc.interctc_layer_idx


def add_sos_eos(text, text_lens, sos: int, eos: int, ignore: int = -1):
    """add_sos_eos.py:12-31 without a host loop: text int [B, Lmax] (padding arbitrary), text_lens [B] -> ys_in, ys_out int32 [B, Lmax+1]."""
    B, Lm = text.shape
    ar = torch.arange(Lm + 1, device=text.device)[None, :]
    tl = text_lens[:, None].to(torch.long)
    t = text.to(torch.long)
    shifted = torch.cat([torch.full((B, 1), sos, device=text.device, dtype=torch.long), t], 1)
    ys_in = torch.where(ar <= tl, shifted, torch.full_like(shifted, eos))
    padded = torch.cat([t, torch.full((B, 1), eos, device=text.device, dtype=torch.long)], 1)
    ys_out = torch.where(ar < tl, padded, torch.where(ar == tl, torch.full_like(padded, eos), torch.full_like(padded, ignore)))
    return ys_in.to(torch.int32).contiguous(), ys_out.to(torch.int32).contiguous()


def train_step_from_feats(feats, flens, text, text_lens, P, W, G, cfg: ModelCfg, ctx: StepCtx, bn_buffers=None,
                          grad_scale: float = 1.0, on_ready=None, want_grad: bool = True):
    """feats f32 [B,T_in,idim] (already normalised), flens int32 [B], text int32 [B,Lmax], text_lens int32 [B] -- all on the
    device.  Runs forward and backward; parameter gradients (times grad_scale) are ACCUMULATED into G.
    Returns dict(loss, loss_ctc, loss_att, acc) of 1-element device tensors (no synchronisation)."""
    B = feats.shape[0]
    V = cfg.vocab
    enc, enc16, olens, s_enc = encoder_fwd(feats, flens, P, W, cfg.enc, ctx, bn_buffers=bn_buffers)
    T, inter = s_enc[10], s_enc[11]
    wc, wa = cfg.ctc_weight * grad_scale, (1.0 - cfg.ctc_weight) * grad_scale
    text = text.to(torch.int32).contiguous()
    tl = text_lens.to(torch.int32)
    # intermediate CTC (espnet_model.py:223-245): same CTC module, same targets, losses averaged over the taps and mixed
    # into the CTC term with interctc_weight; every head writes its own, already weighted, input gradient
    iw = cfg.interctc_weight if inter else 0.0
    loss_ctc, denc = ctc_head(enc16, olens, text, tl, P, W, G, V, B, T, wc * (1.0 - iw), want_grad=want_grad)
    stats_ic, dinter, loss_ic = {}, {}, None
    if iw != 0.0:
        for idx, _, _, i16, _, cond in inter:
            l, dinter[idx] = ctc_head(i16, olens, text, tl, P, W, G, V, B, T, wc * iw / len(inter), want_grad=want_grad,
                                      logits=cond[0] if cond is not None else None)
            stats_ic[f"loss_interctc_layer{idx}"] = l
            loss_ic = l if loss_ic is None else loss_ic + l
    ys_in, ys_out = add_sos_eos(text, tl, V - 1, V - 1)
    ylens = (tl + 1).to(torch.int32)
    logits, s_dec = decoder_fwd(enc16, olens, ys_in, ylens, P, W, cfg.dec, ctx, B, T)
    loss_att, acc, dlogits = att_head(logits, ys_out, V, B, cfg.lsm_weight, wa)
    if want_grad:
        decoder_bwd(dlogits, denc, s_dec, enc16, P, W, G, cfg.dec, B, T)
        del s_dec
        if on_ready is not None:
            on_ready("heads")  # decoder + CTC head gradients are final
        encoder_bwd(denc, s_enc, P, W, G, cfg.enc, on_ready=on_ready, dinter=dinter)
    # stats["loss_ctc"] is the final block's CTC loss alone (espnet_model.py:217, recorded before the mix)
    mixed = loss_ctc if loss_ic is None else (1.0 - iw) * loss_ctc + iw * (loss_ic / len(inter))
    loss = cfg.ctc_weight * mixed + (1.0 - cfg.ctc_weight) * loss_att
    return {"loss": loss, "loss_ctc": loss_ctc, "loss_att": loss_att, "acc": acc, **stats_ic}


def frontend_fwd(speech, speech_lens, melmat, cfg: ModelCfg, draws: Optional[dict] = None):
    """DefaultFrontend -> SpecAug (draws: dict of int32 device arrays or None) -> UtteranceMVN (espnet_model.py:319-338)."""
    feats, flens = K.logmel(speech, speech_lens.to(torch.int32), melmat, cfg.n_fft, cfg.hop)
    d = draws or {}
    feats = K.specaug_mvn(feats, flens, d.get("center"), d.get("warped"), d.get("fmask"), d.get("tmask"), do_mvn=True)
    return feats, flens
