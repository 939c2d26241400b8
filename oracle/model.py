"""Functional CPU restatement of the ESPnet2 Conformer CTC/attention train step.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Everything here is a function of
``(state_dict, inputs)``: parameters are looked up by the *reference's*
state_dict key names so that the same weights drive the reference, this oracle
and the CUDA path.  All arithmetic is plain torch on CPU in the dtype of the
inputs (fp32 by default, fp64 for tight checks); gradients come from torch
autograd over these functions, so only the forward is restated.

Reference paths are relative to /root/reference.
"""
from __future__ import annotations

import functools
import math
from dataclasses import dataclass, field
from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]

# False (default, what every parity test runs): each op spelled out with elementary tensor arithmetic, so that the
# restatement does not lean on the library kernels whose behaviour it pins.  True: LayerNorm, Linear, the depthwise
# convolution, BatchNorm, GLU, the STFT and the bicubic time-warp resize go through the SAME torch library ops the
# reference's nn.Modules call (F.layer_norm, F.linear, F.conv1d(groups), F.batch_norm, F.glu, torch.stft,
# F.interpolate) -- used by bench.py's same-box leg, which times the reference's eager PyTorch step (cuBLAS / cuDNN /
# ATen kernels) on the B200.  tests/test_oracle_golden.py checks that both settings agree.
LIBRARY_OPS = False


# --------------------------------------------------------------------------- #
# configuration
# --------------------------------------------------------------------------- #
@dataclass
class EncCfg:
    """ConformerEncoder hyper-parameters (espnet2/asr/encoder/conformer_encoder.py:82-111)."""

    idim: int = 80
    d: int = 512
    heads: int = 8
    ffn: int = 2048
    blocks: int = 12
    kernel: int = 31
    rel_pos_type: str = "latest"  # or "legacy"


@dataclass
class DecCfg:
    """TransformerDecoder hyper-parameters (espnet2/asr/decoder/transformer_decoder.py:232-280)."""

    vocab: int = 600
    d: int = 512
    heads: int = 8
    ffn: int = 2048
    blocks: int = 6


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
    fs: int = 16000
    interctc_weight: float = 0.0
    interctc_layer_idx: tuple = ()  # 1-based block indices (conformer_encoder.py:283-285)
    interctc_use_conditioning: bool = False


# --------------------------------------------------------------------------- #
# small helpers
# --------------------------------------------------------------------------- #
def pad_mask(lens: torch.Tensor, maxlen: int) -> torch.Tensor:
    """True where PADDED.  espnet/nets/pytorch_backend/nets_utils.py:64-176."""
    ar = torch.arange(maxlen, device=lens.device)
    return ar[None, :] >= lens[:, None]


def layer_norm(x: torch.Tensor, sd: SD, p: str, eps: float = 1e-12) -> torch.Tensor:
    """espnet/nets/pytorch_backend/transformer/layer_norm.py:12-38 (eps=1e-12 at :23)."""
    if LIBRARY_OPS:
        return F.layer_norm(x, x.shape[-1:], sd[p + ".weight"], sd[p + ".bias"], eps)
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * sd[p + ".weight"] + sd[p + ".bias"]


def linear(x: torch.Tensor, sd: SD, p: str) -> torch.Tensor:
    if LIBRARY_OPS:
        return F.linear(x, sd[p + ".weight"], sd.get(p + ".bias"))
    y = x @ sd[p + ".weight"].t()
    if p + ".bias" in sd:
        y = y + sd[p + ".bias"]
    return y


def swish(x: torch.Tensor) -> torch.Tensor:
    """espnet/nets/pytorch_backend/conformer/swish.py:13-18."""
    return x * torch.sigmoid(x)


# --------------------------------------------------------------------------- #
# rel_shift index maps (bit-exact integer work)
# --------------------------------------------------------------------------- #
def rel_shift_index_latest(T: int) -> np.ndarray:
    """Column of the un-shifted (T, 2T-1) matrix that lands at out[i, j].

    espnet/nets/pytorch_backend/transformer/attention.py:240-263:
    out[i, j] = bd[i, T-1-i+j].
    """
    i = np.arange(T)[:, None]
    j = np.arange(T)[None, :]
    return (T - 1 - i + j).astype(np.int64)


def rel_shift_index_legacy(T: int) -> Tuple[np.ndarray, np.ndarray]:
    """Legacy rel_shift (attention.py:145-165) as (row, col) gather maps into the
    un-shifted (T, T) matrix; col == -1 marks the literal zero at j == i+1.

    j <= i   -> bd[i,   T-1-(i-j)]
    j == i+1 -> 0
    j >  i+1 -> bd[i+1, j-i-2]
    """
    i = np.arange(T)[:, None] + np.zeros((1, T), dtype=np.int64)
    j = np.arange(T)[None, :] + np.zeros((T, 1), dtype=np.int64)
    row = np.where(j <= i, i, i + 1)
    col = np.where(j <= i, T - 1 - (i - j), j - i - 2)
    col = np.where(j == i + 1, -1, col)
    row = np.where(j == i + 1, 0, row)
    return row.astype(np.int64), col.astype(np.int64)


# --------------------------------------------------------------------------- #
# frontend: STFT -> power -> log-mel   (fp32 island)
# --------------------------------------------------------------------------- #
def slaney_mel_matrix(fs: int = 16000, n_fft: int = 512, n_mels: int = 80,
                      fmin: float = 0.0, fmax: Optional[float] = None) -> np.ndarray:
    """(n_fft//2+1, n_mels) Slaney-scale, area-normalised mel filterbank.

    The reference takes this constant from librosa.filters.mel(htk=False)
    (espnet2/layers/log_mel.py:36-52); librosa (>=0.8, setup.py:18) is absent
    here, so this restates its published algorithm.  "mel matrix parity vs
    librosa: unpinned" (no reference test stores values); it is cross-checked
    against torchaudio.functional.melscale_fbanks(norm='slaney',
    mel_scale='slaney') in tests/test_oracle_golden.py.
    """
    fmax = fs / 2 if fmax is None else fmax
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = math.log(6.4) / 27.0

    def hz2mel(f):
        f = np.asarray(f, dtype=np.float64)
        lin = f / f_sp
        log = min_log_mel + np.log(np.maximum(f, 1e-30) / min_log_hz) / logstep
        return np.where(f >= min_log_hz, log, lin)

    def mel2hz(m):
        m = np.asarray(m, dtype=np.float64)
        lin = f_sp * m
        log = min_log_hz * np.exp(logstep * (m - min_log_mel))
        return np.where(m >= min_log_mel, log, lin)

    fft_f = np.linspace(0, fs / 2, n_fft // 2 + 1)
    mel_f = mel2hz(np.linspace(hz2mel(fmin), hz2mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fft_f[None, :]
    w = np.zeros((n_mels, n_fft // 2 + 1))
    for m in range(n_mels):
        lower = -ramps[m] / fdiff[m]
        upper = ramps[m + 2] / fdiff[m + 1]
        w[m] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    w *= enorm[:, None]
    return w.T.astype(np.float32)


def frontend_logmel(wav: torch.Tensor, lens: torch.Tensor, melmat: torch.Tensor,
                    n_fft: int = 512, hop: int = 128) -> Tuple[torch.Tensor, torch.Tensor]:
    """DefaultFrontend.forward (espnet2/asr/frontend/default.py:82-117).

    Stft (espnet2/layers/stft.py:83-170): centre reflect-pad n_fft//2, periodic
    Hann window, onesided rFFT, frames past olens zeroed; power re^2+im^2
    (default.py:110); LogMel (espnet2/layers/log_mel.py:62-79): matmul, clamp
    1e-10, natural log, padded frames -> 0.
    """
    B, N = wav.shape
    pad = n_fft // 2
    if LIBRARY_OPS:  # stft.py:94-105: torch.stft with the same arguments
        spec = torch.stft(wav, n_fft, hop, n_fft, torch.hann_window(n_fft, dtype=wav.dtype, device=wav.device), center=True,
                          pad_mode="reflect", normalized=False, onesided=True, return_complex=True).transpose(1, 2)
        olens = (lens + 2 * pad - n_fft) // hop + 1
        padm = pad_mask(olens, spec.shape[1])[:, :, None]
        power = (spec.real ** 2 + spec.imag ** 2).masked_fill(padm, 0.0)
        mel = torch.clamp(power @ melmat.to(power.dtype), min=1e-10).log()
        return mel.masked_fill(padm, 0.0), olens
    x = F.pad(wav[:, None, :], (pad, pad), mode="reflect")[:, 0]
    T_in = N // hop + 1
    idx = torch.arange(T_in, device=wav.device)[:, None] * hop + torch.arange(n_fft, device=wav.device)[None, :]
    frames = x[:, idx]  # (B, T_in, n_fft)
    n = torch.arange(n_fft, dtype=wav.dtype, device=wav.device)
    window = 0.5 - 0.5 * torch.cos(2 * math.pi * n / n_fft)  # torch.hann_window(periodic)
    spec = torch.fft.rfft(frames * window, dim=-1)
    olens = (lens + 2 * pad - n_fft) // hop + 1
    padm = pad_mask(olens, T_in)[:, :, None]
    re = spec.real.masked_fill(padm, 0.0)
    im = spec.imag.masked_fill(padm, 0.0)
    power = re ** 2 + im ** 2
    mel = torch.clamp(power @ melmat.to(power.dtype), min=1e-10).log()
    return mel.masked_fill(padm, 0.0), olens


# --------------------------------------------------------------------------- #
# SpecAug + UtteranceMVN
# --------------------------------------------------------------------------- #
def _cubic_w(t: torch.Tensor, A: float = -0.75):
    """Cubic-convolution weights (Keys, A=-0.75) as used by torch's bicubic
    F.interpolate, which espnet2/layers/time_warp.py:33-38 calls."""
    def c1(x):  # |x| <= 1
        return ((A + 2) * x - (A + 3)) * x * x + 1

    def c2(x):  # 1 < |x| < 2
        return ((A * x - 5 * A) * x + 8 * A) * x - 4 * A

    return c2(t + 1), c1(t), c1(1 - t), c2(2 - t)


def cubic_resize_time(x: torch.Tensor, out_len: int) -> torch.Tensor:
    """(T, F) -> (out_len, F): the time axis of F.interpolate(mode='bicubic',
    align_corners=False) when the freq size is unchanged (then the freq axis is
    the identity: source coordinate j+0.5-0.5=j, weights (0,1,0,0))."""
    if LIBRARY_OPS:  # time_warp.py:33-38
        return F.interpolate(x[None, None], (out_len, x.shape[1]), mode="bicubic", align_corners=False)[0, 0]
    in_len = x.shape[0]
    scale = in_len / out_len
    dst = torch.arange(out_len, dtype=x.dtype, device=x.device)
    src = (dst + 0.5) * scale - 0.5
    i0 = torch.floor(src)
    t = (src - i0)[:, None]
    i0 = i0.long()
    w = _cubic_w(t)
    out = 0
    for k in range(4):
        idx = torch.clamp(i0 - 1 + k, 0, in_len - 1)
        out = out + w[k] * x[idx]
    return out


def time_warp_one(x: torch.Tensor, center: int, warped: int) -> torch.Tensor:
    """espnet2/layers/time_warp.py:9-46 for one un-padded utterance (T, F) with the
    two RNG draws given (center ~ randint(W, T-W), warped ~ randint(c-W, c+W)+1)."""
    left = cubic_resize_time(x[:center], warped)
    right = cubic_resize_time(x[center:], x.shape[0] - warped)
    return torch.cat([left, right], 0)


def spec_augment(feats: torch.Tensor, lens: torch.Tensor, draws: dict, window: int = 5) -> torch.Tensor:
    """SpecAug.forward (espnet2/asr/specaug/specaug.py:89-96) with the RNG draws
    supplied: draws = {center (B,), warped (B,), f_len (B,nf), f_pos (B,nf),
    t_len (B,nt), t_pos (B,nt)}; center < 0 means "no warp for this utterance"
    (time_warp.py:24-25: t - W <= W); center None means the warp stage is off.

    TimeWarp (time_warp.py:65-88) runs per utterance on the UNPADDED slice and
    re-pads with 0; mask_along_axis (mask_along_axis.py:8-68) zeroes
    [pos, pos+len) along freq then along the padded time axis.
    """
    B, T, Fq = feats.shape
    if draws.get("center") is None:  # apply_time_warp=False: no per-utterance slicing, the padding passes through
        out = feats.clone()
    else:
        out = torch.zeros_like(feats)
        for b in range(B):
            L = int(lens[b])
            xb = feats[b, :L]
            c, w = int(draws["center"][b]), int(draws["warped"][b])
            if c >= 0:
                xb = time_warp_one(xb, c, w)
            out[b, :L] = xb
    ar_f = torch.arange(Fq, device=feats.device)[None, None, :]
    fm = ((draws["f_pos"][:, :, None] <= ar_f) & (ar_f < (draws["f_pos"] + draws["f_len"])[:, :, None])).any(1)
    out = out.masked_fill(fm[:, None, :], 0.0)
    ar_t = torch.arange(T, device=feats.device)[None, None, :]
    tm = ((draws["t_pos"][:, :, None] <= ar_t) & (ar_t < (draws["t_pos"] + draws["t_len"])[:, :, None])).any(1)
    out = out.masked_fill(tm[:, :, None], 0.0)
    return out


def utterance_mvn(x: torch.Tensor, lens: torch.Tensor) -> torch.Tensor:
    """espnet2/layers/utterance_mvn.py:45-88 (norm_means=True, norm_vars=False):
    zero the padding, mean over valid frames, subtract from ALL frames."""
    x = x.masked_fill(pad_mask(lens, x.shape[1])[:, :, None], 0.0)
    mean = x.sum(1, keepdim=True) / lens.to(x.dtype)[:, None, None]
    return x - mean


def global_mvn(x: torch.Tensor, lens: torch.Tensor, mean: torch.Tensor, std: torch.Tensor, norm_means: bool = True,
               norm_vars: bool = True) -> torch.Tensor:
    """espnet2/layers/global_mvn.py:66-94: subtract the global mean, zero the padding, divide by the global std."""
    if norm_means:
        x = x - mean
    x = x.masked_fill(pad_mask(lens, x.shape[1])[:, :, None], 0.0)
    if norm_vars:
        x = x / std
    return x


def variable_time_mask_range(T: int, ratio_range) -> Optional[Tuple[int, int]]:
    """MaskAlongAxisVariableMaxWidth.forward, espnet2/layers/mask_along_axis.py:187-203: the width range of the time
    masks follows the padded length; None when the range is empty (no mask is drawn then)."""
    lo = max(0, math.floor(T * ratio_range[0]))
    hi = min(T, math.floor(T * ratio_range[1]))
    return (lo, hi) if hi > lo else None


# --------------------------------------------------------------------------- #
# encoder
# --------------------------------------------------------------------------- #
@functools.lru_cache(maxsize=16)
def rel_pos_table(T: int, d: int, dtype=torch.float32) -> torch.Tensor:
    """RelPositionalEncoding (espnet/nets/pytorch_backend/transformer/embedding.py:196-244):
    rows are relative positions +(T-1) ... 0 ... -(T-1); sin on even, cos on odd."""
    pos = torch.arange(T - 1, -T, -1, dtype=torch.float32)[:, None]
    div = torch.exp(torch.arange(0, d, 2, dtype=torch.float32) * -(math.log(10000.0) / d))
    pe = torch.zeros(2 * T - 1, d)
    pe[:, 0::2] = torch.sin(pos * div)
    pe[:, 1::2] = torch.cos(pos * div)
    return pe.to(dtype)


@functools.lru_cache(maxsize=16)
def legacy_pos_table(T: int, d: int, maxlen: int = 5000, dtype=torch.float32) -> torch.Tensor:
    """LegacyRelPositionalEncoding (embedding.py:137-170, reverse=True at :153):
    position[r] = maxlen-1-r, rows 0..T-1 taken."""
    pos = torch.arange(maxlen - 1, maxlen - 1 - T, -1, dtype=torch.float32)[:, None]
    div = torch.exp(torch.arange(0, d, 2, dtype=torch.float32) * -(math.log(10000.0) / d))
    pe = torch.zeros(T, d)
    pe[:, 0::2] = torch.sin(pos * div)
    pe[:, 1::2] = torch.cos(pos * div)
    return pe.to(dtype)


@functools.lru_cache(maxsize=16)
def abs_pos_table(L: int, d: int, dtype=torch.float32) -> torch.Tensor:
    """PositionalEncoding (embedding.py:59-80)."""
    pos = torch.arange(0, L, dtype=torch.float32)[:, None]
    div = torch.exp(torch.arange(0, d, 2, dtype=torch.float32) * -(math.log(10000.0) / d))
    pe = torch.zeros(L, d)
    pe[:, 0::2] = torch.sin(pos * div)
    pe[:, 1::2] = torch.cos(pos * div)
    return pe.to(dtype)


def subsampled_len(T_in: int) -> int:
    """espnet/nets/pytorch_backend/transformer/subsampling.py:57-59,87."""
    return ((T_in - 1) // 2 - 1) // 2


def subsampled_lens(ilens: torch.Tensor, T_in: int) -> torch.Tensor:
    """olens = masks[:, :, :-2:2][:, :, :-2:2].sum() (subsampling.py:87,
    conformer_encoder.py:365): kept mask positions are 0,4,8,... so
    olens = min(T, ceil(ilens/4))."""
    T = subsampled_len(T_in)
    return torch.clamp((ilens + 3) // 4, max=T)


def conv2d_subsample(x: torch.Tensor, sd: SD, p: str) -> torch.Tensor:
    """Conv2dSubsampling.forward (subsampling.py:67-87) without the pos-enc:
    conv(1->d,3,s2)+ReLU, conv(d->d,3,s2)+ReLU, (b,c,t,f)->(b,t,c*f), Linear."""
    h = F.relu(F.conv2d(x[:, None], sd[p + ".conv.0.weight"], sd[p + ".conv.0.bias"], stride=2))
    h = F.relu(F.conv2d(h, sd[p + ".conv.2.weight"], sd[p + ".conv.2.bias"], stride=2))
    b, c, t, f = h.shape
    h = h.transpose(1, 2).reshape(b, t, c * f)
    return linear(h, sd, p + ".out.0")


def _d(drop, x, tag: str):
    """One torch.nn.Dropout call site of the reference.  ``drop`` is None (p = 0, the configuration of the frozen goldens)
    or a callable (x, tag) -> x * keep / (1 - p) that replays given keep decisions: torch's own Philox stream cannot be
    reproduced by another implementation, so dropout parity is checked with the decisions as an INPUT (the tests feed
    the CUDA path's exported decisions, b200slu_dropout_mask).  Tags name the module that owns the Dropout."""
    return x if drop is None else drop(x, tag)


def mha_core(q, k, v, key_valid, extra_scores=None, causal=False, drop=None, tag: str = ""):
    """forward_attention (attention.py:64-96): scores already scaled; masked with
    finfo.min, softmax, masked entries re-zeroed, dropout (:91), @ v, heads merged."""
    B, H, Tq, dk = q.shape
    Tk = k.shape[2]
    scores = q @ k.transpose(-2, -1)
    if extra_scores is not None:
        scores = scores + extra_scores
    scores = scores / math.sqrt(dk)
    valid = key_valid[:, None, None, :]  # (B,1,1,Tk)
    if causal:
        tri = torch.tril(torch.ones(Tq, Tk, dtype=torch.bool, device=q.device))
        valid = valid & tri[None, None]
    scores = scores.masked_fill(~valid, torch.finfo(scores.dtype).min)
    attn = torch.softmax(scores, dim=-1).masked_fill(~valid, 0.0)
    ctx = _d(drop, attn, tag + ".att") @ v
    return ctx.transpose(1, 2).reshape(B, Tq, H * dk)


@functools.lru_cache(maxsize=16)
def _shift_index_tensors(T: int, mode: str, device):
    """The integer index maps above as tensors on ``device`` (cached: they depend on T only)."""
    if mode == "latest":
        return (torch.from_numpy(rel_shift_index_latest(T)).to(device),)
    row, col = rel_shift_index_legacy(T)
    return (torch.from_numpy(row).to(device), torch.from_numpy(np.maximum(col, 0)).to(device),
            torch.from_numpy(col < 0).to(device))


def rel_pos_mhsa(x, pos_emb, key_valid, sd: SD, p: str, H: int, mode: str = "latest", drop=None):
    """RelPositionMultiHeadedAttention.forward (attention.py:265-308) /
    LegacyRelPositionMultiHeadedAttention.forward (attention.py:167-209).
    rel_shift is applied through the integer index maps above instead of the
    reference's pad/view trick."""
    B, T, d = x.shape
    dk = d // H
    q = linear(x, sd, p + ".linear_q").view(B, T, H, dk)
    k = linear(x, sd, p + ".linear_k").view(B, T, H, dk).transpose(1, 2)
    v = linear(x, sd, p + ".linear_v").view(B, T, H, dk).transpose(1, 2)
    pp = (pos_emb @ sd[p + ".linear_pos.weight"].t()).view(-1, H, dk).transpose(0, 1)  # (H, P, dk)
    qu = (q + sd[p + ".pos_bias_u"]).transpose(1, 2)
    qv = (q + sd[p + ".pos_bias_v"]).transpose(1, 2)
    bd = qv @ pp.transpose(-2, -1)[None]  # (B,H,T,P)
    if mode == "latest":
        col = _shift_index_tensors(T, mode, x.device)[0]
        bd = torch.gather(bd, 3, col[None, None].expand(B, H, T, T))
    else:
        row_t, col_t, zero = _shift_index_tensors(T, mode, x.device)
        bd = bd[:, :, row_t, col_t].masked_fill(zero[None, None], 0.0)
    ctx = mha_core(qu, k, v, key_valid, extra_scores=bd, drop=drop, tag=p)
    return linear(ctx, sd, p + ".linear_out")


def mha(xq, xkv, key_valid, sd: SD, p: str, H: int, causal: bool = False, drop=None):
    """MultiHeadedAttention.forward (attention.py:98-114)."""
    B, Tq, d = xq.shape
    Tk = xkv.shape[1]
    dk = d // H
    q = linear(xq, sd, p + ".linear_q").view(B, Tq, H, dk).transpose(1, 2)
    k = linear(xkv, sd, p + ".linear_k").view(B, Tk, H, dk).transpose(1, 2)
    v = linear(xkv, sd, p + ".linear_v").view(B, Tk, H, dk).transpose(1, 2)
    return linear(mha_core(q, k, v, key_valid, causal=causal, drop=drop, tag=p), sd, p + ".linear_out")


def ffn(x, sd: SD, p: str, act, drop=None):
    """PositionwiseFeedForward.forward (positionwise_feed_forward.py:30-32): w_2(dropout(act(w_1(x))))."""
    return linear(_d(drop, act(linear(x, sd, p + ".w_1")), p + ".inner"), sd, p + ".w_2")


def conv_module(x, sd: SD, p: str, kernel: int, training: bool = True, bn_stats_out: Optional[dict] = None):
    """ConvolutionModule.forward (espnet/nets/pytorch_backend/conformer/convolution.py:56-79):
    pw1 -> GLU(channel halves) -> depthwise(K, same pad) -> BatchNorm1d (training:
    batch statistics over ALL B*T positions incl. padding, biased var, eps 1e-5)
    -> Swish -> pw2.  Works channels-last; no transposes needed."""
    B, T, d = x.shape
    if LIBRARY_OPS:  # convolution.py:56-79 op for op (channels-first, as the reference lays it out)
        h = F.conv1d(x.transpose(1, 2), sd[p + ".pointwise_conv1.weight"], sd[p + ".pointwise_conv1.bias"])
        h = F.glu(h, dim=1)
        h = F.conv1d(h, sd[p + ".depthwise_conv.weight"], sd[p + ".depthwise_conv.bias"], padding=(kernel - 1) // 2, groups=d)
        if training:
            if bn_stats_out is not None:
                n = B * T
                mean = h.mean((0, 2))
                bn_stats_out[p] = (mean.detach(), (((h - mean[None, :, None]) ** 2).mean((0, 2)) * n / max(n - 1, 1)).detach())
            h = F.batch_norm(h, None, None, sd[p + ".norm.weight"], sd[p + ".norm.bias"], True, 0.1, 1e-5)
        else:
            h = F.batch_norm(h, sd[p + ".norm.running_mean"], sd[p + ".norm.running_var"], sd[p + ".norm.weight"],
                             sd[p + ".norm.bias"], False, 0.1, 1e-5)
        h = swish(h)
        return F.conv1d(h, sd[p + ".pointwise_conv2.weight"], sd[p + ".pointwise_conv2.bias"]).transpose(1, 2)
    h = x @ sd[p + ".pointwise_conv1.weight"][:, :, 0].t() + sd[p + ".pointwise_conv1.bias"]
    h = h[..., :d] * torch.sigmoid(h[..., d:])
    hp = F.pad(h, (0, 0, (kernel - 1) // 2, (kernel - 1) // 2))
    w = sd[p + ".depthwise_conv.weight"][:, 0, :]  # (d, K)
    y = sd[p + ".depthwise_conv.bias"].expand(B, T, d).clone()
    for k in range(kernel):
        y = y + hp[:, k:k + T, :] * w[:, k]
    if training:
        mean = y.mean((0, 1))
        var = ((y - mean) ** 2).mean((0, 1))
        if bn_stats_out is not None:
            n = B * T
            bn_stats_out[p] = (mean.detach(), (var * n / max(n - 1, 1)).detach())
    else:
        mean, var = sd[p + ".norm.running_mean"], sd[p + ".norm.running_var"]
    y = (y - mean) / torch.sqrt(var + 1e-5) * sd[p + ".norm.weight"] + sd[p + ".norm.bias"]
    y = swish(y)
    return y @ sd[p + ".pointwise_conv2.weight"][:, :, 0].t() + sd[p + ".pointwise_conv2.bias"]


def conformer_block(x, pos_emb, key_valid, sd: SD, p: str, cfg: EncCfg, bn_stats_out=None, training: bool = True, drop=None):
    """EncoderLayer.forward (espnet/nets/pytorch_backend/conformer/encoder_layer.py:76-157),
    pre-LN macaron; every branch is ``residual + scale * dropout(branch)`` (:101,129,138,144), dropout replayed through
    ``drop`` (see _d; None = p 0); training=False: BatchNorm uses its running statistics."""
    x = x + 0.5 * _d(drop, ffn(layer_norm(x, sd, p + ".norm_ff_macaron"), sd, p + ".feed_forward_macaron", swish, drop),
                     p + ".feed_forward_macaron.outer")
    x = x + _d(drop, rel_pos_mhsa(layer_norm(x, sd, p + ".norm_mha"), pos_emb, key_valid, sd, p + ".self_attn",
                                  cfg.heads, cfg.rel_pos_type, drop), p + ".self_attn.out")
    x = x + _d(drop, conv_module(layer_norm(x, sd, p + ".norm_conv"), sd, p + ".conv_module", cfg.kernel, training=training,
                                 bn_stats_out=bn_stats_out), p + ".conv_module.out")
    x = x + 0.5 * _d(drop, ffn(layer_norm(x, sd, p + ".norm_ff"), sd, p + ".feed_forward", swish, drop),
                     p + ".feed_forward.outer")
    return layer_norm(x, sd, p + ".norm_final")


def conformer_encoder(feats, ilens, sd: SD, cfg: EncCfg, p: str = "encoder", bn_stats_out=None,
                      return_blocks: bool = False, training: bool = True, cond_idx=(), drop=None):
    """ConformerEncoder.forward (espnet2/asr/encoder/conformer_encoder.py:292-368).  cond_idx: 1-based block indices
    after which the CTC posterior of the (after_norm'ed) block output is fed back, self-conditioning :350-358:
    x += conditioning_layer(softmax(ctc_lo(after_norm(x)))); the returned blocks are the outputs BEFORE that addition,
    which is what the intermediate CTC losses see (:339-347)."""
    T_in = feats.shape[1]
    x = _d(drop, conv2d_subsample(feats, sd, p + ".embed") * math.sqrt(cfg.d), p + ".embed.x")  # embedding.py:239,244
    T = x.shape[1]
    olens = subsampled_lens(ilens, T_in)
    key_valid = ~pad_mask(olens, T)
    if cfg.rel_pos_type == "latest":
        pos_emb = rel_pos_table(T, cfg.d, x.dtype).to(x.device)
    else:
        pos_emb = legacy_pos_table(T, cfg.d, dtype=x.dtype).to(x.device)
    pos_emb = _d(drop, pos_emb, p + ".embed.pos")  # embedding.py:244 / :170: dropout on the positional table too
    outs = []
    for n in range(cfg.blocks):
        x = conformer_block(x, pos_emb, key_valid, sd, f"{p}.encoders.{n}", cfg, bn_stats_out, training, drop)
        if return_blocks:
            outs.append(x)
        if n + 1 in cond_idx:
            post = torch.softmax(linear(layer_norm(x, sd, p + ".after_norm"), sd, "ctc.ctc_lo"), dim=-1)
            x = x + linear(post, sd, p + ".conditioning_layer")
    x = layer_norm(x, sd, p + ".after_norm")
    if return_blocks:
        return x, olens, outs
    return x, olens


# --------------------------------------------------------------------------- #
# CTC
# --------------------------------------------------------------------------- #
def ctc_extended_labels(labels: Sequence[int], blank: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """Blank-extended label lattice (espnet/nets/pytorch_backend/gtn_ctc.py:26-36):
    S = 2L+1 states, even = blank, odd = label[(s-1)//2]; skip arc s-2 -> s allowed
    iff s odd, s > 1 and label != previous label."""
    L = len(labels)
    ext = np.full(2 * L + 1, blank, dtype=np.int64)
    ext[1::2] = np.asarray(labels, dtype=np.int64)
    skip = np.zeros(2 * L + 1, dtype=np.uint8)
    for s in range(3, 2 * L + 1, 2):
        skip[s] = ext[s] != ext[s - 2]
    return ext, skip


def ctc_loss_numpy(logits: np.ndarray, hlens, labels_list, blank: int = 0):
    """fp64 log-space alpha/beta CTC over raw logits (T,B,V) with fused log_softmax.

    Restates what espnet2/asr/ctc.py:52-63,82-95 asks of torch.nn.CTCLoss(
    reduction='none', zero_infinity=True): per-utterance nll, and the gradient of
    sum_b nll_b wrt the LOGITS, = softmax - occupancy for t < T_b, 0 beyond and 0 for
    infeasible (inf) utterances.  Small cases only (python loops)."""
    T, B, V = logits.shape
    x = logits.astype(np.float64)
    m = x.max(-1, keepdims=True)
    lp = x - m - np.log(np.exp(x - m).sum(-1, keepdims=True))
    nll = np.zeros(B)
    grad = np.zeros_like(x)
    for b in range(B):
        Tb = int(hlens[b])
        ext, skip = ctc_extended_labels(labels_list[b], blank)
        S = len(ext)
        NEG = -np.inf
        a = np.full((Tb, S), NEG)
        a[0, 0] = lp[0, b, ext[0]]
        if S > 1:
            a[0, 1] = lp[0, b, ext[1]]
        for t in range(1, Tb):
            for s in range(S):
                c = [a[t - 1, s]]
                if s >= 1:
                    c.append(a[t - 1, s - 1])
                if skip[s]:
                    c.append(a[t - 1, s - 2])
                mx = max(c)
                if mx > NEG:
                    a[t, s] = mx + math.log(sum(math.exp(v - mx) for v in c)) + lp[t, b, ext[s]]
        fin = [a[Tb - 1, S - 1]] + ([a[Tb - 1, S - 2]] if S > 1 else [])
        mx = max(fin)
        if mx == NEG:
            nll[b] = 0.0  # zero_infinity
            continue
        ll = mx + math.log(sum(math.exp(v - mx) for v in fin))
        nll[b] = -ll
        bt = np.full((Tb, S), NEG)
        bt[Tb - 1, S - 1] = lp[Tb - 1, b, ext[S - 1]]
        if S > 1:
            bt[Tb - 1, S - 2] = lp[Tb - 1, b, ext[S - 2]]
        for t in range(Tb - 2, -1, -1):
            for s in range(S):
                c = [bt[t + 1, s]]
                if s + 1 < S:
                    c.append(bt[t + 1, s + 1])
                if s + 2 < S and skip[s + 2]:
                    c.append(bt[t + 1, s + 2])
                mx = max(c)
                if mx > NEG:
                    bt[t, s] = mx + math.log(sum(math.exp(v - mx) for v in c)) + lp[t, b, ext[s]]
        for t in range(Tb):
            occ = np.zeros(V)
            for s in range(S):
                ab = a[t, s] + bt[t, s]
                if ab > NEG:
                    occ[ext[s]] += math.exp(ab - ll - lp[t, b, ext[s]])
            grad[t, b] = np.exp(lp[t, b]) - occ
    return nll, grad


def ctc_branch(enc_out, olens, text, text_lens, sd: SD, p: str = "ctc"):
    """CTC.forward (espnet2/asr/ctc.py:72-97): ctc_lo, (T,B,V), log_softmax,
    CTCLoss(none, zero_infinity), sum / B.  torch's own ctc_loss IS the
    reference's arithmetic here (ctc.py:39-41)."""
    logits = linear(enc_out, sd, p + ".ctc_lo").transpose(0, 1)
    lp = logits.log_softmax(2)
    tgt = torch.cat([text[i, :l] for i, l in enumerate(text_lens.tolist())])
    loss = F.ctc_loss(lp, tgt, olens, text_lens, blank=0, reduction="none", zero_infinity=True)
    return loss.sum() / enc_out.shape[0]


# --------------------------------------------------------------------------- #
# decoder + label smoothing
# --------------------------------------------------------------------------- #
def add_sos_eos(text, text_lens, sos: int, eos: int, ignore: int = -1):
    """espnet/nets/pytorch_backend/transformer/add_sos_eos.py:12-31."""
    B = text.shape[0]
    L = int(text_lens.max()) + 1
    ys_in = torch.full((B, L), eos, dtype=torch.long, device=text.device)
    ys_out = torch.full((B, L), ignore, dtype=torch.long, device=text.device)
    for b in range(B):
        n = int(text_lens[b])
        ys_in[b, 0] = sos
        ys_in[b, 1:n + 1] = text[b, :n]
        ys_out[b, :n] = text[b, :n]
        ys_out[b, n] = eos
    return ys_in, ys_out


def transformer_decoder(memory, mem_lens, ys_in, ys_in_lens, sd: SD, cfg: DecCfg, p: str = "decoder", drop=None):
    """TransformerDecoder.forward (espnet2/asr/decoder/transformer_decoder.py:92-145) with
    DecoderLayer.forward (espnet/nets/pytorch_backend/transformer/decoder_layer.py:63-134)."""
    B, L = ys_in.shape
    x = _d(drop, sd[p + ".embed.0.weight"][ys_in] * math.sqrt(cfg.d) + abs_pos_table(L, cfg.d, memory.dtype).to(memory.device), p + ".embed")
    tgt_valid = ~pad_mask(ys_in_lens, L)
    mem_valid = ~pad_mask(mem_lens, memory.shape[1])
    for n in range(cfg.blocks):
        q = f"{p}.decoders.{n}"
        h = layer_norm(x, sd, q + ".norm1")
        x = x + _d(drop, mha(h, h, tgt_valid, sd, q + ".self_attn", cfg.heads, causal=True, drop=drop), q + ".self_attn.out")
        x = x + _d(drop, mha(layer_norm(x, sd, q + ".norm2"), memory, mem_valid, sd, q + ".src_attn", cfg.heads, drop=drop),
                   q + ".src_attn.out")
        x = x + _d(drop, ffn(layer_norm(x, sd, q + ".norm3"), sd, q + ".feed_forward", F.relu, drop), q + ".feed_forward.outer")
    x = layer_norm(x, sd, p + ".after_norm")
    return linear(x, sd, p + ".output_layer")


def decoder_forward_one_step(tgt, memory, sd: SD, cfg: DecCfg, cache=None, p: str = "decoder"):
    """BaseTransformerDecoder.forward_one_step (espnet2/asr/decoder/transformer_decoder.py:147-184) with the cache branch
    of DecoderLayer.forward (decoder_layer.py:84-134): only the LAST prefix position is a query; keys/values are recomputed
    over the whole prefix (layer input = previous layer's cache + its new row) and over the memory (memory_mask None);
    each layer's new row is appended to its cache.  Returns (log-probs of the next token (B, V), new cache list)."""
    B, L = tgt.shape
    x = sd[p + ".embed.0.weight"][tgt] * math.sqrt(cfg.d) + abs_pos_table(L, cfg.d, memory.dtype).to(memory.device)
    tgt_valid = torch.ones(B, L, dtype=torch.bool, device=tgt.device)
    mem_valid = torch.ones(B, memory.shape[1], dtype=torch.bool, device=tgt.device)
    new_cache = []
    for n in range(cfg.blocks):
        q = f"{p}.decoders.{n}"
        c = None if cache is None else cache[n]
        h = layer_norm(x, sd, q + ".norm1")
        if c is None:
            y = x + mha(h, h, tgt_valid, sd, q + ".self_attn", cfg.heads, causal=True)
        else:
            assert c.shape == (B, L - 1, cfg.d), (c.shape, (B, L - 1, cfg.d))
            # the subsequent mask's last row lets the newest position see the whole prefix: no causal mask needed
            y = x[:, -1:] + mha(h[:, -1:], h, tgt_valid, sd, q + ".self_attn", cfg.heads)
        y = y + mha(layer_norm(y, sd, q + ".norm2"), memory, mem_valid, sd, q + ".src_attn", cfg.heads)
        y = y + ffn(layer_norm(y, sd, q + ".norm3"), sd, q + ".feed_forward", F.relu)
        x = y if c is None else torch.cat([c, y], dim=1)
        new_cache.append(x)
    y = layer_norm(x[:, -1], sd, p + ".after_norm")
    return torch.log_softmax(linear(y, sd, p + ".output_layer"), dim=-1), new_cache


def label_smoothing_loss(logits, target, smoothing: float, ignore: int = -1):
    """LabelSmoothingLoss.forward (label_smoothing_loss.py:41-63), normalize_length=False:
    KL(true_dist || softmax) summed over non-ignored rows, / batch size."""
    B, L, V = logits.shape
    x = logits.reshape(-1, V)
    t = target.reshape(-1)
    ign = t == ignore
    tt = t.masked_fill(ign, 0)
    true = torch.full_like(x, smoothing / (V - 1))
    true.scatter_(1, tt[:, None], 1.0 - smoothing)
    kl = true * (torch.log(true) - torch.log_softmax(x, 1))
    return kl.masked_fill(ign[:, None], 0.0).sum() / B


def th_accuracy(logits, target, ignore: int = -1) -> float:
    """espnet/nets/pytorch_backend/nets_utils.py:299-320."""
    pred = logits.argmax(-1)
    m = target != ignore
    return float(((pred == target) & m).sum()) / float(m.sum())


# --------------------------------------------------------------------------- #
# whole step
# --------------------------------------------------------------------------- #
def asr_forward_from_feats(feats, feat_lens, text, text_lens, sd: SD, cfg: ModelCfg, drop=None):
    """ESPnetASRModel.forward from the encoder input on
    (espnet2/asr/espnet_model.py:202-297): returns (loss, stats)."""
    enc, olens, blocks = conformer_encoder(feats, feat_lens, sd, cfg.enc, return_blocks=True,
                                           cond_idx=cfg.interctc_layer_idx if cfg.interctc_use_conditioning else (),
                                           drop=drop)
    loss_ctc = ctc_branch(enc, olens, text, text_lens, sd)
    stats_ic = {"loss_ctc": loss_ctc.detach()}  # espnet_model.py:217: recorded before the intermediate losses are mixed in
    if cfg.interctc_weight != 0.0 and cfg.interctc_layer_idx:
        # intermediate CTC (conformer_encoder.py:336-347: the block output goes through after_norm; espnet_model.py:223-245:
        # the same CTC module scores it against the same targets, the losses are averaged and mixed into loss_ctc)
        loss_ic = 0.0
        for idx in cfg.interctc_layer_idx:
            l = ctc_branch(layer_norm(blocks[idx - 1], sd, "encoder.after_norm"), olens, text, text_lens, sd)
            stats_ic[f"loss_interctc_layer{idx}"] = l.detach()
            loss_ic = loss_ic + l
        loss_ic = loss_ic / len(cfg.interctc_layer_idx)
        loss_ctc = (1 - cfg.interctc_weight) * loss_ctc + cfg.interctc_weight * loss_ic
    sos = eos = cfg.vocab - 1  # espnet_model.py:75-82
    ys_in, ys_out = add_sos_eos(text, text_lens, sos, eos)
    dec = transformer_decoder(enc, olens, ys_in, text_lens + 1, sd, cfg.dec, drop=drop)
    loss_att = label_smoothing_loss(dec, ys_out, cfg.lsm_weight)
    acc = th_accuracy(dec, ys_out)
    loss = cfg.ctc_weight * loss_ctc + (1 - cfg.ctc_weight) * loss_att  # espnet_model.py:284
    return loss, {"loss_att": loss_att.detach(), "acc": acc, "loss": loss.detach(), **stats_ic}


def asr_forward(speech, speech_lens, text, text_lens, sd: SD, cfg: ModelCfg, melmat, draws=None):
    """ESPnetASRModel.forward from the raw waveform (espnet_model.py:169-297,
    encode 319-377): frontend -> SpecAug (draws given; None = eval/no-aug) ->
    UtteranceMVN -> encoder -> CTC + attention losses."""
    feats, flens = frontend_logmel(speech, speech_lens, melmat, cfg.n_fft, cfg.hop)
    if draws is not None:
        feats = spec_augment(feats, flens, draws)
    feats = utterance_mvn(feats, flens)
    return asr_forward_from_feats(feats, flens, text, text_lens, sd, cfg)
