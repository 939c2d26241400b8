"""Drop-in modules for the reference's plugin registry.

Each class keeps the reference class's constructor keywords, forward signature, state_dict names / shapes / default
initialisation and error behaviour, and runs on the B200 kernels of ``libb200slu.so``.  A maintainer registers them in
``espnet2/tasks/asr.py``; the recipe YAMLs, checkpoints and ``--init_param`` keep working.

    reference class (file:line)                                            replacement here
    ConformerEncoder   espnet2/asr/encoder/conformer_encoder.py:39-368     ConformerEncoder
    TransformerDecoder espnet2/asr/decoder/transformer_decoder.py:232-280  TransformerDecoder
    CTC                espnet2/asr/ctc.py:9-127                            CTC
    DefaultFrontend    espnet2/asr/frontend/default.py:18-131              DefaultFrontend
    SpecAug            espnet2/asr/specaug/specaug.py:12-96                SpecAug
    UtteranceMVN       espnet2/layers/utterance_mvn.py:9-88                UtteranceMVN
    GlobalMVN          espnet2/layers/global_mvn.py:13-94                  GlobalMVN
    ESPnetASRModel     espnet2/asr/espnet_model.py:39-297                  ESPnetASRModel (whole step fused: one
                                                                           forward+backward pass over the kernels)

The torch sub-modules created in the constructors (nn.Linear, nn.Conv2d, nn.LayerNorm, ...) are parameter containers
only: they give the reference's names, shapes, default init and BatchNorm buffers; their own ``forward`` is never
called.  Options the kernels do not implement raise ``NotImplementedError`` in the constructor -- there is no CPU or
eager fallback, and a module on a non-CUDA device raises at the first call.
"""
from __future__ import annotations

import math
import weakref
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

import torch
import torch.nn as nn

from . import melbank
from . import step as S
from .params import ParamPack


class TooShortUttError(Exception):
    """espnet/nets/pytorch_backend/transformer/subsampling.py:14-28 (same constructor and attributes)."""

    def __init__(self, message, actual_size, limit):
        super().__init__(message)
        self.actual_size = actual_size
        self.limit = limit


def check_short_utt(ins, size):
    """subsampling.py:31-46 for the layers this package implements (Conv2dSubsampling: kernel 3, stride 2, twice)."""
    if isinstance(ins, Conv2dSubsampling) and size < 7:
        return True, 7
    return False, -1


def _unsupported(cond: bool, what: str):
    if cond:
        raise NotImplementedError(f"espnet_slurp_b200: {what} is outside the B200 hot path (SURVEY.md 8a); "
                                  "use the reference module for this configuration")


def _need_cuda(t: torch.Tensor, who: str):
    if t.device.type != "cuda":
        raise RuntimeError(f"{who}: espnet_slurp_b200 modules run on a CUDA device (sm_100a) only; there is no CPU fallback")


class _Packed(nn.Module):
    """Shared plumbing: the module's parameters live in one flat fp32 buffer (+ bf16 copy + flat gradient buffer)."""

    _prefix = ""

    def _named(self) -> Dict[str, torch.Tensor]:
        return {self._prefix + n: p for n, p in self.named_parameters()}

    def _pack(self) -> ParamPack:
        pp: Optional[ParamPack] = self.__dict__.get("_pp")
        named = self._named()
        first = next(iter(named.values()))
        _need_cuda(first, type(self).__name__)
        stale = pp is None or pp.device != first.device
        if not stale:  # .to() / .cuda() / load of new Parameter objects re-creates storage: re-adopt
            for n, p in named.items():
                if p.data_ptr() != pp.P[n].data_ptr():
                    stale = True
                    break
        if stale:
            pp = ParamPack(named, device=first.device, adopt=True)
            self.__dict__["_pp"] = pp
        return pp

    def _infer_pack(self) -> ParamPack:
        """Pack for the no-grad scoring entry points (encode / forward_one_step / log_softmax ...): a sub-module that sits
        inside an ESPnetASRModel reads the MODEL's flat buffers (same names), so a decoding loop that alternates
        model.encode(), decoder.batch_score() and ctc.log_softmax() never re-packs the parameters."""
        owner = self.__dict__.get("_owner")
        owner = owner() if owner is not None else None
        if owner is not None and not any(sub is self for sub in (owner.encoder, owner.decoder, owner.ctc)):
            owner = None  # a copy of the sub-module that still carries the original model's link
        holder = owner if owner is not None else self
        pp = holder._pack()
        # the bf16 copies are refreshed only when a parameter changed since the last scoring call (a decoding loop calls
        # this once per hypothesis step: re-casting 109 M parameters each time would dominate it)
        ver = sum(int(p._version) for p in holder.parameters()) + 7919 * id(pp)
        if holder.__dict__.get("_infer_ver") != ver:
            pp.refresh()
            holder.__dict__["_infer_ver"] = ver
        return pp

    def __getstate__(self):  # pickling / deepcopy: the flat buffers and the owner link are rebuilt on first use
        return {k: v for k, v in self.__dict__.items() if k not in ("_pp", "_owner")}

    def _bn_buffers(self) -> Dict[str, torch.Tensor]:
        return {self._prefix + n: b for n, b in self.named_buffers()}

    def _ctx(self) -> S.StepCtx:
        # one dropout seed per call, drawn from torch's generator so torch.manual_seed() makes runs reproducible
        seed = int(torch.randint(0, 2 ** 31 - 1, (1,)).item()) if self.training else 0
        return S.StepCtx(seed=seed, training=self.training)


def _grads_out(pp: ParamPack, names: Sequence[str], scale) -> List[torch.Tensor]:
    """Parameter gradients in the reference layout, as fresh tensors (autograd may keep them as .grad)."""
    pp.finalize_grads()
    return [pp.Gref[n] * scale for n in names]


# ----------------------------------------------------------------------------------------------------------------------
# frontend / specaug / normalize  (no parameters, no gradient: the waveform carries none)
# ----------------------------------------------------------------------------------------------------------------------
class DefaultFrontend(nn.Module):
    """STFT (n_fft 512, hann, center, reflect pad) -> power -> Slaney mel -> log; default.py:82-131, stft.py:56-170,
    log_mel.py:57-84.  Same constructor keywords as default.py:25-44."""

    def __init__(self, fs: Union[int, str] = 16000, n_fft: int = 512, win_length: int = None, hop_length: int = 128,
                 window: Optional[str] = "hann", center: bool = True, normalized: bool = False, onesided: bool = True,
                 n_mels: int = 80, fmin: int = None, fmax: int = None, htk: bool = False, frontend_conf=None,
                 apply_stft: bool = True):
        super().__init__()
        if isinstance(fs, str):
            fs = int(float(fs.lower().replace("k", "")) * (1000 if "k" in fs.lower() else 1))
        _unsupported(win_length not in (None, n_fft), "win_length != n_fft")
        _unsupported(window != "hann" or not center or normalized or not onesided, "a non-default STFT window/centering")
        _unsupported(htk, "HTK mel scale")
        _unsupported(bool(frontend_conf), "WPE / beamforming frontend_conf")
        _unsupported(not apply_stft, "apply_stft=False")
        self.fs, self.n_fft, self.hop_length, self.n_mels = fs, n_fft, hop_length, n_mels
        fmin = 0.0 if fmin is None else fmin
        fmax = fs / 2 if fmax is None else fmax
        # the reference keeps the filterbank in its state_dict as "frontend.logmel.melmat" (log_mel.py:52), (n_freq, n_mels)
        self.logmel = nn.Module()
        self.logmel.register_buffer("melmat", melbank.mel_matrix(fs, n_fft, n_mels, float(fmin), float(fmax)).float())

    def output_size(self) -> int:
        return self.n_mels

    def forward(self, input: torch.Tensor, input_lengths: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        _need_cuda(input, "DefaultFrontend")
        if self.logmel.melmat.device != input.device:
            self.logmel.melmat = self.logmel.melmat.to(input.device)
        feats, flens = S.K.logmel(input.float().contiguous(), input_lengths.to(device=input.device, dtype=torch.int32),
                                  self.logmel.melmat, self.n_fft, self.hop_length)
        return feats, flens.to(torch.long)


def specaug_draws(flens: torch.Tensor, T_in: int, n_mels: int, W: Optional[int], F: Optional[Tuple[int, int]], n_f: int,
                  Tm: Optional[Tuple[int, int]], n_t: int, gen=None) -> Dict[str, Optional[torch.Tensor]]:
    """The random draws SpecAug makes, in the reference's order: time_warp.py:28-29 once per utterance, then
    mask_along_axis.py:33-43 for the frequency masks and for the time masks (one (B, n) draw each, shared max width)."""
    B = flens.numel()
    out: Dict[str, Optional[torch.Tensor]] = {"center": None, "warped": None, "fmask": None, "tmask": None}
    if W is not None:
        center = torch.full((B,), -1, dtype=torch.int32)
        warped = torch.zeros(B, dtype=torch.int32)
        fl = flens.tolist()
        for b in range(B):
            t = int(fl[b])
            if t - W > W:
                c = int(torch.randint(W, t - W, (1,), generator=gen))
                center[b] = c
                warped[b] = int(torch.randint(c - W, c + W, (1,), generator=gen)) + 1
        out["center"], out["warped"] = center, warped
    if F is not None:
        f_len = torch.randint(F[0], F[1], (B, n_f), generator=gen)
        f_pos = torch.randint(0, max(1, n_mels - int(f_len.max())), (B, n_f), generator=gen)
        out["fmask"] = torch.stack([f_pos, f_len], -1).int().contiguous()
    if Tm is not None:
        t_len = torch.randint(Tm[0], Tm[1], (B, n_t), generator=gen)
        t_pos = torch.randint(0, max(1, T_in - int(t_len.max())), (B, n_t), generator=gen)
        out["tmask"] = torch.stack([t_pos, t_len], -1).int().contiguous()
    return out


def _range(x) -> Optional[Tuple[int, int]]:
    if x is None:
        return None
    if isinstance(x, int):
        return (0, x)
    return (int(x[0]), int(x[1]))


class SpecAug(nn.Module):
    """Time warp (bicubic, window W) + frequency masks + time masks, specaug.py:89-96.  Same keywords as :27-45."""

    def __init__(self, apply_time_warp: bool = True, time_warp_window: int = 5, time_warp_mode: str = "bicubic",
                 apply_freq_mask: bool = True, freq_mask_width_range: Union[int, Sequence[int]] = (0, 20),
                 num_freq_mask: int = 2, apply_time_mask: bool = True,
                 time_mask_width_range: Optional[Union[int, Sequence[int]]] = None,
                 time_mask_width_ratio_range: Optional[Union[float, Sequence[float]]] = None, num_time_mask: int = 2):
        super().__init__()
        if not apply_time_warp and not apply_time_mask and not apply_freq_mask:
            raise ValueError("Either one of time_warp, time_mask, or freq_mask should be applied")
        if apply_time_mask and time_mask_width_range is not None and time_mask_width_ratio_range is not None:
            raise ValueError('Either one of "time_mask_width_range" or "time_mask_width_ratio_range" can be used')
        _unsupported(apply_time_warp and time_warp_mode != "bicubic", "time_warp_mode != bicubic")
        if apply_time_mask and time_mask_width_range is None and time_mask_width_ratio_range is None:
            raise ValueError('Either one of "time_mask_width_range" or "time_mask_width_ratio_range" should be used.')
        self.W = time_warp_window if apply_time_warp else None
        self.F = _range(freq_mask_width_range) if apply_freq_mask else None
        self.Tm = _range(time_mask_width_range) if apply_time_mask and time_mask_width_range is not None else None
        # MaskAlongAxisVariableMaxWidth (mask_along_axis.py:132-204): the width range follows the padded length
        self.Tratio = None
        if apply_time_mask and time_mask_width_range is None:
            r = time_mask_width_ratio_range
            r = (0.0, float(r)) if isinstance(r, float) else (float(r[0]), float(r[1]))
            if not r[1] > r[0]:
                raise ValueError(f"time_mask_width_ratio_range must be increasing: {r}")
            self.Tratio = r
        self.n_f, self.n_t = num_freq_mask, num_time_mask

    def forward(self, x: torch.Tensor, x_lengths: Optional[torch.Tensor] = None):
        _need_cuda(x, "SpecAug")
        B, T, Fd = x.shape
        lens = x_lengths if x_lengths is not None else torch.full((B,), T, dtype=torch.long)
        Tm = self.Tm
        if self.Tratio is not None:  # mask_along_axis.py:187-203: no mask at all when the range is empty
            lo, hi = max(0, math.floor(T * self.Tratio[0])), min(T, math.floor(T * self.Tratio[1]))
            Tm = (lo, hi) if hi > lo else None
        d = specaug_draws(lens.cpu(), T, Fd, self.W, self.F, self.n_f, Tm, self.n_t)
        dev = {k: (v.to(x.device) if v is not None else None) for k, v in d.items()}
        y = S.K.specaug_mvn(x.float().contiguous(), lens.to(device=x.device, dtype=torch.int32), dev["center"], dev["warped"],
                            dev["fmask"], dev["tmask"], do_mvn=False)
        return y, x_lengths


class UtteranceMVN(nn.Module):
    """Per-utterance mean normalisation over the valid frames, utterance_mvn.py:45-88 (norm_vars=False as in the recipes)."""

    def __init__(self, norm_means: bool = True, norm_vars: bool = False, eps: float = 1.0e-20):
        super().__init__()
        _unsupported(norm_vars or not norm_means, "UtteranceMVN with norm_vars=True or norm_means=False")
        self.norm_means, self.norm_vars, self.eps = norm_means, norm_vars, eps

    def extra_repr(self):
        return f"norm_means={self.norm_means}, norm_vars={self.norm_vars}"

    def forward(self, x: torch.Tensor, ilens: Optional[torch.Tensor] = None):
        _need_cuda(x, "UtteranceMVN")
        B, T, _ = x.shape
        lens = ilens if ilens is not None else torch.full((B,), T, dtype=torch.long)
        y = S.K.specaug_mvn(x.float().contiguous(), lens.to(device=x.device, dtype=torch.int32), do_mvn=True)
        return y, ilens


class GlobalMVN(nn.Module):
    """Global mean / variance normalisation from a statistics file, global_mvn.py:13-94 (asr.sh default feats_normalize).
    Same keywords; the Kaldi-style .npy and the {count, sum, sum_square} .npz statistics are both read (:41-53)."""

    def __init__(self, stats_file, norm_means: bool = True, norm_vars: bool = True, eps: float = 1.0e-20):
        super().__init__()
        import numpy as np

        self.norm_means, self.norm_vars, self.eps, self.stats_file = norm_means, norm_vars, eps, str(stats_file)
        stats = np.load(self.stats_file)
        if isinstance(stats, np.ndarray):
            count = stats[0].flatten()[-1]
            mean = stats[0, :-1] / count
            var = stats[1, :-1] / count - mean * mean
        else:
            count = stats["count"]
            mean = stats["sum"] / count
            var = stats["sum_square"] / count - mean * mean
        std = np.sqrt(np.maximum(var, eps))
        self.register_buffer("mean", torch.from_numpy(np.asarray(mean)))
        self.register_buffer("std", torch.from_numpy(np.asarray(std)))

    def extra_repr(self):
        return f"stats_file={self.stats_file}, norm_means={self.norm_means}, norm_vars={self.norm_vars}"

    def forward(self, x: torch.Tensor, ilens: Optional[torch.Tensor] = None):
        _need_cuda(x, "GlobalMVN")
        B, T, F = x.shape
        _unsupported(F % 4 != 0, "GlobalMVN with a feature dimension that is not a multiple of 4")
        lens = ilens if ilens is not None else torch.full((B,), T, dtype=torch.long)
        mean = self.mean.to(device=x.device, dtype=torch.float32).contiguous() if self.norm_means else None
        inv_std = (1.0 / self.std.to(device=x.device, dtype=torch.float32)).contiguous() if self.norm_vars else None
        y = S.K.global_mvn(x.float().contiguous(), lens.to(device=x.device, dtype=torch.int32), mean, inv_std)
        return y, ilens


# ----------------------------------------------------------------------------------------------------------------------
# parameter containers with the reference's attribute names (=> identical state_dict keys, shapes and default init)
# ----------------------------------------------------------------------------------------------------------------------
class Conv2dSubsampling(nn.Module):
    """subsampling.py:49-65: conv = [Conv2d(1,C,3,2), ReLU, Conv2d(C,C,3,2), ReLU]; out = [Linear(C*F2, C), pos_enc]."""

    def __init__(self, idim: int, odim: int):
        super().__init__()
        self.conv = nn.Sequential(nn.Conv2d(1, odim, 3, 2), nn.ReLU(), nn.Conv2d(odim, odim, 3, 2), nn.ReLU())
        self.out = nn.Sequential(nn.Linear(odim * (((idim - 1) // 2 - 1) // 2), odim), nn.Identity())


class _RelPosAttnParams(nn.Module):
    """attention.py:24-38 + 217-238 (linear_pos without bias, pos_bias_u / pos_bias_v xavier-uniform)."""

    def __init__(self, n_head: int, n_feat: int):
        super().__init__()
        self.linear_q = nn.Linear(n_feat, n_feat)
        self.linear_k = nn.Linear(n_feat, n_feat)
        self.linear_v = nn.Linear(n_feat, n_feat)
        self.linear_out = nn.Linear(n_feat, n_feat)
        self.linear_pos = nn.Linear(n_feat, n_feat, bias=False)
        self.pos_bias_u = nn.Parameter(torch.Tensor(n_head, n_feat // n_head))
        self.pos_bias_v = nn.Parameter(torch.Tensor(n_head, n_feat // n_head))
        nn.init.xavier_uniform_(self.pos_bias_u)
        nn.init.xavier_uniform_(self.pos_bias_v)


class _AttnParams(nn.Module):
    def __init__(self, n_feat: int):
        super().__init__()
        self.linear_q = nn.Linear(n_feat, n_feat)
        self.linear_k = nn.Linear(n_feat, n_feat)
        self.linear_v = nn.Linear(n_feat, n_feat)
        self.linear_out = nn.Linear(n_feat, n_feat)


class _FFNParams(nn.Module):
    def __init__(self, idim: int, hidden: int):
        super().__init__()
        self.w_1 = nn.Linear(idim, hidden)
        self.w_2 = nn.Linear(hidden, idim)


class _ConvModuleParams(nn.Module):
    """convolution.py:22-54."""

    def __init__(self, channels: int, kernel_size: int):
        super().__init__()
        assert (kernel_size - 1) % 2 == 0
        self.pointwise_conv1 = nn.Conv1d(channels, 2 * channels, 1)
        self.depthwise_conv = nn.Conv1d(channels, channels, kernel_size, padding=(kernel_size - 1) // 2, groups=channels)
        self.norm = nn.BatchNorm1d(channels)
        self.pointwise_conv2 = nn.Conv1d(channels, channels, 1)


class _LayerNorm(nn.LayerNorm):
    def __init__(self, n: int):
        super().__init__(n, eps=1e-12)  # layer_norm.py:23


class _EncoderLayerParams(nn.Module):
    """conformer/encoder_layer.py:45-74 attribute names."""

    def __init__(self, size: int, heads: int, ffn: int, kernel: int):
        super().__init__()
        self.self_attn = _RelPosAttnParams(heads, size)
        self.feed_forward = _FFNParams(size, ffn)
        self.feed_forward_macaron = _FFNParams(size, ffn)
        self.conv_module = _ConvModuleParams(size, kernel)
        self.norm_ff = _LayerNorm(size)
        self.norm_mha = _LayerNorm(size)
        self.norm_ff_macaron = _LayerNorm(size)
        self.norm_conv = _LayerNorm(size)
        self.norm_final = _LayerNorm(size)


class _DecoderLayerParams(nn.Module):
    """transformer/decoder_layer.py:40-61 attribute names."""

    def __init__(self, size: int, ffn: int):
        super().__init__()
        self.self_attn = _AttnParams(size)
        self.src_attn = _AttnParams(size)
        self.feed_forward = _FFNParams(size, ffn)
        self.norm1 = _LayerNorm(size)
        self.norm2 = _LayerNorm(size)
        self.norm3 = _LayerNorm(size)


# ----------------------------------------------------------------------------------------------------------------------
# autograd glue: the kernels' hand-written backward behind torch.autograd.Function
# ----------------------------------------------------------------------------------------------------------------------
def _i32(t: torch.Tensor, device) -> torch.Tensor:
    return t.to(device=device, dtype=torch.int32).contiguous()


class _EncoderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod: "ConformerEncoder", xs: torch.Tensor, ilens: torch.Tensor, *params):
        pp = mod._pack()
        pp.refresh()
        B = xs.shape[0]
        bn = mod._bn_buffers()
        enc, _, olens, saved = S.encoder_fwd(xs.float().contiguous(), _i32(ilens, xs.device), pp.P, pp.W, mod._cfg, mod._ctx(),
                                             p="encoder", bn_buffers=bn, want_inter_f32=True)
        ctx.mod, ctx.saved = mod, saved
        olens = olens.to(torch.long)
        ctx.mark_non_differentiable(olens)
        d = mod._cfg.d
        # intermediate-CTC taps (fp32, after_norm'ed) are further differentiable outputs, in block order
        return (enc.view(B, -1, d), olens, *[tap[4].view(B, -1, d) for tap in saved[11]])

    @staticmethod
    def backward(ctx, denc, _dolens, *dtaps):
        mod = ctx.mod
        pp = mod._pp
        pp.zero_grad()
        d = mod._cfg.d
        dinter = {tap[0]: g.float().contiguous().view(-1, d) for tap, g in zip(ctx.saved[11], dtaps) if g is not None}
        denc = denc.float().contiguous().view(-1, d) if denc is not None else torch.zeros_like(ctx.saved[6])
        S.encoder_bwd(denc, ctx.saved, pp.P, pp.W, pp.G, mod._cfg, p="encoder", dinter=dinter)
        ctx.saved = None
        return (None, None, None, *_grads_out(pp, mod._names, 1.0))


class _DecoderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod: "TransformerDecoder", hs: torch.Tensor, hlens, ys_in, ys_in_lens, *params):
        pp = mod._pack()
        pp.refresh()
        B, T, d = hs.shape
        mem16 = S.K.cast_bf16(hs.float().contiguous().view(B * T, d))
        logits, saved = S.decoder_fwd(mem16, _i32(hlens, hs.device), _i32(ys_in, hs.device), _i32(ys_in_lens, hs.device), pp.P,
                                      pp.W, mod._cfg, mod._ctx(), B, T, p="decoder")
        ctx.mod, ctx.saved, ctx.mem16, ctx.dims = mod, saved, mem16, (B, T, d, ys_in.shape[1])
        V = mod._cfg.vocab
        return logits[:, :V].reshape(B, ys_in.shape[1], V)

    @staticmethod
    def backward(ctx, dlogits):
        mod = ctx.mod
        pp = mod._pp
        pp.zero_grad()
        B, T, d, L = ctx.dims
        V = mod._cfg.vocab
        g16 = torch.zeros(B * L, S.K.round8(V), device=dlogits.device, dtype=torch.bfloat16)
        g16[:, :V] = dlogits.reshape(B * L, V)
        dmem = torch.zeros(B * T, d, device=dlogits.device, dtype=torch.float32)
        S.decoder_bwd(g16, dmem, ctx.saved, ctx.mem16, pp.P, pp.W, pp.G, mod._cfg, B, T, p="decoder")
        ctx.saved = ctx.mem16 = None
        return (None, dmem.view(B, T, d), None, None, None, *_grads_out(pp, mod._names, 1.0))


class _CTCFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod: "CTC", hs: torch.Tensor, hlens, ys_pad, ys_lens, *params):
        pp = mod._pack()
        pp.refresh()
        pp.zero_grad()
        B, T, d = hs.shape
        enc16 = S.K.cast_bf16(hs.float().contiguous().view(B * T, d))
        # the loss and its gradients (w.r.t. the encoder output and ctc_lo) come out of one pass over the lattice
        loss, denc = S.ctc_head(enc16, _i32(hlens, hs.device), _i32(ys_pad, hs.device), _i32(ys_lens, hs.device), pp.P, pp.W,
                                pp.G, mod._odim, B, T, 1.0, p="ctc", want_grad=True)
        # ctc_lo's gradients are copied out of the shared buffers now (1.2 MB at V = 600): the reference calls the same CTC
        # module once per intermediate-CTC tap before any backward (espnet_model.py:230-232)
        ctx.denc, ctx.dims, ctx.pgrads = denc, (B, T, d), _grads_out(pp, mod._names, 1.0)
        return loss.reshape(())

    @staticmethod
    def backward(ctx, g):
        B, T, d = ctx.dims
        return (None, (ctx.denc * g).view(B, T, d), None, None, None, *[p * g for p in ctx.pgrads])


class _ModelFn(torch.autograd.Function):
    """Whole train step (espnet_model.py:202-297 from the normalised features on): forward AND backward run inside
    forward() -- the backward kernels consume activations while they are hot and free them block by block --
    and autograd's backward() only scales the stored gradients by d(total)/d(loss)."""

    @staticmethod
    def forward(ctx, model: "ESPnetASRModel", want: bool, feats, flens, text, text_lens, *params):
        # want: gradients requested (decided by the caller: grad mode is always off inside Function.forward)
        pp = model._pack()
        pp.refresh()
        if want:
            pp.zero_grad()
        out = S.train_step_from_feats(feats, _i32(flens, feats.device), _i32(text, feats.device), _i32(text_lens, feats.device),
                                      pp.P, pp.W, pp.G, model._cfg, model._ctx(), bn_buffers=model._bn_buffers(),
                                      want_grad=want)
        ctx.model = model
        ctx.gen = model._gen = model.__dict__.get("_gen", 0) + 1
        ic = [out[f"loss_interctc_layer{i}"].reshape(()) for i in (model._cfg.enc.interctc_layer_idx if model._cfg.interctc_weight
                                                                   != 0.0 else ())]
        ret = (out["loss"].reshape(()), out["loss_ctc"].reshape(()), out["loss_att"].reshape(()), out["acc"].reshape(()), *ic)
        ctx.mark_non_differentiable(*ret[1:])
        return ret

    @staticmethod
    def backward(ctx, g, *_):
        model = ctx.model
        if ctx.gen != model._gen:
            raise RuntimeError("ESPnetASRModel: a later forward() overwrote this step's gradient buffers; call backward() "
                               "before the next forward()")
        return (None, None, None, None, None, None, *_grads_out(model._pp, model._names, g))


# ----------------------------------------------------------------------------------------------------------------------
# the plugin classes
# ----------------------------------------------------------------------------------------------------------------------
class ConformerEncoder(_Packed):
    """conformer_encoder.py:39-368.  Implemented configuration: conv2d input layer, rel_pos / rel_selfattn
    (latest and legacy), macaron FFN (linear, swish), conv module, pre-norm; the rest raises NotImplementedError."""

    _prefix = "encoder."

    def __init__(self, input_size: int, output_size: int = 256, attention_heads: int = 4, linear_units: int = 2048,
                 num_blocks: int = 6, dropout_rate: float = 0.1, positional_dropout_rate: float = 0.1,
                 attention_dropout_rate: float = 0.0, input_layer: str = "conv2d", normalize_before: bool = True,
                 concat_after: bool = False, positionwise_layer_type: str = "linear", positionwise_conv_kernel_size: int = 3,
                 macaron_style: bool = False, rel_pos_type: str = "legacy", pos_enc_layer_type: str = "rel_pos",
                 selfattention_layer_type: str = "rel_selfattn", activation_type: str = "swish", use_cnn_module: bool = True,
                 zero_triu: bool = False, cnn_module_kernel: int = 31, padding_idx: int = -1,
                 interctc_layer_idx: List[int] = [], interctc_use_conditioning: bool = False,
                 stochastic_depth_rate: Union[float, List[float]] = 0.0, layer_drop_rate: float = 0.0,
                 max_pos_emb_len: int = 5000):
        super().__init__()
        self._output_size = output_size
        if rel_pos_type == "legacy":  # conformer_encoder.py:116-120
            if pos_enc_layer_type == "rel_pos":
                pos_enc_layer_type = "legacy_rel_pos"
            if selfattention_layer_type == "rel_selfattn":
                selfattention_layer_type = "legacy_rel_selfattn"
        elif rel_pos_type != "latest":
            raise ValueError("unknown rel_pos_type: " + rel_pos_type)
        _unsupported(input_layer != "conv2d", f"input_layer={input_layer!r}")
        _unsupported(not normalize_before or concat_after, "post-norm / concat_after blocks")
        _unsupported(positionwise_layer_type != "linear", f"positionwise_layer_type={positionwise_layer_type!r}")
        _unsupported(not macaron_style or not use_cnn_module, "blocks without the macaron FFN or the conv module")
        _unsupported(pos_enc_layer_type not in ("rel_pos", "legacy_rel_pos") or
                     selfattention_layer_type not in ("rel_selfattn", "legacy_rel_selfattn"), "non-relative attention")
        if (pos_enc_layer_type == "rel_pos") != (selfattention_layer_type == "rel_selfattn"):
            raise ValueError("pos_enc_layer_type and selfattention_layer_type must both be legacy or both be latest")
        _unsupported(activation_type != "swish", f"activation_type={activation_type!r}")
        _unsupported(zero_triu, "zero_triu")
        if len(interctc_layer_idx) > 0:
            assert 0 < min(interctc_layer_idx) and max(interctc_layer_idx) < num_blocks  # conformer_encoder.py:284-285
        sd = stochastic_depth_rate
        _unsupported((any(sd) if isinstance(sd, (list, tuple)) else sd != 0.0) or layer_drop_rate != 0.0,
                     "stochastic depth / layer drop")
        _unsupported(output_size % attention_heads != 0 or output_size % 8 != 0 or (output_size // attention_heads) % 8 != 0,
                     "head sizes that are not multiples of 8")
        self.embed = Conv2dSubsampling(input_size, output_size)
        self.encoders = nn.ModuleList(_EncoderLayerParams(output_size, attention_heads, linear_units, cnn_module_kernel)
                                      for _ in range(num_blocks))
        self.after_norm = _LayerNorm(output_size)
        self.interctc_layer_idx = interctc_layer_idx
        self.interctc_use_conditioning = interctc_use_conditioning
        self.conditioning_layer = None
        self._cfg = S.EncCfg(input_size, output_size, attention_heads, linear_units, num_blocks, cnn_module_kernel,
                             "latest" if pos_enc_layer_type == "rel_pos" else "legacy", dropout_rate, positional_dropout_rate,
                             attention_dropout_rate, tuple(sorted(set(int(i) for i in interctc_layer_idx))),
                             bool(interctc_use_conditioning))
        self._names = [self._prefix + n for n, _ in self.named_parameters()]

    def output_size(self) -> int:
        return self._output_size

    def forward(self, xs_pad: torch.Tensor, ilens: torch.Tensor, prev_states: torch.Tensor = None, ctc=None
                ) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
        _need_cuda(xs_pad, "ConformerEncoder")
        # self-conditioning reads the CTC projection inside the encoder: available in the fused ESPnetASRModel / TrainStep
        # (one parameter pack), not through this stand-alone forward
        _unsupported(self.interctc_use_conditioning and len(self.interctc_layer_idx) > 0,
                     "a stand-alone ConformerEncoder.forward with interctc_use_conditioning (use ESPnetASRModel)")
        short, limit = check_short_utt(self.embed, xs_pad.size(1))
        if short:  # conformer_encoder.py:320-327
            raise TooShortUttError(f"has {xs_pad.size(1)} frames and is too short for subsampling "
                                   f"(it needs more than {limit} frames), return empty results", xs_pad.size(1), limit)
        xs, olens, *taps = _EncoderFn.apply(self, xs_pad, ilens, *self.parameters())
        if taps:  # conformer_encoder.py:366-367
            return (xs, list(zip(self._cfg.interctc_layer_idx, taps))), olens, None
        return xs, olens, None


class TransformerDecoder(_Packed):
    """transformer_decoder.py:232-280 / 92-145 (training forward).  Implemented: embed input layer, output layer,
    pre-norm blocks, sinusoidal PositionalEncoding."""

    _prefix = "decoder."

    def __init__(self, vocab_size: int, encoder_output_size: int, attention_heads: int = 4, linear_units: int = 2048,
                 num_blocks: int = 6, dropout_rate: float = 0.1, positional_dropout_rate: float = 0.1,
                 self_attention_dropout_rate: float = 0.0, src_attention_dropout_rate: float = 0.0, input_layer: str = "embed",
                 use_output_layer: bool = True, pos_enc_class=None, normalize_before: bool = True, concat_after: bool = False,
                 layer_drop_rate: float = 0.0):
        super().__init__()
        _unsupported(input_layer != "embed" or not use_output_layer, "a decoder without embed input / output layer")
        _unsupported(pos_enc_class is not None and getattr(pos_enc_class, "__name__", "") != "PositionalEncoding",
                     "a non-default pos_enc_class")
        _unsupported(not normalize_before or concat_after or layer_drop_rate != 0.0, "post-norm / concat_after / layer drop")
        d = encoder_output_size
        self.embed = nn.Sequential(nn.Embedding(vocab_size, d), nn.Identity())
        self.decoders = nn.ModuleList(_DecoderLayerParams(d, linear_units) for _ in range(num_blocks))
        self.after_norm = _LayerNorm(d)
        self.output_layer = nn.Linear(d, vocab_size)
        self._cfg = S.DecCfg(vocab_size, d, attention_heads, linear_units, num_blocks, dropout_rate, positional_dropout_rate,
                             self_attention_dropout_rate, src_attention_dropout_rate)
        self._names = [self._prefix + n for n, _ in self.named_parameters()]

    def forward(self, hs_pad: torch.Tensor, hlens: torch.Tensor, ys_in_pad: torch.Tensor, ys_in_lens: torch.Tensor
                ) -> Tuple[torch.Tensor, torch.Tensor]:
        _need_cuda(hs_pad, "TransformerDecoder")
        x = _DecoderFn.apply(self, hs_pad, hlens, ys_in_pad, ys_in_lens, *self.parameters())
        return x, ys_in_lens.to(hs_pad.device)

    # ---- incremental scoring (next row N2): BatchScorerInterface as asr_inference.py / batch_beam_search.py use it ----
    def forward_one_step(self, tgt: torch.Tensor, tgt_mask: torch.Tensor, memory: torch.Tensor,
                         cache: List[torch.Tensor] = None) -> Tuple[torch.Tensor, List[torch.Tensor]]:
        """transformer_decoder.py:147-184.  tgt int64 (batch, maxlen_out) prefix, memory (batch, maxlen_in, feat) without
        a mask; returns (log-probs of the next token (batch, vocab), per-layer outputs (batch, maxlen_out, size)).

        With ``cache`` (per layer, the outputs of the earlier positions, decoder_layer.py:84-98) only the LAST position is a
        query: one row per hypothesis goes through the projections, both attentions, the FFN and the output layer, the
        layer's output row is appended to its cache.  Keys / values of the prefix are projected from the cached layer
        inputs like the reference does (:100-110); the keys / values of the MEMORY -- maxlen_in rows per layer, which the
        reference projects again at every step (:112-123) and which dominate its cost -- are projected once per memory
        tensor and kept (`_mem_kv`).  Without a cache (first step, or a caller that passes none) the whole prefix runs
        causally through the training kernels with dropout off; both give the same rows (the invariant of
        test/test_transformer_decode.py)."""
        _need_cuda(memory, "TransformerDecoder")
        B, L = tgt.shape
        T, d = memory.shape[1], memory.shape[2]
        n_layers = len(self.decoders)
        if cache is not None:
            assert len(cache) == n_layers and all(c.shape == (B, L - 1, d) for c in cache), \
                ([tuple(c.shape) for c in cache], (B, L - 1, d))
        dev = memory.device
        if cache is not None and L > 1:
            return self._one_step_cached(tgt, memory, cache)
        with torch.no_grad():
            pp = self._infer_pack()
            mem16 = S.K.cast_bf16(memory.float().contiguous().view(B * T, d))
            mlens = torch.full((B,), T, device=dev, dtype=torch.int32)
            ylens = torch.full((B,), L, device=dev, dtype=torch.int32)
            logits, saved = S.decoder_fwd(mem16, mlens, _i32(tgt, dev).contiguous(), ylens, pp.P, pp.W, self._cfg,
                                          S.StepCtx(seed=0, training=False), B, T, p="decoder")
            layers, x_last = saved[2], saved[3]
            # a layer's output is the next layer's saved input
            outs = [layers[n + 1][0] for n in range(n_layers - 1)] + [x_last]
            new_cache = [o.view(B, L, d) for o in outs]
            V = self._cfg.vocab
            y = S.K.vocab_log_softmax(logits.view(B, L, -1)[:, -1].contiguous(), V)
        return y, new_cache

    def _attend(self, q, k, v, lens, B: int, Tq: int, Tk: int):
        """softmax(q k^T / sqrt(d_k)) v over the first lens[b] keys (attention.py:98-114), no dropout: the fused kernel
        when the keys fit it, else the batched contractions + softmax kernel."""
        K, cfg = S.K, self._cfg
        H, scale = cfg.heads, 1.0 / math.sqrt(cfg.d // cfg.heads)
        if K.relattn_supported(Tq, Tk, cfg.d // H, "none"):
            return K.relattn_fwd(q, None, k, v, None, lens, B, Tq, Tk, H, "none", False, scale)[2]
        ac = K.attn_f1(q, k, B, Tq, Tk, H)
        pr, _ = K.softmax_fwd(ac, None, lens, Tk, "none", False, scale)
        return K.attn_f2(pr, v, B, Tq, Tk, H)

    def _memory_kv(self, memory: torch.Tensor, pp):
        """Per layer, linear_k / linear_v of the encoder memory (bf16 [B*T, 2d]); computed once per memory tensor."""
        key = (memory.data_ptr(), tuple(memory.shape), tuple(memory.stride()), int(memory._version), id(pp),
               self.__dict__.get("_infer_ver"))
        memo = self.__dict__.get("_mem_kv")
        if memo is None or memo[0] != key:
            B, T, d = memory.shape
            mem16 = S.K.cast_bf16(memory.float().contiguous().view(B * T, d))
            kvs = []
            for n in range(len(self.decoders)):
                a = f"decoder.decoders.{n}.src_attn"
                kvs.append(S.K.gemm(mem16, pp.W[a + ".linear_kv.weight"], bias=pp.P[a + ".linear_kv.bias"]))
            memo = self.__dict__["_mem_kv"] = (key, kvs)
        return memo[1]

    def _one_step_cached(self, tgt: torch.Tensor, memory: torch.Tensor, cache: List[torch.Tensor]):
        K, cfg = S.K, self._cfg
        B, L = tgt.shape
        T, d = memory.shape[1], memory.shape[2]
        dev = memory.device
        with torch.no_grad():
            pp = self._infer_pack()
            P, W = pp.P, pp.W
            mkv = self._memory_kv(memory, pp)
            mlens = torch.full((B,), T, device=dev, dtype=torch.int32)
            ylens = torch.full((B,), L, device=dev, dtype=torch.int32)
            # the embedded prefix (transformer_decoder.py:166, embedding.py:82-94): layer 0's input; a gather, no cache needed
            x_emb = K.embed_fwd(P["decoder.embed.0.weight"], _i32(tgt, dev).contiguous().view(-1), S.pos_table("abs", L, d, dev), L,
                                math.sqrt(d))                                           # [B*L, d]
            x_new = x_emb.view(B, L, d)[:, -1].contiguous()                            # [B, d]
            new_cache = []
            for n in range(len(self.decoders)):
                p = f"decoder.decoders.{n}"
                # this layer's input over the whole prefix: the embedded prefix for layer 0, else the previous layer's cached
                # outputs + its new row (exactly what new_cache[n - 1] holds)
                x_full = x_emb if n == 0 else new_cache[n - 1].view(B * L, d)
                ln1, _, _ = K.layernorm_fwd(x_full, P[p + ".norm1.weight"], P[p + ".norm1.bias"])
                wq = W[p + ".self_attn.linear_qkv.weight"]
                bq = P[p + ".self_attn.linear_qkv.bias"]
                kv = K.gemm(ln1, wq[d:], bias=bq[d:])                                   # prefix keys / values [B*L, 2d]
                q = K.gemm(ln1.view(B, L, d)[:, -1], wq[:d], bias=bq[:d])              # the last position is the only query
                c1 = self._attend(q, kv[:, :d], kv[:, d:], ylens, B, 1, L)
                x1 = K.gemm(c1, W[p + ".self_attn.linear_out.weight"], bias=P[p + ".self_attn.linear_out.bias"],
                            residual=x_new, out_dtype=torch.float32)
                ln2, _, _ = K.layernorm_fwd(x1, P[p + ".norm2.weight"], P[p + ".norm2.bias"])
                q2 = K.gemm(ln2, W[p + ".src_attn.linear_q.weight"], bias=P[p + ".src_attn.linear_q.bias"])
                c2 = self._attend(q2, mkv[n][:, :d], mkv[n][:, d:], mlens, B, 1, T)
                x2 = K.gemm(c2, W[p + ".src_attn.linear_out.weight"], bias=P[p + ".src_attn.linear_out.bias"], residual=x1,
                            out_dtype=torch.float32)
                ln3, _, _ = K.layernorm_fwd(x2, P[p + ".norm3.weight"], P[p + ".norm3.bias"])
                hdn = K.gemm(ln3, W[p + ".feed_forward.w_1.weight"], bias=P[p + ".feed_forward.w_1.bias"], act="relu")
                x3 = K.gemm(hdn, W[p + ".feed_forward.w_2.weight"], bias=P[p + ".feed_forward.w_2.bias"], residual=x2,
                            out_dtype=torch.float32)
                new_cache.append(torch.cat([cache[n].to(torch.float32), x3.view(B, 1, d)], 1))
                x_new = x3
            lnf, _, _ = K.layernorm_fwd(x_new, P["decoder.after_norm.weight"], P["decoder.after_norm.bias"])
            V = cfg.vocab
            logits = torch.empty(B, K.round8(V), device=dev, dtype=torch.float32)
            K.gemm(lnf, W["decoder.output_layer.weight"], bias=P["decoder.output_layer.bias"], out=logits)
            y = K.vocab_log_softmax(logits, V)
        return y, new_cache

    def score(self, ys: torch.Tensor, state, x: torch.Tensor):
        """transformer_decoder.py:186-192: one hypothesis (ys (ylen,), x (xlen, feat))."""
        logp, state = self.forward_one_step(ys.unsqueeze(0), None, x.unsqueeze(0), cache=state)
        return logp.squeeze(0), state

    def batch_score(self, ys: torch.Tensor, states: List[Any], xs: torch.Tensor) -> Tuple[torch.Tensor, List[Any]]:
        """transformer_decoder.py:194-229: states[b][layer] (ylen-1, size) <-> batch layout [layer] (n_batch, ylen-1, size)."""
        n_batch, n_layers = len(ys), len(self.decoders)
        if states[0] is None:
            batch_state = None
        else:
            batch_state = [torch.stack([states[b][i] for b in range(n_batch)]) for i in range(n_layers)]
        logp, states = self.forward_one_step(ys, None, xs, cache=batch_state)
        return logp, [[states[i][b] for i in range(n_layers)] for b in range(n_batch)]


class CTC(_Packed):
    """ctc.py:9-127: ctc_lo projection + log_softmax + CTC loss (sum / batch, zero_infinity)."""

    _prefix = "ctc."

    def __init__(self, odim: int, encoder_output_size: int, dropout_rate: float = 0.0, ctc_type: str = "builtin",
                 reduce: bool = True, ignore_nan_grad: bool = None, zero_infinity: bool = True):
        super().__init__()
        _unsupported(ctc_type != "builtin", f"ctc_type={ctc_type!r}")
        _unsupported(not reduce, "reduce=False")
        if ignore_nan_grad is not None:
            zero_infinity = ignore_nan_grad
        _unsupported(not zero_infinity, "zero_infinity=False")
        _unsupported(dropout_rate != 0.0, "CTC dropout")
        self.dropout_rate, self.ctc_type, self.reduce = dropout_rate, ctc_type, reduce
        self.ctc_lo = nn.Linear(encoder_output_size, odim)
        self._odim = odim
        self._names = [self._prefix + n for n, _ in self.named_parameters()]

    def forward(self, hs_pad, hlens, ys_pad, ys_lens):
        _need_cuda(hs_pad, "CTC")
        return _CTCFn.apply(self, hs_pad, hlens, ys_pad, ys_lens, *self.parameters())

    def _logits(self, hs_pad):
        pp = self._infer_pack()
        B, T, d = hs_pad.shape
        V = self._odim
        out = torch.empty(B * T, S.K.round8(V), device=hs_pad.device, dtype=torch.float32)
        S.K.gemm(S.K.cast_bf16(hs_pad.float().contiguous().view(B * T, d)), pp.W["ctc.ctc_lo.weight"],
                 bias=pp.P["ctc.ctc_lo.bias"], out=out)
        return out

    def softmax(self, hs_pad):
        return self.log_softmax(hs_pad).exp()

    def log_softmax(self, hs_pad):
        """ctc.py:109-117 on the fp32 log-softmax kernel (no torch op on the scoring path)."""
        B, T, _ = hs_pad.shape
        with torch.no_grad():
            return S.K.vocab_log_softmax(self._logits(hs_pad), self._odim).view(B, T, self._odim)

    def argmax(self, hs_pad):
        B, T, _ = hs_pad.shape
        with torch.no_grad():
            return torch.argmax(self._logits(hs_pad)[:, :self._odim], dim=1).view(B, T)


class ESPnetASRModel(_Packed):
    """espnet_model.py:39-297, CTC/attention hybrid branch, with the whole forward+backward fused (``_ModelFn``)."""

    _prefix = ""

    def __init__(self, vocab_size: int, token_list: Union[Tuple[str, ...], List[str]], frontend, specaug, normalize,
                 preencoder, encoder: ConformerEncoder, postencoder, decoder: TransformerDecoder, ctc: CTC, joint_network=None,
                 ctc_weight: float = 0.5, interctc_weight: float = 0.0, ignore_id: int = -1, lsm_weight: float = 0.0,
                 length_normalized_loss: bool = False, report_cer: bool = True, report_wer: bool = True,
                 sym_space: str = "<space>", sym_blank: str = "<blank>", sym_sos: str = "<sos/eos>", sym_eos: str = "<sos/eos>",
                 extract_feats_in_collect_stats: bool = True, lang_token_id: int = -1):
        assert 0.0 <= ctc_weight <= 1.0, ctc_weight
        assert 0.0 <= interctc_weight < 1.0, interctc_weight
        super().__init__()
        _unsupported(preencoder is not None or postencoder is not None or joint_network is not None,
                     "pre/post-encoders and transducer joint networks")
        _unsupported(ctc_weight in (0.0, 1.0), "pure attention (ctc_weight 0) or pure CTC (ctc_weight 1) training")
        _unsupported(length_normalized_loss or lang_token_id != -1, "length-normalised loss / language tokens")
        _unsupported(not isinstance(encoder, ConformerEncoder) or not isinstance(decoder, TransformerDecoder) or
                     not isinstance(ctc, CTC), "mixing reference encoder/decoder/ctc modules into the fused model")
        token_list = list(token_list)
        self.blank_id = token_list.index(sym_blank) if sym_blank in token_list else 0
        _unsupported(self.blank_id != 0, "a blank symbol that is not token 0")
        self.sos = token_list.index(sym_sos) if sym_sos in token_list else vocab_size - 1
        self.eos = token_list.index(sym_eos) if sym_eos in token_list else vocab_size - 1
        _unsupported(self.sos != vocab_size - 1 or self.eos != vocab_size - 1, "sos/eos other than the last token")
        self.vocab_size, self.ignore_id, self.ctc_weight = vocab_size, ignore_id, ctc_weight
        self.interctc_weight = interctc_weight
        self.token_list = token_list
        self.frontend, self.specaug, self.normalize = frontend, specaug, normalize
        self.preencoder = self.postencoder = None
        self.encoder, self.decoder, self.ctc = encoder, decoder, ctc
        if not hasattr(self.encoder, "interctc_use_conditioning"):
            self.encoder.interctc_use_conditioning = False
        if self.encoder.interctc_use_conditioning:  # espnet_model.py:96-101
            _unsupported(vocab_size % 8 != 0, "self-conditioned intermediate CTC with a vocabulary size that is not a multiple "
                                              "of 8 (the posterior is a GEMM operand)")
            self.encoder.conditioning_layer = nn.Linear(vocab_size, self.encoder.output_size())
            self.encoder._cfg.vocab = vocab_size
        self.extract_feats_in_collect_stats = extract_feats_in_collect_stats
        # intermediate CTC (espnet_model.py:223-245) needs both the weight and taps in the encoder; either alone is a no-op
        self._cfg = S.ModelCfg(enc=encoder._cfg, dec=decoder._cfg, vocab=vocab_size, ctc_weight=ctc_weight, lsm_weight=lsm_weight,
                               interctc_weight=interctc_weight if encoder._cfg.interctc_layer_idx else 0.0)
        self._names = [n for n, _ in self.named_parameters()]
        self._link_owner()

    def _link_owner(self):
        for sub in (self.encoder, self.decoder, self.ctc):  # weak: no module cycle
            sub.__dict__["_owner"] = weakref.ref(self)

    # espnet_model.py:319-377 (the entry asr_inference.py uses): frontend -> [SpecAug when training] -> normalize -> encoder
    def encode(self, speech: torch.Tensor, speech_lengths: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        _need_cuda(speech, "ESPnetASRModel")
        self._link_owner()  # (again after a deepcopy / unpickle)
        feats, flens = self._extract_feats(speech, speech_lengths)
        if self.specaug is not None and self.training:
            feats, flens = self.specaug(feats, flens)
        if self.normalize is not None:
            feats, flens = self.normalize(feats, flens)
        short, limit = check_short_utt(self.encoder.embed, feats.size(1))
        if short:
            raise TooShortUttError(f"has {feats.size(1)} frames and is too short for subsampling "
                                   f"(it needs more than {limit} frames), return empty results", feats.size(1), limit)
        B = feats.shape[0]
        with torch.no_grad():
            pp = self._infer_pack()
            enc, _, olens, _ = S.encoder_fwd(feats.float().contiguous(), _i32(flens, feats.device), pp.P, pp.W, self._cfg.enc,
                                             self._ctx(), p="encoder", bn_buffers=self._bn_buffers())
        encoder_out, encoder_out_lens = enc.view(B, -1, self._cfg.enc.d), olens.to(torch.long)
        assert encoder_out.size(0) == speech.size(0), (encoder_out.size(), speech.size(0))
        return encoder_out, encoder_out_lens

    # espnet_model.py:379-396
    def _extract_feats(self, speech, speech_lengths):
        assert speech_lengths.dim() == 1, speech_lengths.shape
        speech = speech[:, : speech_lengths.max()]
        if self.frontend is not None:
            return self.frontend(speech, speech_lengths)
        return speech, speech_lengths

    def collect_feats(self, speech, speech_lengths, text, text_lengths, **kwargs) -> Dict[str, torch.Tensor]:
        feats, feats_lengths = self._extract_feats(speech, speech_lengths)
        return {"feats": feats, "feats_lengths": feats_lengths}

    def forward(self, speech: torch.Tensor, speech_lengths: torch.Tensor, text: torch.Tensor, text_lengths: torch.Tensor,
                **kwargs) -> Tuple[torch.Tensor, Dict[str, torch.Tensor], torch.Tensor]:
        assert text_lengths.dim() == 1, text_lengths.shape
        assert speech.shape[0] == speech_lengths.shape[0] == text.shape[0] == text_lengths.shape[0], \
            (speech.shape, speech_lengths.shape, text.shape, text_lengths.shape)  # espnet_model.py:186-193
        _need_cuda(speech, "ESPnetASRModel")
        batch_size = speech.shape[0]
        text = text[:, : text_lengths.max()]
        feats, flens = self._extract_feats(speech, speech_lengths)
        if self.specaug is not None and self.training:
            feats, flens = self.specaug(feats, flens)
        if self.normalize is not None:
            feats, flens = self.normalize(feats, flens)
        short, limit = check_short_utt(self.encoder.embed, feats.size(1))
        if short:
            raise TooShortUttError(f"has {feats.size(1)} frames and is too short for subsampling "
                                   f"(it needs more than {limit} frames), return empty results", feats.size(1), limit)
        want = torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters())
        loss, loss_ctc, loss_att, acc, *ic = _ModelFn.apply(self, want, feats.float().contiguous(), flens, text, text_lengths,
                                                            *self.parameters())
        stats = dict(loss_ctc=loss_ctc.detach()[None], cer_ctc=None)
        for idx, l in zip(self._cfg.enc.interctc_layer_idx, ic):  # espnet_model.py:236-239
            stats[f"loss_interctc_layer{idx}"] = l.detach()[None]
            stats[f"cer_interctc_layer{idx}"] = None
        stats.update(loss_att=loss_att.detach()[None], acc=acc.detach()[None], cer=None, wer=None, loss=loss.detach()[None])
        weight = torch.tensor([batch_size], dtype=torch.long, device=loss.device)
        return loss[None], stats, weight
