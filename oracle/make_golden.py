"""Generate tests/golden/*.npz from the UNMODIFIED reference (build container only).

TEST INFRASTRUCTURE (see oracle/__init__.py).  Run:  python -m oracle.make_golden
The reference cannot travel to the GPU box, so its outputs on small seeded inputs
are frozen here; tests/test_oracle_golden.py replays them against oracle/model.py.
Fixture shapes follow the reference's own tests.
"""
from __future__ import annotations

import math
import os
import warnings

import numpy as np
import torch

from oracle import ref_shim

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

TINY_ENC = dict(output_size=16, attention_heads=2, linear_units=32, num_blocks=2, dropout_rate=0.0,
                positional_dropout_rate=0.0, attention_dropout_rate=0.0, input_layer="conv2d",
                normalize_before=True, macaron_style=True, rel_pos_type="latest", pos_enc_layer_type="rel_pos",
                selfattention_layer_type="rel_selfattn", activation_type="swish", use_cnn_module=True,
                cnn_module_kernel=7)
TINY_DEC = dict(attention_heads=2, linear_units=32, num_blocks=2, dropout_rate=0.0, positional_dropout_rate=0.0,
                self_attention_dropout_rate=0.0, src_attention_dropout_rate=0.0)
TINY_V = 20


def _np(d):
    return {k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in d.items()}


def randomize_(model, seed):
    """Give every parameter (incl. zero-init biases, BN affine) a non-trivial value."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for n, p in model.named_parameters():
            if p.dim() == 1:
                p.add_(0.1 * torch.randn(p.shape, generator=g))


def tiny_batch(seed=0, B=3, N=6000, V=TINY_V):
    g = torch.Generator().manual_seed(seed)
    lens = torch.tensor([N, N - 800, N - 1900])[:B]
    wav = torch.randn(B, N, generator=g) * 0.1
    for b in range(B):
        wav[b, lens[b]:] = 0
    tl = torch.tensor([4, 2, 3])[:B]
    text = torch.full((B, int(tl.max())), -1, dtype=torch.long)
    for b in range(B):
        text[b, :tl[b]] = torch.randint(2, V - 1, (int(tl[b]),), generator=g)
    return wav, lens, text, tl


def gen_e2e(rel_pos_type: str, name: str, interctc=None, cond=False, V=TINY_V):
    """Whole train step from raw waveform: loss, stats, encoder output, all grads.  interctc = (weight, [block idx])
    adds intermediate CTC losses (next-row N4; 3 blocks so that two of them can carry one)."""
    torch.manual_seed(1)
    enc_conf = dict(TINY_ENC, rel_pos_type=rel_pos_type)
    if rel_pos_type == "legacy":
        enc_conf.update(pos_enc_layer_type="rel_pos", selfattention_layer_type="rel_selfattn")
    if interctc is not None:
        enc_conf.update(num_blocks=3, interctc_layer_idx=list(interctc[1]), interctc_use_conditioning=cond)
    m = ref_shim.build_reference_model(V, enc_conf, TINY_DEC, interctc_weight=interctc[0] if interctc else 0.0)
    randomize_(m, 2)
    m.train()
    wav, lens, text, tl = tiny_batch(V=V)
    loss, stats, _ = m(wav, lens, text.clone(), tl)
    loss.backward()
    with torch.no_grad():
        feats, flens = m._extract_feats(wav, lens)
        feats_n, _ = m.normalize(feats.clone(), flens)
        # note: BN running stats updated twice
        enc_out, enc_lens, _ = m.encoder(feats_n, flens, ctc=m.ctc) if cond else m.encoder(feats_n, flens)
        if isinstance(enc_out, tuple):
            enc_out = enc_out[0]
    out = {"wav": wav, "lens": lens, "text": text, "text_lens": tl, "loss": loss.detach(),
           "loss_ctc": stats["loss_ctc"], "loss_att": stats["loss_att"], "acc": stats["acc"],
           "feats": feats, "feat_lens": flens, "enc_out": enc_out, "enc_lens": enc_lens}
    for k, v in stats.items():
        if k.startswith("loss_interctc"):
            out[k] = v
    for k, v in m.state_dict().items():
        if "running_" in k or "num_batches" in k:
            continue
        out["sd/" + k] = v
    for k, p in m.named_parameters():
        out["grad/" + k] = p.grad if p.grad is not None else torch.zeros_like(p)
    np.savez_compressed(os.path.join(OUT, name), **_np(out))
    print(name, float(loss))


def gen_bn_running():
    """BatchNorm running-stat update after ONE training forward (momentum 0.1, unbiased var)."""
    torch.manual_seed(3)
    m = ref_shim.build_reference_model(TINY_V, TINY_ENC, TINY_DEC, frontend=False)
    randomize_(m, 4)
    m.train()
    g = torch.Generator().manual_seed(5)
    feats = torch.randn(2, 40, 80, generator=g)
    flens = torch.tensor([40, 33])
    sd0 = {k: v.clone() for k, v in m.state_dict().items()}
    with torch.no_grad():
        enc_out, enc_lens, _ = m.encoder(feats, flens)
    out = {"feats": feats, "feat_lens": flens, "enc_out": enc_out, "enc_lens": enc_lens}
    for k, v in sd0.items():
        if k.startswith("encoder."):
            out["sd/" + k] = v
    for k, v in m.state_dict().items():
        if "running_" in k or "num_batches" in k:
            out["after/" + k] = v
    np.savez_compressed(os.path.join(OUT, "encoder_bn.npz"), **_np(out))


def gen_rel_shift():
    """Index maps of both rel_shift variants, traced through the reference functions
    with a matrix whose entries encode their own (row, col)."""
    ref_shim.install()
    from espnet.nets.pytorch_backend.transformer.attention import (
        LegacyRelPositionMultiHeadedAttention, RelPositionMultiHeadedAttention)

    new = RelPositionMultiHeadedAttention(1, 4, 0.0)
    old = LegacyRelPositionMultiHeadedAttention(1, 4, 0.0)
    out = {}
    for T in (1, 2, 3, 7, 16, 33, 64):
        P = 2 * T - 1
        code = (torch.arange(T)[:, None] * 10000 + torch.arange(P)[None, :] + 1).double()[None, None]
        out[f"latest_{T}"] = new.rel_shift(code)[0, 0].long()
        code = (torch.arange(T)[:, None] * 10000 + torch.arange(T)[None, :] + 1).double()[None, None]
        out[f"legacy_{T}"] = old.rel_shift(code)[0, 0].long()
    np.savez_compressed(os.path.join(OUT, "rel_shift.npz"), **_np(out))


def gen_ctc():
    """torch.nn.CTCLoss(reduction='none', zero_infinity=True) -- the arithmetic the
    reference delegates to (espnet2/asr/ctc.py:39-41,54-55) -- in fp64 on the
    reference tests' length sets (test/test_loss.py:12-48, test/espnet2/asr/test_ctc.py:7-14)
    plus repeats and an infeasible utterance."""
    cases = {
        "loss_a": ([11, 17, 15], [4, 2, 3], 7),
        "loss_b": ([4], [1], 7),
        "ctc_fixture": ([10, 8], [5, 2], 10),
        "infeasible": ([6, 3, 9], [2, 3, 4], 6),  # utt 1: T=3 < 2*... with repeats -> inf
    }
    out = {}
    g = torch.Generator().manual_seed(11)
    for name, (hl, yl, V) in cases.items():
        T, B = max(hl), len(hl)
        logits = torch.randn(T, B, V, generator=g, dtype=torch.float64, requires_grad=True)
        ys = [torch.randint(1, V, (n,), generator=g) for n in yl]
        if name == "infeasible":
            ys[1] = torch.tensor([2, 2, 2])
            ys[2] = torch.tensor([3, 3, 1, 1])
        lp = logits.log_softmax(2)
        loss = torch.nn.functional.ctc_loss(lp, torch.cat(ys), torch.tensor(hl), torch.tensor(yl),
                                            blank=0, reduction="none", zero_infinity=True)
        loss.sum().backward()
        out[name + "/logits"] = logits.detach()
        out[name + "/hlens"] = torch.tensor(hl)
        out[name + "/ylens"] = torch.tensor(yl)
        out[name + "/ys"] = torch.cat(ys)
        out[name + "/nll"] = loss.detach()
        out[name + "/grad"] = logits.grad
    # the espnet2 CTC module end to end (ctc_lo + loss/B), fixture of test_ctc.py:7-14
    ref_shim.install()
    from espnet2.asr.ctc import CTC

    torch.manual_seed(12)
    ctc = CTC(odim=10, encoder_output_size=10)
    h = torch.randn(2, 10, 10, requires_grad=True)
    y = torch.randint(1, 5, (2, 5))
    loss = ctc(h, torch.tensor([10, 8]), y, torch.tensor([5, 2]))
    loss.backward()
    out.update({"module/h": h.detach(), "module/y": y, "module/w": ctc.ctc_lo.weight.detach(),
                "module/b": ctc.ctc_lo.bias.detach(), "module/loss": loss.detach(), "module/grad_h": h.grad,
                "module/grad_w": ctc.ctc_lo.weight.grad})
    np.savez_compressed(os.path.join(OUT, "ctc.npz"), **_np(out))


def gen_specaug():
    """SpecAug with the reference's RNG stream replayed: the draws are re-derived here by
    issuing the same torch.randint calls in the same order (time_warp.py:28-29 per
    utterance, then mask_along_axis.py:33-43 for freq, then time)."""
    ref_shim.install()
    from espnet2.asr.specaug.specaug import SpecAug

    conf = dict(apply_time_warp=True, time_warp_window=5, time_warp_mode="bicubic", apply_freq_mask=True,
                freq_mask_width_range=(0, 30), num_freq_mask=2, apply_time_mask=True,
                time_mask_width_range=(0, 40), num_time_mask=2)
    sa = SpecAug(**conf)
    g = torch.Generator().manual_seed(21)
    B, T, Fq = 3, 60, 80
    x = torch.randn(B, T, Fq, generator=g)
    lens = torch.tensor([60, 47, 31])
    for b in range(B):
        x[b, lens[b]:] = 0
    torch.manual_seed(22)
    y, _ = sa(x.clone(), lens)
    # replay
    torch.manual_seed(22)
    W = 5
    center, warped = [], []
    for b in range(B):
        t = int(lens[b])
        if t - W <= W:
            center.append(-1), warped.append(-1)
            continue
        c = torch.randint(W, t - W, (1,))[0]
        w = torch.randint(c - W, c + W, (1,))[0] + 1
        center.append(int(c)), warped.append(int(w))
    f_len = torch.randint(0, 30, (B, 2))
    f_pos = torch.randint(0, max(1, Fq - int(f_len.max())), (B, 2))
    t_len = torch.randint(0, 40, (B, 2))
    t_pos = torch.randint(0, max(1, T - int(t_len.max())), (B, 2))
    out = {"x": x, "lens": lens, "y": y, "center": np.array(center), "warped": np.array(warped),
           "f_len": f_len, "f_pos": f_pos, "t_len": t_len, "t_pos": t_pos}
    np.savez_compressed(os.path.join(OUT, "specaug.npz"), **_np(out))


def gen_frontend():
    """DefaultFrontend + UtteranceMVN on the length set probed in SURVEY.md Appendix A.2
    (scaled down) and the STFT shape fixture of test/espnet2/layers/test_stft.py:10-16."""
    ref_shim.install()
    from espnet2.asr.frontend.default import DefaultFrontend
    from espnet2.layers.stft import Stft
    from espnet2.layers.utterance_mvn import UtteranceMVN

    g = torch.Generator().manual_seed(31)
    lens = torch.tensor([6400, 6000, 5000, 4800])
    wav = torch.randn(4, 6400, generator=g) * 0.1
    for b in range(4):
        wav[b, lens[b]:] = 0
    fe = DefaultFrontend(frontend_conf=None)
    feats, flens = fe(wav, lens)
    mvn, _ = UtteranceMVN()(feats.clone(), flens)
    st = Stft(n_fft=4, win_length=4, hop_length=2)
    s_out, s_lens = st(torch.randn(2, 30, generator=g), torch.tensor([30, 15]))
    out = {"wav": wav, "lens": lens, "feats": feats, "feat_lens": flens, "mvn": mvn,
           "melmat": fe.logmel.melmat, "stft_shape": np.array(s_out.shape), "stft_lens": s_lens}
    np.savez_compressed(os.path.join(OUT, "frontend.npz"), **_np(out))


def gen_norm_variants():
    """N4 rows: GlobalMVN (global_mvn.py, both statistics file formats) and SpecAug with time_mask_width_ratio_range
    (MaskAlongAxisVariableMaxWidth, the LibriSpeech conformer YAMLs) with the reference's RNG stream replayed."""
    ref_shim.install()
    import tempfile
    from espnet2.asr.specaug.specaug import SpecAug
    from espnet2.layers.global_mvn import GlobalMVN

    g = torch.Generator().manual_seed(41)
    B, T, Fq = 3, 50, 80
    x = torch.randn(B, T, Fq, generator=g) * 2 + 1
    lens = torch.tensor([50, 33, 41])
    count = 1234.0
    ssum = torch.randn(Fq, generator=g).double() * count
    ssq = (torch.rand(Fq, generator=g).double() * 3 + 0.5) * count + ssum * ssum / count
    out = {"x": x, "lens": lens, "count": np.array(count), "sum": ssum, "sum_square": ssq}
    with tempfile.TemporaryDirectory() as td:
        npz = os.path.join(td, "feats_stats.npz")
        np.savez(npz, count=np.array(count), sum=ssum.numpy(), sum_square=ssq.numpy())
        kaldi = np.zeros((2, Fq + 1))
        kaldi[0, :-1], kaldi[0, -1], kaldi[1, :-1] = ssum.numpy(), count, ssq.numpy()
        npy = os.path.join(td, "cmvn.npy")
        np.save(npy, kaldi)
        for nm, nv in ((True, True), (True, False), (False, True)):
            y, _ = GlobalMVN(npz, norm_means=nm, norm_vars=nv)(x.clone(), lens)
            out[f"y_{int(nm)}{int(nv)}"] = y
        mk = GlobalMVN(npy)
        out["y_kaldi"], _ = mk(x.clone(), lens)
        out["mean"], out["std"] = mk.mean, mk.std
    # SpecAug, ratio-range time masks (train_asr_conformer10_hop_length160.yaml:62-76 style), no time warp
    sa = SpecAug(apply_time_warp=False, apply_freq_mask=True, freq_mask_width_range=(0, 27), num_freq_mask=2,
                 apply_time_mask=True, time_mask_width_ratio_range=(0.0, 0.2), num_time_mask=3)
    xs = torch.randn(B, T, Fq, generator=g)
    torch.manual_seed(43)
    ys, _ = sa(xs.clone(), lens)
    torch.manual_seed(43)
    f_len = torch.randint(0, 27, (B, 2))
    f_pos = torch.randint(0, max(1, Fq - int(f_len.max())), (B, 2))
    hi = math.floor(T * 0.2)
    t_len = torch.randint(0, hi, (B, 3))
    t_pos = torch.randint(0, max(1, T - int(t_len.max())), (B, 3))
    out.update({"sa_x": xs, "sa_y": ys, "f_len": f_len, "f_pos": f_pos, "t_len": t_len, "t_pos": t_pos})
    np.savez_compressed(os.path.join(OUT, "norm_variants.npz"), **_np(out))


def gen_decode():
    """Next-row N2: eval-mode encode() (BN running statistics), CTC.log_softmax / argmax and the decoder's incremental
    scoring interface (batch_score over a growing prefix with the per-layer cache; score for one hypothesis)."""
    torch.manual_seed(7)
    m = ref_shim.build_reference_model(TINY_V, TINY_ENC, TINY_DEC)
    randomize_(m, 8)
    g = torch.Generator().manual_seed(9)
    with torch.no_grad():
        for n, b in m.named_buffers():
            if n.endswith("running_mean"):
                b.copy_(0.2 * torch.randn(b.shape, generator=g))
            elif n.endswith("running_var"):
                b.copy_(0.5 + torch.rand(b.shape, generator=g))
    m.eval()
    wav, lens, _, _ = tiny_batch(seed=10)
    with torch.no_grad():
        enc_out, enc_lens = m.encode(wav, lens)
        ctc_logp = m.ctc.log_softmax(enc_out)
        ctc_arg = m.ctc.argmax(enc_out)
        x = enc_out[0, : int(enc_lens[0])]
        n_hyp, L = 3, 5
        ys = torch.randint(2, TINY_V - 1, (n_hyp, L), generator=g)
        ys[:, 0] = TINY_V - 1  # sos
        xs = x.unsqueeze(0).expand(n_hyp, -1, -1)
        states = [None] * n_hyp
        logps = []
        for l in range(1, L + 1):
            logp, states = m.decoder.batch_score(ys[:, :l], states, xs)
            logps.append(logp)
        cache = torch.stack([torch.stack([states[b][i] for b in range(n_hyp)]) for i in range(len(m.decoder.decoders))])
        st, one = None, []
        for l in range(1, 4):
            lp, st = m.decoder.score(ys[1, :l], st, x)
            one.append(lp)
        # the full (teacher-forced) forward gives the same rows: transformer_decoder.py:92-145
        full, _ = m.decoder(xs, torch.full((n_hyp,), xs.shape[1]), ys, torch.full((n_hyp,), L))
    out = {"wav": wav, "lens": lens, "enc_out": enc_out, "enc_lens": enc_lens, "ctc_logp": ctc_logp, "ctc_argmax": ctc_arg,
           "ys": ys, "logp_steps": torch.stack(logps), "cache": cache, "logp_one": torch.stack(one),
           "full_logp": torch.log_softmax(full, dim=-1)}
    for k, v in m.state_dict().items():
        out["sd/" + k] = v
    np.savez_compressed(os.path.join(OUT, "decode.npz"), **_np(out))
    print("decode.npz", tuple(enc_out.shape), float(torch.stack(logps).abs().max()),
          float((torch.stack(logps)[-1] - torch.log_softmax(full, -1)[:, -1]).abs().max()))


def main():
    warnings.filterwarnings("ignore")
    os.makedirs(OUT, exist_ok=True)
    gen_rel_shift()
    gen_ctc()
    gen_specaug()
    gen_frontend()
    gen_norm_variants()
    gen_decode()
    gen_bn_running()
    gen_e2e("latest", "e2e_latest.npz")
    gen_e2e("legacy", "e2e_legacy.npz")
    gen_e2e("latest", "e2e_interctc.npz", interctc=(0.4, [1, 2]))
    gen_e2e("latest", "e2e_selfcond.npz", interctc=(0.4, [1, 2]), cond=True, V=24)  # V % 8 == 0: a GEMM K dimension
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
