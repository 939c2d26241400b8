"""Drop-in modules (espnet_slurp_b200.modules) on the GPU: same fixtures as the reference's own module tests
(test/espnet2/asr/encoder/test_conformer_encoder.py:40-63, test/espnet2/asr/decoder/test_transformer_decoder.py:34-50,
test/espnet2/asr/test_ctc.py:7-14, test/espnet2/asr/test_espnet_model.py:13-38), checked against the CPU oracle with
the bf16 tolerances stated in tests/test_step_gpu.py."""
import numpy as np
import pytest
import torch

from oracle import model as om
from test_step_gpu import grad_report, make_sd

pytestmark = pytest.mark.gpu


def M():
    import espnet_slurp_b200.modules as m
    return m


def _load(module, sd, prefix):
    own = module.state_dict()
    new = {k: (sd[prefix + k] if prefix + k in sd else v) for k, v in own.items()}
    module.load_state_dict(new)


def _grads(module, prefix):
    return {prefix + n: (p.grad if p.grad is not None else torch.zeros_like(p)) for n, p in module.named_parameters()}


def _batch(B=3, Tin=150, V=50, seed=5):
    g = torch.Generator().manual_seed(seed)
    feats = torch.randn(B, Tin, 80, generator=g)
    flens = torch.tensor([Tin, Tin - 19, Tin - 60][:B])
    for b in range(B):
        feats[b, flens[b]:] = 0.3
    tl = torch.tensor([9, 4, 6][:B])
    text = torch.full((B, 9), -1, dtype=torch.long)
    for b in range(B):
        text[b, : tl[b]] = torch.randint(1, V - 1, (int(tl[b]),), generator=g)
    return feats, flens, text, tl


@pytest.mark.parametrize("rel", ["latest", "legacy"])
def test_conformer_encoder_forward_backward(rel):
    m = M()
    d, H, F, K, nb = 256, 4, 512, 15, 2
    enc = m.ConformerEncoder(80, output_size=d, attention_heads=H, linear_units=F, num_blocks=nb, dropout_rate=0.0,
                             positional_dropout_rate=0.0, attention_dropout_rate=0.0, macaron_style=True, rel_pos_type=rel,
                             cnn_module_kernel=K).cuda()
    ocfg = om.EncCfg(80, d, H, F, nb, K, rel)
    sd = make_sd(ocfg, om.DecCfg(20, d, H, F, 0), 20, seed=1)
    _load(enc, sd, "encoder.")
    feats, flens, _, _ = _batch()
    x = feats.cuda()
    y, olens, _none = enc(x, flens.cuda())
    assert _none is None and y.shape == (3, 36, d) and olens.dtype == torch.long
    w = torch.randn(y.shape, generator=torch.Generator().manual_seed(2))
    (y * w.cuda()).sum().backward()
    osd = {k: v.clone().requires_grad_(True) for k, v in sd.items() if k.startswith("encoder.")}
    oy, oolens = om.conformer_encoder(feats, flens, osd, ocfg)
    (oy * w).sum().backward()
    assert oolens.tolist() == olens.tolist()
    err = (y.detach().cpu() - oy.detach()).abs().max().item() / oy.abs().max().item()
    assert err < 0.03, err
    want = {k: (v.grad.numpy() if v.grad is not None else np.zeros(v.shape, np.float32)) for k, v in osd.items()}
    rows = grad_report(_grads(enc, "encoder."), want)
    assert rows[0][0] < 0.08, rows[:6]
    # BatchNorm running statistics moved (train mode) and eval mode runs on them
    assert float(enc.encoders[0].conv_module.norm.num_batches_tracked) == 1
    enc.eval()
    with torch.no_grad():
        ye, _, _ = enc(x, flens.cuda())
    assert torch.isfinite(ye).all()


def test_encoder_too_short_and_unsupported():
    m = M()
    enc = m.ConformerEncoder(80, 64, 4, 128, 1, macaron_style=True, rel_pos_type="latest").cuda()
    with pytest.raises(m.TooShortUttError) as e:
        enc(torch.randn(2, 6, 80).cuda(), torch.tensor([6, 5]).cuda())
    assert e.value.actual_size == 6 and e.value.limit == 7


def test_transformer_decoder_forward_backward():
    m = M()
    d, H, F, V, nb = 256, 4, 512, 50, 2
    dec = m.TransformerDecoder(V, d, attention_heads=H, linear_units=F, num_blocks=nb, dropout_rate=0.0,
                               positional_dropout_rate=0.0).cuda()
    ocfg = om.DecCfg(V, d, H, F, nb)
    sd = make_sd(om.EncCfg(80, d, H, F, 0, 15, "latest"), ocfg, V, seed=4)
    _load(dec, sd, "decoder.")
    g = torch.Generator().manual_seed(9)
    B, T, L = 3, 40, 10
    mem = torch.randn(B, T, d, generator=g)
    mlens = torch.tensor([40, 33, 20])
    ys = torch.randint(1, V - 1, (B, L), generator=g)
    ylens = torch.tensor([10, 7, 5])
    memc = mem.cuda().requires_grad_(True)
    out, olens = dec(memc, mlens.cuda(), ys.cuda(), ylens.cuda())
    assert out.shape == (B, L, V) and olens.tolist() == ylens.tolist()
    w = torch.randn(B, L, V, generator=g) * (torch.arange(L)[None, :, None] < ylens[:, None, None])
    (out * w.cuda()).sum().backward()
    osd = {k: v.clone().requires_grad_(True) for k, v in sd.items() if k.startswith("decoder.")}
    omem = mem.clone().requires_grad_(True)
    oout = om.transformer_decoder(omem, mlens, ys, ylens, osd, ocfg)
    oout = oout[0] if isinstance(oout, tuple) else oout
    (oout * w).sum().backward()
    valid = (torch.arange(L)[None, :] < ylens[:, None])
    err = ((out.detach().cpu() - oout.detach()).abs() * valid[..., None]).max().item() / oout.abs().max().item()
    assert err < 0.03, err
    want = {k: (v.grad.numpy() if v.grad is not None else np.zeros(v.shape, np.float32)) for k, v in osd.items()}
    rows = grad_report(_grads(dec, "decoder."), want)
    assert rows[0][0] < 0.08, rows[:6]
    mvalid = (torch.arange(T)[None, :] < mlens[:, None])[..., None]
    gm, go = memc.grad.cpu() * mvalid, omem.grad * mvalid
    assert ((gm - go).norm() / go.norm()).item() < 0.05


@pytest.mark.parametrize("hl,yl", [([11, 17, 15], [4, 2, 3]), ([4], [1])])  # test/test_loss.py:12-48 length sets, V=7
def test_ctc_module(hl, yl):
    m = M()
    V, d, B, T = 7, 64, len(hl), max(hl)
    ctc = m.CTC(V, d).cuda()
    g = torch.Generator().manual_seed(3)
    hs = torch.randn(B, T, d, generator=g)
    ys = torch.full((B, max(yl)), -1, dtype=torch.long)
    for b in range(B):
        ys[b, : yl[b]] = torch.randint(1, V, (yl[b],), generator=g)
    hsc = hs.cuda().requires_grad_(True)
    loss = ctc(hsc, torch.tensor(hl).cuda(), ys.cuda(), torch.tensor(yl).cuda())
    (loss * 2.0).backward()
    W = ctc.ctc_lo.weight.detach().cpu().bfloat16().float().requires_grad_(True)
    bvec = ctc.ctc_lo.bias.detach().cpu().requires_grad_(True)
    hso = hs.bfloat16().float().requires_grad_(True)
    lp = torch.log_softmax(hso @ W.t() + bvec, -1).transpose(0, 1)
    tgt = torch.cat([ys[b, : yl[b]] for b in range(B)])
    ref = torch.nn.functional.ctc_loss(lp, tgt, torch.tensor(hl), torch.tensor(yl), reduction="sum", zero_infinity=True) / B
    (ref * 2.0).backward()
    np.testing.assert_allclose(loss.item(), ref.item(), rtol=2e-3)
    assert ((hsc.grad.cpu() - hso.grad).norm() / hso.grad.norm()).item() < 0.02
    assert ((ctc.ctc_lo.weight.grad.cpu() - W.grad).norm() / W.grad.norm()).item() < 0.02
    assert ((ctc.ctc_lo.bias.grad.cpu() - bvec.grad).norm() / bvec.grad.norm()).item() < 0.02
    assert ctc.argmax(hs.cuda()).shape == (B, T) and ctc.log_softmax(hs.cuda()).shape == (B, T, V)


def test_espnet_asr_model_from_waveform():
    m = M()
    d, H, F, K, V = 256, 4, 512, 15, 50
    enc = m.ConformerEncoder(80, d, H, F, 2, 0.0, 0.0, 0.0, macaron_style=True, rel_pos_type="latest", cnn_module_kernel=K)
    dec = m.TransformerDecoder(V, d, H, F, 1, 0.0, 0.0)
    ctc = m.CTC(V, d)
    tokens = ["<blank>", "<unk>"] + [f"t{i}" for i in range(V - 3)] + ["<sos/eos>"]
    model = m.ESPnetASRModel(V, tokens, m.DefaultFrontend(), None, m.UtteranceMVN(), None, enc, None, dec, ctc, None,
                             ctc_weight=0.3, lsm_weight=0.1).cuda()
    ocfg = om.ModelCfg(enc=om.EncCfg(80, d, H, F, 2, K, "latest"), dec=om.DecCfg(V, d, H, F, 1), vocab=V)
    sd = make_sd(ocfg.enc, ocfg.dec, V, seed=7)
    _load(model, sd, "")
    g = torch.Generator().manual_seed(11)
    B = 3
    lens = torch.tensor([16000, 14000, 9000])
    wav = torch.randn(B, 16000, generator=g) * 0.1
    for b in range(B):
        wav[b, lens[b]:] = 0
    _, _, text, tl = _batch(V=V)
    loss, stats, weight = model(wav.cuda(), lens.cuda(), text.cuda(), tl.cuda())
    assert loss.shape == (1,) and weight.tolist() == [B] and set(stats) >= {"loss", "loss_att", "loss_ctc", "acc", "cer", "wer"}
    (loss * 3.0).backward()
    osd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    mel = torch.from_numpy(om.slaney_mel_matrix()) if isinstance(om.slaney_mel_matrix(), np.ndarray) else om.slaney_mel_matrix()
    oloss, ostats = om.asr_forward(wav, lens, text, tl, osd, ocfg, mel)
    (oloss * 3.0).backward()
    np.testing.assert_allclose(stats["loss_ctc"].item(), ostats["loss_ctc"].item(), rtol=2e-2)
    np.testing.assert_allclose(stats["loss_att"].item(), ostats["loss_att"].item(), rtol=2e-2)
    np.testing.assert_allclose(loss.item(), oloss.item(), rtol=2e-2)
    want = {k: (v.grad.numpy() if v.grad is not None else np.zeros(v.shape, np.float32)) for k, v in osd.items()}
    rows = grad_report(_grads(model, ""), want)
    assert rows[0][0] < 0.08, rows[:6]
    # a stale backward is an error, not a silent wrong gradient
    l1, _, _ = model(wav.cuda(), lens.cuda(), text.cuda(), tl.cuda())
    l2, _, _ = model(wav.cuda(), lens.cuda(), text.cuda(), tl.cuda())
    with pytest.raises(RuntimeError):
        l1.sum().backward()
    # eval / no_grad forward
    model.eval()
    with torch.no_grad():
        le, se, _ = model(wav.cuda(), lens.cuda(), text.cuda(), tl.cuda())
    assert torch.isfinite(le).all()


def test_encode_and_incremental_scoring_golden(golden_dir):
    """Next-row N2 against the unmodified reference (tests/golden/decode.npz): eval-mode ESPnetASRModel.encode from raw
    waveforms (BatchNorm running statistics), CTC.log_softmax / argmax, TransformerDecoder.batch_score over a growing
    prefix with the per-layer state lists, score for one hypothesis; bf16 tolerances as in tests/test_step_gpu.py."""
    import os
    m = M()
    z = np.load(os.path.join(golden_dir, "decode.npz"))
    g = {k: z[k] for k in z.files}
    sd = {k[3:]: torch.tensor(v) for k, v in g.items() if k.startswith("sd/")}
    V = 20
    enc = m.ConformerEncoder(80, 16, 2, 32, 2, 0.1, 0.1, 0.1, macaron_style=True, rel_pos_type="latest", cnn_module_kernel=7)
    dec = m.TransformerDecoder(V, 16, 2, 32, 2, 0.1, 0.1)
    ctc = m.CTC(V, 16)
    tokens = ["<blank>", "<unk>"] + [f"t{i}" for i in range(V - 3)] + ["<sos/eos>"]
    model = m.ESPnetASRModel(V, tokens, m.DefaultFrontend(), m.SpecAug(time_mask_width_range=(0, 40)), m.UtteranceMVN(), None, enc,
                             None, dec, ctc, None)
    model.load_state_dict(sd)
    model = model.cuda().eval()
    enc_out, enc_lens = model.encode(torch.tensor(g["wav"]).cuda(), torch.tensor(g["lens"]).cuda())
    assert enc_lens.dtype == torch.long and enc_lens.tolist() == g["enc_lens"].tolist()
    assert tuple(enc_out.shape) == g["enc_out"].shape
    for b, L in enumerate(g["enc_lens"]):  # rows past the length are padding the reference does not define either
        err = np.abs(enc_out[b, :L].cpu().numpy() - g["enc_out"][b, :L]).max()
        assert err < 0.06 * max(1.0, np.abs(g["enc_out"][b, :L]).max()), err
    want = torch.tensor(g["enc_out"]).cuda()
    lp = model.ctc.log_softmax(want)
    np.testing.assert_allclose(lp.cpu().numpy(), g["ctc_logp"], atol=0.05)
    am = model.ctc.argmax(want).cpu().numpy()
    # bf16 logits may flip near-ties: the chosen class must be within the tolerance of the best one
    ref_lp = torch.tensor(g["ctc_logp"])
    chosen = ref_lp.gather(2, torch.tensor(am)[..., None])[..., 0]
    assert float((ref_lp.max(2).values - chosen).max()) < 0.05
    assert (am == g["ctc_argmax"]).mean() > 0.9
    # the model's flat buffers serve the sub-modules' scoring calls: no second pack
    assert "_pp" not in model.decoder.__dict__ and "_pp" not in model.ctc.__dict__

    ys = torch.tensor(g["ys"]).cuda()
    x = want[0, : int(g["enc_lens"][0])]
    xs = x.unsqueeze(0).expand(ys.shape[0], -1, -1)
    states = [None] * ys.shape[0]
    for l in range(1, ys.shape[1] + 1):
        logp, states = model.decoder.batch_score(ys[:, :l], states, xs)
        assert logp.shape == (ys.shape[0], V) and len(states) == ys.shape[0] and len(states[0]) == 2
        assert states[0][0].shape == (l, 16)
        np.testing.assert_allclose(logp.cpu().numpy(), g["logp_steps"][l - 1], atol=0.06)
        np.testing.assert_allclose(logp.exp().sum(-1).cpu().numpy(), 1.0, rtol=1e-4)
    cache = torch.stack([torch.stack([states[b][i] for b in range(ys.shape[0])]) for i in range(2)])
    err = (cache.cpu() - torch.tensor(g["cache"])).abs().max().item()
    assert err < 0.06 * max(1.0, float(np.abs(g["cache"]).max())), err
    st = None
    for l in range(1, 4):
        lp1, st = model.decoder.score(ys[1, :l], st, x)
        assert lp1.shape == (V,)
        np.testing.assert_allclose(lp1.cpu().numpy(), g["logp_one"][l - 1], atol=0.06)
    # a cache of the wrong length is refused rather than ignored
    with pytest.raises(AssertionError):
        model.decoder.forward_one_step(ys[:, :3], None, xs, cache=[torch.zeros(3, 1, 16).cuda()] * 2)
    # stand-alone decoder (not inside a model) packs its own parameters
    dec2 = m.TransformerDecoder(V, 16, 2, 32, 2).cuda().eval()
    _load(dec2, {k: v.cuda() for k, v in sd.items()}, "decoder.")
    lp2, _ = dec2.forward_one_step(ys, None, xs)
    np.testing.assert_allclose(lp2.cpu().numpy(), g["logp_steps"][-1], atol=0.06)
    # training mode keeps working after the scoring calls (same flat buffers)
    model.train()
    _, _, text, tl = _batch(V=V)
    loss, _, _ = model(torch.tensor(g["wav"]).cuda(), torch.tensor(g["lens"]).cuda(), text[:, :4].clamp(max=V - 2).cuda(),
                       torch.tensor([4, 4, 4]).cuda())
    loss.backward()
    assert torch.isfinite(loss).all() and all(p.grad is not None for p in model.decoder.parameters())


def test_intermediate_ctc_modules_golden(golden_dir):
    """Next-row N4 at the module level against the unmodified reference (tests/golden/e2e_interctc.npz): the fused model with
    interctc_weight, and the stand-alone encoder returning ((out, [(idx, tap), ...]), olens, None) with gradients
    flowing back through the taps."""
    import os
    m = M()
    z = np.load(os.path.join(golden_dir, "e2e_interctc.npz"))
    g = {k: z[k] for k in z.files}
    sd = {k[3:]: torch.tensor(v) for k, v in g.items() if k.startswith("sd/")}
    V = 20

    def build():
        enc = m.ConformerEncoder(80, 16, 2, 32, 3, 0.0, 0.0, 0.0, macaron_style=True, rel_pos_type="latest", cnn_module_kernel=7,
                                 interctc_layer_idx=[1, 2])
        dec = m.TransformerDecoder(V, 16, 2, 32, 2, 0.0, 0.0)
        tokens = ["<blank>", "<unk>"] + [f"t{i}" for i in range(V - 3)] + ["<sos/eos>"]
        model = m.ESPnetASRModel(V, tokens, m.DefaultFrontend(), None, m.UtteranceMVN(), None, enc, None, dec, m.CTC(V, 16), None,
                                 ctc_weight=0.3, interctc_weight=0.4, lsm_weight=0.1)
        _load(model, sd, "")
        return model.cuda()

    model = build()
    wav, lens = torch.tensor(g["wav"]).cuda(), torch.tensor(g["lens"]).cuda()
    text, tl = torch.tensor(g["text"]).cuda(), torch.tensor(g["text_lens"]).cuda()
    loss, stats, _ = model(wav, lens, text, tl)
    loss.backward()
    for k in ("loss", "loss_ctc", "loss_att", "loss_interctc_layer1", "loss_interctc_layer2"):
        np.testing.assert_allclose(stats[k].item(), g[k].item(), rtol=2e-2, err_msg=k)
    assert stats["cer_interctc_layer1"] is None and list(stats).index("loss_interctc_layer1") < list(stats).index("loss_att")
    want = {k[5:]: v for k, v in g.items() if k.startswith("grad/")}
    rows = grad_report(_grads(model, ""), want)
    assert rows[0][0] < 0.10, rows[:6]

    # module by module, composed the way the reference's ESPnetASRModel.forward does (espnet_model.py:202-245)
    model2 = build()
    feats, flens = model2.normalize(*model2.frontend(wav, lens))
    (enc_out, taps), olens, _ = model2.encoder(feats, flens)
    assert [i for i, _ in taps] == [1, 2] and taps[0][1].shape == enc_out.shape
    l_ctc = model2.ctc(enc_out, olens, text, tl)
    l_ic = sum(model2.ctc(t, olens, text, tl) for _, t in taps) / len(taps)
    np.testing.assert_allclose(l_ctc.item(), g["loss_ctc"].item(), rtol=2e-2)
    np.testing.assert_allclose(l_ic.item(), 0.5 * (g["loss_interctc_layer1"].item() + g["loss_interctc_layer2"].item()), rtol=2e-2)
    ((1 - 0.4) * l_ctc + 0.4 * l_ic).backward()
    # encoder gradients of the CTC terms alone: compare with the oracle's autograd on the same expression
    osd = {k: v.clone().requires_grad_(v.dtype == torch.float32) for k, v in sd.items()}
    ocfg = om.ModelCfg(enc=om.EncCfg(80, 16, 2, 32, 3, 7, "latest"), dec=om.DecCfg(20, 16, 2, 32, 2), vocab=V)
    of, ofl = om.frontend_logmel(torch.tensor(g["wav"]), torch.tensor(g["lens"]), torch.from_numpy(om.slaney_mel_matrix()))
    oenc, oolens, blocks = om.conformer_encoder(om.utterance_mvn(of, ofl), ofl, osd, ocfg.enc, return_blocks=True)
    ot, otl = torch.tensor(g["text"]), torch.tensor(g["text_lens"])
    o_ic = sum(om.ctc_branch(om.layer_norm(blocks[i - 1], osd, "encoder.after_norm"), oolens, ot, otl, osd) for i in (1, 2)) / 2
    (0.6 * om.ctc_branch(oenc, oolens, ot, otl, osd) + 0.4 * o_ic).backward()
    want2 = {k: (v.grad.numpy() if v.grad is not None else np.zeros(tuple(v.shape), np.float32)) for k, v in osd.items()
             if k.startswith(("encoder.", "ctc.")) and v.dtype == torch.float32}
    got2 = {k: v for k, v in _grads(model2, "").items() if k in want2}
    rows = grad_report(got2, want2)
    assert rows[0][0] < 0.10, rows[:6]


def test_self_conditioned_model_golden(golden_dir):
    """ESPnetASRModel with interctc_use_conditioning: the model creates encoder.conditioning_layer like the reference
    (espnet_model.py:96-101), the reference's state_dict loads, loss / stats / gradients follow tests/golden/e2e_selfcond.npz."""
    import os
    m = M()
    z = np.load(os.path.join(golden_dir, "e2e_selfcond.npz"))
    g = {k: z[k] for k in z.files}
    sd = {k[3:]: torch.tensor(v) for k, v in g.items() if k.startswith("sd/")}
    V = 24
    enc = m.ConformerEncoder(80, 16, 2, 32, 3, 0.0, 0.0, 0.0, macaron_style=True, rel_pos_type="latest", cnn_module_kernel=7,
                             interctc_layer_idx=[1, 2], interctc_use_conditioning=True)
    dec = m.TransformerDecoder(V, 16, 2, 32, 2, 0.0, 0.0)
    tokens = ["<blank>", "<unk>"] + [f"t{i}" for i in range(V - 3)] + ["<sos/eos>"]
    model = m.ESPnetASRModel(V, tokens, m.DefaultFrontend(), None, m.UtteranceMVN(), None, enc, None, dec, m.CTC(V, 16), None,
                             ctc_weight=0.3, interctc_weight=0.4, lsm_weight=0.1)
    assert tuple(model.encoder.conditioning_layer.weight.shape) == (16, V)
    assert set(sd) <= set(model.state_dict())
    _load(model, sd, "")
    model = model.cuda()
    loss, stats, _ = model(torch.tensor(g["wav"]).cuda(), torch.tensor(g["lens"]).cuda(), torch.tensor(g["text"]).cuda(),
                           torch.tensor(g["text_lens"]).cuda())
    loss.backward()
    for k in ("loss", "loss_ctc", "loss_att", "loss_interctc_layer1", "loss_interctc_layer2"):
        np.testing.assert_allclose(stats[k].item(), g[k].item(), rtol=2e-2, err_msg=k)
    rows = grad_report(_grads(model, ""), {k[5:]: v for k, v in g.items() if k.startswith("grad/")})
    assert rows[0][0] < 0.10, rows[:6]
    with pytest.raises(NotImplementedError):
        model.encoder(torch.zeros(1, 40, 80).cuda(), torch.tensor([40]).cuda())
    with pytest.raises(NotImplementedError):  # vocabulary must be a multiple of 8 for the posterior GEMM
        m.ESPnetASRModel(20, tokens[:20], None, None, None, None,
                         m.ConformerEncoder(80, 16, 2, 32, 3, macaron_style=True, rel_pos_type="latest", cnn_module_kernel=7,
                                            interctc_layer_idx=[1], interctc_use_conditioning=True),
                         None, m.TransformerDecoder(20, 16, 2, 32, 1), m.CTC(20, 16), None, ctc_weight=0.3)


def test_global_mvn_and_ratio_specaug_modules(golden_dir, tmp_path):
    """GlobalMVN / SpecAug(time_mask_width_ratio_range) drop-ins against outputs of the unmodified reference."""
    import os
    import numpy as np
    m = M()
    g = np.load(os.path.join(golden_dir, "norm_variants.npz"))
    x, lens = torch.tensor(g["x"]).cuda(), torch.tensor(g["lens"]).cuda()
    npz = str(tmp_path / "feats_stats.npz")
    np.savez(npz, count=g["count"], sum=g["sum"], sum_square=g["sum_square"])
    kaldi = np.zeros((2, g["sum"].shape[0] + 1))
    kaldi[0, :-1], kaldi[0, -1], kaldi[1, :-1] = g["sum"], float(g["count"]), g["sum_square"]
    npy = str(tmp_path / "cmvn.npy")
    np.save(npy, kaldi)
    for nm, nv in ((True, True), (True, False), (False, True)):
        y, yl = m.GlobalMVN(npz, norm_means=nm, norm_vars=nv)(x, lens)
        assert yl is lens
        np.testing.assert_allclose(y.cpu().numpy(), g[f"y_{int(nm)}{int(nv)}"], rtol=1e-5, atol=2e-6)
    y, _ = m.GlobalMVN(npy)(x, lens)
    np.testing.assert_allclose(y.cpu().numpy(), g["y_kaldi"], rtol=1e-5, atol=2e-6)
    # the module draws in the reference's order: same seed, same masks, same output
    sa = m.SpecAug(apply_time_warp=False, apply_freq_mask=True, freq_mask_width_range=(0, 27), num_freq_mask=2,
                   apply_time_mask=True, time_mask_width_ratio_range=(0.0, 0.2), num_time_mask=3)
    torch.manual_seed(43)
    ys, _ = sa(torch.tensor(g["sa_x"]).cuda(), lens)
    np.testing.assert_allclose(ys.cpu().numpy(), g["sa_y"], rtol=1e-6, atol=1e-7)
    with pytest.raises(ValueError):
        m.SpecAug(apply_time_mask=True, time_mask_width_range=None, time_mask_width_ratio_range=None)


def test_specaug_and_mvn_modules():
    m = M()
    torch.manual_seed(0)
    sa = m.SpecAug(time_warp_window=5, freq_mask_width_range=(0, 30), time_mask_width_range=(0, 40))
    x = torch.randn(4, 300, 80).cuda()
    lens = torch.tensor([300, 280, 200, 120]).cuda()
    y, yl = sa(x, lens)
    assert y.shape == x.shape and yl is lens
    assert (y == 0).float().mean().item() > 0.01          # masks landed
    z, _ = m.UtteranceMVN()(x, lens)
    for b in range(4):
        assert z[b, : int(lens[b])].mean(0).abs().max().item() < 1e-4


def test_cached_one_step_matches_prefix_rerun_at_real_sizes():
    """N2: forward_one_step with the per-layer cache (last position is the only query, memory keys / values projected once)
    against the uncached prefix re-run and against the oracle's cached step, d = 512 / 8 heads (the fused attention kernel
    with one query row), beam of 5 hypotheses over a 155-frame memory, 7 steps."""
    m = M()
    V, d, H, F, NB = 60, 512, 8, 2048, 2
    ocfg = om.DecCfg(V, d, H, F, NB)
    sd = make_sd(om.EncCfg(80, d, H, F, 0, 15, "latest"), ocfg, V, seed=9)
    dec = m.TransformerDecoder(V, d, H, F, NB).cuda().eval()
    _load(dec, {k: v.cuda() for k, v in sd.items()}, "decoder.")
    g = torch.Generator().manual_seed(1)
    nb, T, L = 5, 155, 7
    mem = torch.randn(1, T, d, generator=g).cuda().expand(nb, -1, -1)
    ys = torch.randint(1, V - 1, (nb, L), generator=g).cuda()
    cache, ocache = None, None
    osd = {k: v for k, v in sd.items() if k.startswith("decoder.")}
    for l in range(1, L + 1):
        lp, cache = dec.forward_one_step(ys[:, :l], None, mem, cache=cache)
        lp_full, cache_full = dec.forward_one_step(ys[:, :l], None, mem, cache=None)
        olp, ocache = om.decoder_forward_one_step(ys[:, :l].cpu(), mem.cpu().contiguous(), osd, ocfg, cache=ocache)
        assert lp.dtype == torch.float32 and lp.shape == (nb, V)
        np.testing.assert_allclose(lp.cpu().numpy(), lp_full.cpu().numpy(), atol=0.05)
        np.testing.assert_allclose(lp.cpu().numpy(), olp.numpy(), atol=0.08)
        for a, b in zip(cache, cache_full):
            assert a.shape == (nb, l, d)
            assert (a - b).abs().max().item() < 0.05 * max(1.0, b.abs().max().item())
    assert "_mem_kv" in dec.__dict__  # the memory projections were computed once and reused
