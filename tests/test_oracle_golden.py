"""Pin the CPU oracle (oracle/model.py) against outputs of the UNMODIFIED reference
frozen in tests/golden/*.npz by oracle/make_golden.py.  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import model as om

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    z = np.load(os.path.join(GOLDEN, name))
    return {k: z[k] for k in z.files}


def tiny_cfg(rel):
    return om.ModelCfg(enc=om.EncCfg(80, 16, 2, 32, 2, 7, rel), dec=om.DecCfg(20, 16, 2, 32, 2), vocab=20)


@pytest.mark.parametrize("T", [1, 2, 3, 7, 16, 33, 64])
def test_rel_shift_latest_bit_exact(T):
    g = load("rel_shift.npz")
    code = g[f"latest_{T}"]  # entries = row*10000 + col + 1 of the un-shifted matrix
    col = om.rel_shift_index_latest(T)
    row = np.arange(T)[:, None] + np.zeros((1, T), dtype=np.int64)
    assert np.array_equal(code, row * 10000 + col + 1)


@pytest.mark.parametrize("T", [1, 2, 3, 7, 16, 33, 64])
def test_rel_shift_legacy_bit_exact(T):
    g = load("rel_shift.npz")
    code = g[f"legacy_{T}"]
    row, col = om.rel_shift_index_legacy(T)
    want = np.where(col < 0, 0, row * 10000 + col + 1)
    assert np.array_equal(code, want)


@pytest.mark.parametrize("case", ["loss_a", "loss_b", "ctc_fixture", "infeasible"])
def test_ctc_numpy_vs_torch_fp64(case):
    g = load("ctc.npz")
    ys, yl = g[case + "/ys"], g[case + "/ylens"]
    off = np.concatenate([[0], np.cumsum(yl)])
    labels = [ys[off[i]:off[i + 1]].tolist() for i in range(len(yl))]
    nll, grad = om.ctc_loss_numpy(g[case + "/logits"], g[case + "/hlens"], labels)
    np.testing.assert_allclose(nll, g[case + "/nll"], rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(grad, g[case + "/grad"], rtol=1e-9, atol=1e-10)


def test_ctc_lattice_bit_exact():
    ext, skip = om.ctc_extended_labels([3, 3, 5, 1, 1, 2])
    assert ext.tolist() == [0, 3, 0, 3, 0, 5, 0, 1, 0, 1, 0, 2, 0]
    assert skip.tolist() == [0, 0, 0, 0, 0, 1, 0, 1, 0, 0, 0, 1, 0]


def test_ctc_module():
    g = load("ctc.npz")
    h = torch.tensor(g["module/h"], requires_grad=True)
    sd = {"ctc.ctc_lo.weight": torch.tensor(g["module/w"], requires_grad=True), "ctc.ctc_lo.bias": torch.tensor(g["module/b"])}
    loss = om.ctc_branch(h, torch.tensor([10, 8]), torch.tensor(g["module/y"]), torch.tensor([5, 2]), sd)
    loss.backward()
    np.testing.assert_allclose(loss.item(), g["module/loss"], rtol=1e-6)
    np.testing.assert_allclose(h.grad.numpy(), g["module/grad_h"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(sd["ctc.ctc_lo.weight"].grad.numpy(), g["module/grad_w"], rtol=1e-5, atol=1e-7)


def test_frontend_and_mvn():
    g = load("frontend.npz")
    mel = torch.from_numpy(om.slaney_mel_matrix())
    np.testing.assert_allclose(mel.numpy(), g["melmat"], rtol=0, atol=0)
    feats, flens = om.frontend_logmel(torch.tensor(g["wav"]), torch.tensor(g["lens"]), mel)
    assert flens.tolist() == g["feat_lens"].tolist()
    np.testing.assert_allclose(feats.numpy(), g["feats"], rtol=2e-4, atol=2e-4)
    mvn = om.utterance_mvn(torch.tensor(g["feats"]), flens)
    np.testing.assert_allclose(mvn.numpy(), g["mvn"], rtol=1e-5, atol=1e-5)
    # STFT shape/length fixture of test/espnet2/layers/test_stft.py:10-16
    assert g["stft_shape"].tolist() == [2, 16, 3, 2] and g["stft_lens"].tolist() == [16, 8]
    f2, l2 = om.frontend_logmel(torch.randn(2, 30), torch.tensor([30, 15]),
                                torch.ones(3, 1), n_fft=4, hop=2)
    assert f2.shape[:2] == (2, 16) and l2.tolist() == [16, 8]


def test_mel_matrix_vs_torchaudio():
    ta = pytest.importorskip("torchaudio")
    ref = ta.functional.melscale_fbanks(257, 0.0, 8000.0, 80, 16000, norm="slaney", mel_scale="slaney")
    np.testing.assert_allclose(om.slaney_mel_matrix(), ref.numpy(), rtol=1e-5, atol=1e-7)


def test_specaug_replay():
    g = load("specaug.npz")
    draws = {k: torch.tensor(g[k]) for k in ("center", "warped", "f_len", "f_pos", "t_len", "t_pos")}
    y = om.spec_augment(torch.tensor(g["x"]), torch.tensor(g["lens"]), draws)
    np.testing.assert_allclose(y.numpy(), g["y"], rtol=1e-5, atol=1e-5)


def test_global_mvn_and_ratio_time_masks():
    """N4 rows against the unmodified reference: GlobalMVN (both statistics formats, the three norm settings) and
    SpecAug with time_mask_width_ratio_range (draws replayed)."""
    g = load("norm_variants.npz")
    x, lens = torch.tensor(g["x"]), torch.tensor(g["lens"])
    mean, std = torch.tensor(g["mean"]).float(), torch.tensor(g["std"]).float()
    cnt = float(g["count"])
    np.testing.assert_allclose(mean.numpy(), g["sum"] / cnt, rtol=1e-6)
    np.testing.assert_allclose(std.numpy(), np.sqrt(np.maximum(g["sum_square"] / cnt - (g["sum"] / cnt) ** 2, 1e-20)), rtol=1e-6)
    for nm, nv in ((True, True), (True, False), (False, True)):
        y = om.global_mvn(x, lens, mean, std, nm, nv)
        np.testing.assert_allclose(y.numpy(), g[f"y_{int(nm)}{int(nv)}"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(g["y_kaldi"], g["y_11"], rtol=1e-6, atol=1e-7)
    T = g["sa_x"].shape[1]
    assert om.variable_time_mask_range(T, (0.0, 0.2)) == (0, 10)
    assert om.variable_time_mask_range(4, (0.0, 0.2)) is None
    draws = {k: torch.tensor(g[k]) for k in ("f_len", "f_pos", "t_len", "t_pos")}
    y = om.spec_augment(torch.tensor(g["sa_x"]), lens, draws)
    np.testing.assert_allclose(y.numpy(), g["sa_y"], rtol=1e-6, atol=1e-7)


def test_encoder_and_bn_running_stats():
    g = load("encoder_bn.npz")
    sd = {k[3:]: torch.tensor(v) for k, v in g.items() if k.startswith("sd/")}
    stats = {}
    enc, olens = om.conformer_encoder(torch.tensor(g["feats"]), torch.tensor(g["feat_lens"]), sd,
                                      tiny_cfg("latest").enc, bn_stats_out=stats)
    assert olens.tolist() == g["enc_lens"].tolist()
    np.testing.assert_allclose(enc.numpy(), g["enc_out"], rtol=1e-4, atol=1e-5)
    for p, (mean, var_unb) in stats.items():
        rm = 0.9 * sd[p + ".norm.running_mean"] + 0.1 * mean
        rv = 0.9 * sd[p + ".norm.running_var"] + 0.1 * var_unb
        np.testing.assert_allclose(rm.numpy(), g["after/" + p + ".norm.running_mean"], rtol=1e-4, atol=1e-6)
        np.testing.assert_allclose(rv.numpy(), g["after/" + p + ".norm.running_var"], rtol=1e-4, atol=1e-6)


def test_eval_encode_and_incremental_decoding():
    """Next-row N2: eval-mode encode (BN running statistics), CTC log_softmax / argmax, and the decoder's cached
    one-step scoring against batch_score / score of the unmodified reference (tests/golden/decode.npz)."""
    g = load("decode.npz")
    sd = {k[3:]: torch.tensor(v) for k, v in g.items() if k.startswith("sd/")}
    cfg = tiny_cfg("latest")
    mel = torch.from_numpy(om.slaney_mel_matrix())
    lens = torch.tensor(g["lens"])
    feats, flens = om.frontend_logmel(torch.tensor(g["wav"]), lens, mel)
    enc, olens = om.conformer_encoder(om.utterance_mvn(feats, flens), flens, sd, cfg.enc, training=False)
    assert olens.tolist() == g["enc_lens"].tolist()
    np.testing.assert_allclose(enc.numpy(), g["enc_out"], rtol=1e-4, atol=2e-5)
    logp = torch.log_softmax(om.linear(torch.tensor(g["enc_out"]), sd, "ctc.ctc_lo"), dim=2)
    np.testing.assert_allclose(logp.numpy(), g["ctc_logp"], rtol=1e-5, atol=1e-5)
    assert np.array_equal(logp.argmax(2).numpy(), g["ctc_argmax"])
    ys = torch.tensor(g["ys"])
    x = torch.tensor(g["enc_out"])[0, : int(g["enc_lens"][0])]
    xs = x.unsqueeze(0).expand(ys.shape[0], -1, -1)
    cache = None
    for l in range(1, ys.shape[1] + 1):
        lp, cache = om.decoder_forward_one_step(ys[:, :l], xs, sd, cfg.dec, cache)
        np.testing.assert_allclose(lp.numpy(), g["logp_steps"][l - 1], rtol=1e-4, atol=1e-5)
        # test/espnet2/asr/decoder/test_transformer_decoder.py / test/test_transformer_decode.py invariant:
        # the cached step equals the un-cached one and the teacher-forced forward's row l-1
        lp0, _ = om.decoder_forward_one_step(ys[:, :l], xs, sd, cfg.dec, None)
        np.testing.assert_allclose(lp0.numpy(), lp.numpy(), rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(lp.numpy(), g["full_logp"][:, l - 1], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(torch.stack(cache).numpy(), g["cache"], rtol=1e-4, atol=1e-5)
    st = None
    for l in range(1, 4):
        lp, st = om.decoder_forward_one_step(ys[1:2, :l], x[None], sd, cfg.dec, st)
        np.testing.assert_allclose(lp[0].numpy(), g["logp_one"][l - 1], rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("rel", ["latest", "legacy"])
def test_whole_step_loss_and_grads(rel):
    g = load(f"e2e_{rel}.npz")
    sd = {k[3:]: torch.tensor(v).requires_grad_(v.dtype == np.float32) for k, v in g.items() if k.startswith("sd/")}
    cfg = tiny_cfg(rel)
    mel = torch.from_numpy(om.slaney_mel_matrix())
    loss, stats = om.asr_forward(torch.tensor(g["wav"]), torch.tensor(g["lens"]), torch.tensor(g["text"]),
                                 torch.tensor(g["text_lens"]), sd, cfg, mel)
    np.testing.assert_allclose(loss.item(), g["loss"].item(), rtol=2e-5)
    np.testing.assert_allclose(stats["loss_ctc"].item(), g["loss_ctc"].item(), rtol=2e-5)
    np.testing.assert_allclose(stats["loss_att"].item(), g["loss_att"].item(), rtol=2e-5)
    assert abs(stats["acc"] - float(g["acc"].item())) < 1e-6
    loss.backward()
    worst = 0.0
    for k, v in g.items():
        if not k.startswith("grad/"):
            continue
        got = sd[k[5:]].grad
        got = torch.zeros_like(sd[k[5:]]) if got is None else got
        # linear_k.bias / depthwise bias grads are mathematically 0 (softmax shift, BN): abs floor
        scale = max(np.abs(v).max(), 1e-3)
        worst = max(worst, float(np.abs(got.numpy() - v).max() / scale))
    assert worst < 2e-3, worst


def test_intermediate_ctc_step():
    """Next-row N4: intermediate CTC on blocks 1 and 2 of 3 (interctc_weight 0.4), loss, per-block stats and every
    gradient against the unmodified reference (tests/golden/e2e_interctc.npz)."""
    g = load("e2e_interctc.npz")
    sd = {k[3:]: torch.tensor(v).requires_grad_(v.dtype == np.float32) for k, v in g.items() if k.startswith("sd/")}
    cfg = om.ModelCfg(enc=om.EncCfg(80, 16, 2, 32, 3, 7, "latest"), dec=om.DecCfg(20, 16, 2, 32, 2), vocab=20,
                      interctc_weight=0.4, interctc_layer_idx=(1, 2))
    mel = torch.from_numpy(om.slaney_mel_matrix())
    loss, stats = om.asr_forward(torch.tensor(g["wav"]), torch.tensor(g["lens"]), torch.tensor(g["text"]),
                                 torch.tensor(g["text_lens"]), sd, cfg, mel)
    for k in ("loss_ctc", "loss_att", "loss_interctc_layer1", "loss_interctc_layer2"):
        np.testing.assert_allclose(float(stats[k]), g[k].item(), rtol=2e-5)
    np.testing.assert_allclose(loss.item(), g["loss"].item(), rtol=2e-5)
    loss.backward()
    worst = 0.0
    for k, v in g.items():
        if k.startswith("grad/"):
            got = sd[k[5:]].grad
            got = torch.zeros_like(sd[k[5:]]) if got is None else got
            worst = max(worst, float(np.abs(got.numpy() - v).max() / max(np.abs(v).max(), 1e-3)))
    assert worst < 2e-3, worst


def test_self_conditioned_intermediate_ctc_step():
    """Self-conditioned intermediate CTC (conformer_encoder.py:350-358, interctc_use_conditioning): the CTC posterior of
    blocks 1 and 2 is fed back through encoder.conditioning_layer; loss, stats and every gradient against the unmodified
    reference (tests/golden/e2e_selfcond.npz, V = 24)."""
    g = load("e2e_selfcond.npz")
    sd = {k[3:]: torch.tensor(v).requires_grad_(v.dtype == np.float32) for k, v in g.items() if k.startswith("sd/")}
    assert sd["encoder.conditioning_layer.weight"].shape == (16, 24)
    cfg = om.ModelCfg(enc=om.EncCfg(80, 16, 2, 32, 3, 7, "latest"), dec=om.DecCfg(24, 16, 2, 32, 2), vocab=24,
                      interctc_weight=0.4, interctc_layer_idx=(1, 2), interctc_use_conditioning=True)
    mel = torch.from_numpy(om.slaney_mel_matrix())
    loss, stats = om.asr_forward(torch.tensor(g["wav"]), torch.tensor(g["lens"]), torch.tensor(g["text"]),
                                 torch.tensor(g["text_lens"]), sd, cfg, mel)
    for k in ("loss_ctc", "loss_att", "loss_interctc_layer1", "loss_interctc_layer2"):
        np.testing.assert_allclose(float(stats[k]), g[k].item(), rtol=2e-5)
    np.testing.assert_allclose(loss.item(), g["loss"].item(), rtol=2e-5)
    loss.backward()
    worst = 0.0
    for k, v in g.items():
        if k.startswith("grad/"):
            got = sd[k[5:]].grad
            got = torch.zeros_like(sd[k[5:]]) if got is None else got
            worst = max(worst, float(np.abs(got.numpy() - v).max() / max(np.abs(v).max(), 1e-3)))
    assert worst < 2e-3, worst


def test_library_ops_switch_agrees_with_the_spelled_out_arithmetic(monkeypatch):
    """oracle.LIBRARY_OPS = True routes LayerNorm / Linear / depthwise conv / BatchNorm / GLU / STFT / bicubic resize through
    the torch library ops the reference's modules call (bench.py's same-box eager-PyTorch leg); same loss and gradients."""
    g = load("e2e_latest.npz")
    sd0 = {k[3:]: torch.tensor(v) for k, v in g.items() if k.startswith("sd/") and v.dtype == np.float32}
    cfg = om.ModelCfg(enc=om.EncCfg(80, 16, 2, 32, 2, 7, "latest"), dec=om.DecCfg(20, 16, 2, 32, 2), vocab=20)
    mel = torch.from_numpy(om.slaney_mel_matrix())
    wav, lens = torch.tensor(g["wav"]), torch.tensor(g["lens"])
    draws = {"center": torch.tensor([20, 30, -1][: wav.shape[0]]), "warped": torch.tensor([22, 27, 0][: wav.shape[0]]),
             "f_pos": torch.tensor([[3, 40]] * wav.shape[0]), "f_len": torch.tensor([[5, 9]] * wav.shape[0]),
             "t_pos": torch.tensor([[2, 30]] * wav.shape[0]), "t_len": torch.tensor([[4, 7]] * wav.shape[0])}
    res = []
    for flag in (False, True):
        monkeypatch.setattr(om, "LIBRARY_OPS", flag)
        sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
        loss, _ = om.asr_forward(wav, lens, torch.tensor(g["text"]), torch.tensor(g["text_lens"]), sd, cfg, mel, draws=draws)
        loss.backward()
        res.append((loss.item(), {k: v.grad.clone() for k, v in sd.items() if v.grad is not None}))
    np.testing.assert_allclose(res[0][0], res[1][0], rtol=2e-5)
    gmax = max(v.abs().max().item() for v in res[0][1].values())
    for k, v in res[0][1].items():  # (gradients that are mathematically zero -- BatchNorm cancels the depthwise bias -- are noise)
        assert (res[1][1][k] - v).abs().max() <= 2e-4 * max(v.abs().max().item(), 1e-3 * gmax), k
