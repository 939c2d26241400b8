"""CPU check of the host logic: the step orchestration (forward + hand-written backward, buffer layouts, fused
projections, gradient re-layouts) run over the plain-torch emulation of the kernels (tests/emul_ops.py) must reproduce
the UNMODIFIED reference's loss and every parameter gradient frozen in tests/golden/e2e_{latest,legacy}.npz."""
import os

import numpy as np
import pytest
import torch

import emul_ops
from espnet_slurp_b200 import params as bparams
from espnet_slurp_b200 import step as bstep
from oracle import model as om

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture()
def emulated(monkeypatch):
    monkeypatch.setattr(bstep, "K", emul_ops)
    monkeypatch.setattr(emul_ops, "exact", True)
    yield
    bstep._PE_CACHE.clear()


def tiny_cfg(rel):
    enc = bstep.EncCfg(80, 16, 2, 32, 2, 7, rel, 0.0, 0.0, 0.0)
    dec = bstep.DecCfg(20, 16, 2, 32, 2, 0.0, 0.0, 0.0, 0.0)
    return bstep.ModelCfg(enc=enc, dec=dec, vocab=20)


def run_step(g, rel):
    sd = {k[3:]: torch.tensor(v) for k, v in g.items() if k.startswith("sd/") and v.dtype == np.float32}
    pack = bparams.ParamPack(sd, lowp=torch.float32)
    pack.refresh()
    pack.zero_grad()
    cfg = tiny_cfg(rel)
    mel = torch.from_numpy(om.slaney_mel_matrix())
    feats, flens = bstep.frontend_fwd(torch.tensor(g["wav"]), torch.tensor(g["lens"]), mel, cfg)
    out = bstep.train_step_from_feats(feats, flens, torch.tensor(g["text"]), torch.tensor(g["text_lens"]), pack.P, pack.W,
                                      pack.G, cfg, bstep.StepCtx(0, True))
    pack.finalize_grads()
    return out, pack


@pytest.mark.parametrize("rel", ["latest", "legacy"])
def test_emulated_step_matches_reference_golden(emulated, rel):
    z = np.load(os.path.join(GOLDEN, f"e2e_{rel}.npz"))
    g = {k: z[k] for k in z.files}
    out, pack = run_step(g, rel)
    np.testing.assert_allclose(out["loss_ctc"].item(), g["loss_ctc"].item(), rtol=5e-5)
    np.testing.assert_allclose(out["loss_att"].item(), g["loss_att"].item(), rtol=5e-5)
    np.testing.assert_allclose(out["loss"].item(), g["loss"].item(), rtol=5e-5)
    assert abs(out["acc"].item() - float(g["acc"].item())) < 1e-6
    worst, worst_name = 0.0, ""
    for k, v in g.items():
        if not k.startswith("grad/"):
            continue
        got = pack.grads()[k[5:]].numpy()
        err = float(np.abs(got - v).max() / max(np.abs(v).max(), 1e-3))
        if err > worst:
            worst, worst_name = err, k
    assert worst < 3e-3, (worst, worst_name)


def test_emulated_step_with_intermediate_ctc(emulated):
    """Next-row N4: intermediate CTC taps on blocks 1 and 2 of 3, against the unmodified reference."""
    z = np.load(os.path.join(GOLDEN, "e2e_interctc.npz"))
    g = {k: z[k] for k in z.files}
    sd = {k[3:]: torch.tensor(v) for k, v in g.items() if k.startswith("sd/") and v.dtype == np.float32}
    pack = bparams.ParamPack(sd, lowp=torch.float32)
    pack.refresh()
    pack.zero_grad()
    enc = bstep.EncCfg(80, 16, 2, 32, 3, 7, "latest", 0.0, 0.0, 0.0, interctc_layer_idx=(1, 2))
    dec = bstep.DecCfg(20, 16, 2, 32, 2, 0.0, 0.0, 0.0, 0.0)
    cfg = bstep.ModelCfg(enc=enc, dec=dec, vocab=20, interctc_weight=0.4)
    mel = torch.from_numpy(om.slaney_mel_matrix())
    feats, flens = bstep.frontend_fwd(torch.tensor(g["wav"]), torch.tensor(g["lens"]), mel, cfg)
    out = bstep.train_step_from_feats(feats, flens, torch.tensor(g["text"]), torch.tensor(g["text_lens"]), pack.P, pack.W,
                                      pack.G, cfg, bstep.StepCtx(0, True))
    pack.finalize_grads()
    for k in ("loss", "loss_ctc", "loss_att", "loss_interctc_layer1", "loss_interctc_layer2"):
        np.testing.assert_allclose(out[k].item(), g[k].item(), rtol=5e-5, err_msg=k)
    worst, worst_name = 0.0, ""
    for k, v in g.items():
        if k.startswith("grad/"):
            got = pack.grads()[k[5:]].numpy()
            err = float(np.abs(got - v).max() / max(np.abs(v).max(), 1e-3))
            if err > worst:
                worst, worst_name = err, k
    assert worst < 3e-3, (worst, worst_name)
    # interctc_weight 0 with taps configured: the taps are computed but carry no loss and no gradient
    pack.zero_grad()
    cfg0 = bstep.ModelCfg(enc=enc, dec=dec, vocab=20, interctc_weight=0.0)
    out0 = bstep.train_step_from_feats(feats, flens, torch.tensor(g["text"]), torch.tensor(g["text_lens"]), pack.P, pack.W,
                                       pack.G, cfg0, bstep.StepCtx(0, True))
    np.testing.assert_allclose(out0["loss_ctc"].item(), g["loss_ctc"].item(), rtol=5e-5)
    assert "loss_interctc_layer1" not in out0


def test_emulated_step_with_self_conditioning(emulated):
    """Self-conditioned intermediate CTC (interctc_use_conditioning) against the unmodified reference, V = 24."""
    z = np.load(os.path.join(GOLDEN, "e2e_selfcond.npz"))
    g = {k: z[k] for k in z.files}
    sd = {k[3:]: torch.tensor(v) for k, v in g.items() if k.startswith("sd/") and v.dtype == np.float32}
    pack = bparams.ParamPack(sd, lowp=torch.float32)
    pack.refresh()
    pack.zero_grad()
    enc = bstep.EncCfg(80, 16, 2, 32, 3, 7, "latest", 0.0, 0.0, 0.0, interctc_layer_idx=(1, 2), interctc_use_conditioning=True,
                       vocab=24)
    dec = bstep.DecCfg(24, 16, 2, 32, 2, 0.0, 0.0, 0.0, 0.0)
    cfg = bstep.ModelCfg(enc=enc, dec=dec, vocab=24, interctc_weight=0.4)
    mel = torch.from_numpy(om.slaney_mel_matrix())
    feats, flens = bstep.frontend_fwd(torch.tensor(g["wav"]), torch.tensor(g["lens"]), mel, cfg)
    out = bstep.train_step_from_feats(feats, flens, torch.tensor(g["text"]), torch.tensor(g["text_lens"]), pack.P, pack.W,
                                      pack.G, cfg, bstep.StepCtx(0, True))
    pack.finalize_grads()
    for k in ("loss", "loss_ctc", "loss_att", "loss_interctc_layer1", "loss_interctc_layer2"):
        np.testing.assert_allclose(out[k].item(), g[k].item(), rtol=5e-5, err_msg=k)
    worst, worst_name = 0.0, ""
    for k, v in g.items():
        if k.startswith("grad/"):
            got = pack.grads()[k[5:]].numpy()
            err = float(np.abs(got - v).max() / max(np.abs(v).max(), 1e-3))
            if err > worst:
                worst, worst_name = err, k
    assert worst < 3e-3, (worst, worst_name)


def test_add_sos_eos_matches_oracle():
    text = torch.tensor([[5, 6, 7, 8], [3, 4, -1, -1], [9, 2, 11, -1]])
    tl = torch.tensor([4, 2, 3])
    a, b = bstep.add_sos_eos(text, tl, 19, 19)
    ra, rb = om.add_sos_eos(text, tl, 19, 19)
    assert torch.equal(a.long(), ra) and torch.equal(b.long(), rb)


def test_param_pack_fusions_are_views():
    d = 16
    sd = {}
    for n in ("q", "k", "v", "out"):
        sd[f"decoder.decoders.0.self_attn.linear_{n}.weight"] = torch.randn(d, d)
        sd[f"decoder.decoders.0.self_attn.linear_{n}.bias"] = torch.randn(d)
        sd[f"decoder.decoders.0.src_attn.linear_{n}.weight"] = torch.randn(d, d)
        sd[f"decoder.decoders.0.src_attn.linear_{n}.bias"] = torch.randn(d)
    pack = bparams.ParamPack(sd)
    qkv = torch.cat([sd[f"decoder.decoders.0.self_attn.linear_{n}.weight"] for n in "qkv"])
    pack.flat16.copy_(pack.flat32.bfloat16())
    assert torch.equal(pack.W["decoder.decoders.0.self_attn.linear_qkv.weight"].float(), qkv.bfloat16().float())
    kvb = torch.cat([sd[f"decoder.decoders.0.src_attn.linear_{n}.bias"] for n in "kv"])
    assert torch.equal(pack.P["decoder.decoders.0.src_attn.linear_kv.bias"], kvb)
    pack.G["decoder.decoders.0.self_attn.linear_qkv.weight"][d:2 * d] += 1.0
    assert torch.all(pack.grads()["decoder.decoders.0.self_attn.linear_k.weight"] == 1.0)
    assert torch.all(pack.grads()["decoder.decoders.0.self_attn.linear_q.weight"] == 0.0)


@pytest.mark.parametrize("rel", ["latest", "legacy"])
def test_emulated_step_with_dropout_matches_oracle_with_replayed_masks(emulated, rel):
    """Dropout ON at every site (encoder 0.1 / 0.1 / 0.1, decoder incl. the self / source attention dropouts the SLURP
    recipes leave at 0): the step's forward and hand-written backward must use the SAME decisions at every site, which
    the oracle replays through its Dropout hooks (tests/dropout_ref.py).  A site or element-index mismatch between a
    forward and its backward gives finite, plausible, wrong gradients -- this is the test that sees it."""
    import dropout_ref
    z = np.load(os.path.join(GOLDEN, f"e2e_{rel}.npz"))
    g = {k: z[k] for k in z.files}
    sd = {k[3:]: torch.tensor(v) for k, v in g.items() if k.startswith("sd/") and v.dtype == np.float32}
    pack = bparams.ParamPack(sd, lowp=torch.float32)
    pack.refresh()
    pack.zero_grad()
    enc = bstep.EncCfg(80, 16, 2, 32, 2, 7, rel, 0.1, 0.1, 0.1)
    dec = bstep.DecCfg(20, 16, 2, 32, 2, 0.1, 0.1, 0.1, 0.1)
    cfg = bstep.ModelCfg(enc=enc, dec=dec, vocab=20)
    mel = torch.from_numpy(om.slaney_mel_matrix())
    feats, flens = bstep.frontend_fwd(torch.tensor(g["wav"]), torch.tensor(g["lens"]), mel, cfg)
    seed = 77
    ctx = bstep.StepCtx(seed, True)
    text, tl = torch.tensor(g["text"]), torch.tensor(g["text_lens"])
    out = bstep.train_step_from_feats(feats, flens, text, tl, pack.P, pack.W, pack.G, cfg, ctx)
    pack.finalize_grads()
    assert len(ctx.sites) == 2 + 2 * 7 + 1 + 2 * 6 and all(p == 0.1 for p, _ in ctx.sites.values())
    assert len({s for _, s in ctx.sites.values()}) == len(ctx.sites)  # one site id per Dropout module
    replay = dropout_ref.DropReplay(ctx.sites, seed)
    osd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    ocfg = om.ModelCfg(enc=om.EncCfg(80, 16, 2, 32, 2, 7, rel), dec=om.DecCfg(20, 16, 2, 32, 2), vocab=20)
    oloss, ostats = om.asr_forward_from_feats(feats, flens.long(), text, tl, osd, ocfg, drop=replay)
    oloss.backward()
    assert replay.used == set(ctx.sites)  # every Dropout of the reference graph has its site in the step, and vice versa
    np.testing.assert_allclose(out["loss"].item(), oloss.item(), rtol=5e-5)
    np.testing.assert_allclose(out["loss_att"].item(), ostats["loss_att"].item(), rtol=5e-5)
    # dropout really changed the result (the goldens are p = 0)
    assert abs(out["loss"].item() - g["loss"].item()) > 1e-3 * abs(g["loss"].item())
    worst, worst_name = 0.0, ""
    for k, v in osd.items():
        want = v.grad.numpy() if v.grad is not None else np.zeros(tuple(v.shape), np.float32)
        got = pack.grads()[k].numpy()
        err = float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-3))
        if err > worst:
            worst, worst_name = err, k
    assert worst < 3e-3, (worst, worst_name)


def test_dropout_ref_keep_rate_and_determinism():
    import dropout_ref
    a = dropout_ref.keep_mask(0.1, 5, 3, 100003)
    assert np.array_equal(a, dropout_ref.keep_mask(0.1, 5, 3, 100003))
    assert abs(a.mean() - 0.9) < 5e-3
    assert not np.array_equal(a, dropout_ref.keep_mask(0.1, 5, 4, 100003))
    assert not np.array_equal(a, dropout_ref.keep_mask(0.1, 6, 3, 100003))
    assert dropout_ref.keep_mask(0.0, 1, 1, 64).all()


@pytest.mark.parametrize("p_drop", [0.0, 0.1])
def test_emulated_step_on_the_fused_attention_formulation(emulated, p_drop):
    """d_k = 64 routes the step through the FUSED attention forward / backward entry points (relattn_fwd / relattn_bwd:
    delta = <dO, O>, dropout(p) regenerated from p, dq_u / dk / dv from dS): the formulation the CUDA kernels implement,
    checked here against autograd over the oracle (with the dropout decisions replayed)."""
    import dropout_ref
    from espnet_slurp_b200.testing import make_sd
    enc = bstep.EncCfg(80, 128, 2, 64, 2, 7, "latest", p_drop, p_drop, p_drop)
    dec = bstep.DecCfg(20, 128, 2, 64, 1, p_drop, p_drop, p_drop, p_drop)
    cfg = bstep.ModelCfg(enc=enc, dec=dec, vocab=20)
    ocfg = om.ModelCfg(enc=om.EncCfg(80, 128, 2, 64, 2, 7, "latest"), dec=om.DecCfg(20, 128, 2, 64, 1), vocab=20)
    sd = make_sd(ocfg.enc, ocfg.dec, 20, seed=11)
    g = torch.Generator().manual_seed(3)
    B, Tin = 3, 70
    feats = torch.randn(B, Tin, 80, generator=g)
    flens = torch.tensor([70, 61, 44])
    tl = torch.tensor([5, 2, 4])
    text = torch.full((B, 5), -1, dtype=torch.long)
    for b in range(B):
        text[b, : tl[b]] = torch.randint(1, 18, (int(tl[b]),), generator=g)
    calls = []
    orig = emul_ops.relattn_bwd
    emul_ops.relattn_bwd = lambda *a, **k: (calls.append(1), orig(*a, **k))[1]
    try:
        pack = bparams.ParamPack(sd, lowp=torch.float32)
        pack.refresh()
        pack.zero_grad()
        ctx = bstep.StepCtx(5, True)
        out = bstep.train_step_from_feats(feats, flens.int(), text, tl, pack.P, pack.W, pack.G, cfg, ctx)
        pack.finalize_grads()
    finally:
        emul_ops.relattn_bwd = orig
    assert len(calls) == 2 + 2  # two encoder blocks, decoder self- and source-attention
    osd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    replay = dropout_ref.DropReplay(ctx.sites, 5) if p_drop > 0 else None
    oloss, _ = om.asr_forward_from_feats(feats, flens, text, tl, osd, ocfg, drop=replay)
    oloss.backward()
    np.testing.assert_allclose(out["loss"].item(), oloss.item(), rtol=5e-5)
    worst, worst_name = 0.0, ""
    for k, v in osd.items():
        want = v.grad.numpy() if v.grad is not None else np.zeros(tuple(v.shape), np.float32)
        err = float(np.abs(pack.grads()[k].numpy() - want).max() / max(np.abs(want).max(), 1e-3))
        if err > worst:
            worst, worst_name = err, k
    assert worst < 3e-3, (worst, worst_name)
