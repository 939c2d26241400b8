"""Dropout parity on the GPU (the benchmarked configuration runs dropout 0.1 at 123 sites; masks are never stored but
regenerated in the backward from (seed, site, element index) in five different kernels).

1. b200slu_dropout_mask (the decisions every kernel derives) == tests/dropout_ref.py, bit for bit.
2. Every kernel that applies dropout, forward and backward form, against its plain-torch statement fed the same decisions.
3. The whole step with dropout ON at every site against the CPU oracle replaying the decisions (oracle/model.py ``drop``
   hook): loss and every parameter gradient within the bars of tests/test_step_gpu.py.
"""
import math

import numpy as np
import pytest
import torch

import dropout_ref
import emul_ops as E
from oracle import model as om

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


def K():
    from espnet_slurp_b200 import ops
    return ops


@pytest.fixture(autouse=True)
def _exact():
    old = E.exact
    E.exact = True
    yield
    E.exact = old


def rnd(*shape, seed=0, scale=1.0, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed + sum(shape))
    return (torch.randn(*shape, generator=g) * scale).to(dtype)


def close(got, want, tol, what=""):
    got, want = got.detach().float().cpu(), want.detach().float().cpu()
    assert got.shape == want.shape, (what, got.shape, want.shape)
    assert torch.isfinite(got).all(), what
    err = (got - want).abs().max().item() / (want.abs().max().item() + 1e-6)
    assert err < tol, (what, err)


@pytest.mark.parametrize("p,seed,site,n", [(0.1, 0, 1, 4096), (0.1, 1234567890123, 123, 100003), (0.5, 2 ** 63 + 5, 7, 17),
                                           (0.05, 99, 4000000000, 8 * 4099), (0.0, 3, 3, 64)])
def test_exported_decisions_match_cpu_restatement_bit_exact(p, seed, site, n):
    got = K().dropout_mask((p, seed, site), n).cpu().numpy() if p > 0 else np.ones(n, np.uint8)
    want = dropout_ref.keep_mask(p, seed, site, n)
    assert np.array_equal(got, want)


def test_exported_decisions_honour_registered_device_seed():
    sp = torch.zeros(3, device="cuda", dtype=torch.int64)
    sp[0] = 0x1234ABCD5678
    try:
        K().set_step_params(sp)
        got = K().dropout_mask((0.1, 42, 9), 8000).cpu().numpy()
    finally:
        K().set_step_params(None)
    assert np.array_equal(got, dropout_ref.keep_mask(0.1, 42, 9, 8000, dev_seed=0x1234ABCD5678))
    assert not np.array_equal(got, dropout_ref.keep_mask(0.1, 42, 9, 8000))


@pytest.mark.parametrize("M,N,K_,act", [(300, 2048, 512, "swish"), (1000, 512, 2048, None), (257, 512, 512, "relu"),
                                        (9920, 2048, 512, "swish")])
def test_gemm_epilogue_dropout_forward_and_saved_derivative(M, N, K_, act):
    a, b, bias = rnd(M, K_, seed=1).to(BF), (rnd(N, K_, seed=2) / math.sqrt(K_)).to(BF), rnd(N, seed=3)
    res = rnd(M, N, seed=4)
    drop = (0.1, 31337, 17)
    # (the saved derivative rides with a bf16 output of the same row stride: the FFN w_1 form; the fp32 + residual form
    # -- FFN w_2 / attention out / conv pw2 -- is checked right after)
    y, pre = K().gemm(a.cuda(), b.cuda(), bias=bias.cuda(), act=act, act_grad_out=True, drop=drop)
    wy, wpre = E.gemm(a, b, bias=bias, act=act, act_grad_out=True, drop=drop, out_dtype=torch.float32)
    close(y, wy, 1e-2, "gemm y")
    close(pre, wpre, 1e-2, "gemm saved derivative")
    y2 = K().gemm(a.cuda(), b.cuda(), bias=bias.cuda(), drop=(0.1, 31337, 18), alpha=0.5, residual=res.cuda(), out_dtype=torch.float32)
    wy2 = E.gemm(a, b, bias=bias, drop=(0.1, 31337, 18), alpha=0.5, residual=res, out_dtype=torch.float32)
    close(y2, wy2, 2e-3, "gemm y (fp32 + residual)")
    # the dropped elements are EXACTLY the exported ones
    keep = torch.from_numpy(dropout_ref.keep_mask(*drop, M * N).reshape(M, N)).bool()
    prec = pre.float().cpu()
    assert (prec[~keep] == 0).all() and (prec[keep & (wpre != 0)] != 0).all()
    # the dgrad consumes the saved derivative (mask included): dh = (g . W) * pre
    g = rnd(M, 64, seed=5).to(BF)
    w2 = (rnd(64, N, seed=6) / 8).to(BF)
    dh = K().gemm(g.cuda(), w2.cuda(), trans_b=True, act_grad_input=pre, act_grad="mul")
    close(dh, E.gemm(g, w2, trans_b=True, act_grad_input=wpre, act_grad="mul"), 2e-2, "dgrad with saved derivative")


def test_layernorm_bwd_branch_head_and_cast_dropout():
    M, D = 1000, 512
    x, ga = rnd(M, D, seed=1) * 2 + 1, rnd(D, seed=2) + 1
    dy, add = rnd(M, D, seed=4), rnd(M, D, seed=5)
    _, _, st = K().layernorm_fwd(x.cuda(), ga.cuda(), torch.zeros(D, device="cuda"))
    drop = (0.1, 5, 44)
    cs, dg, db = torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda")
    dx, g16 = K().layernorm_bwd(dy.cuda(), x.cuda(), ga.cuda(), st, dg, db, dx_add=add.cuda(), want_g=True, g_alpha=0.5,
                                g_drop=drop, g_colsum_out=cs)
    wcs, wg, wb = torch.zeros(D), torch.zeros(D), torch.zeros(D)
    wdx, wg16 = E.layernorm_bwd(dy, x, ga, st.cpu(), wg, wb, dx_add=add, want_g=True, g_alpha=0.5, g_drop=drop, g_colsum_out=wcs)
    close(dx, wdx, 2e-5, "ln dx")
    close(g16, wg16, 1e-2, "ln branch head g")
    close(cs, wcs, 1e-3, "ln branch head colsum")
    wcs2 = torch.zeros(D)
    close(K().cast_bf16(dx, alpha=0.5, drop=drop), E.cast_bf16(wdx, alpha=0.5, drop=drop, colsum_out=wcs2), 1e-2, "cast")
    pe = rnd(247, 512, seed=9)
    close(K().cast_bf16(pe.cuda(), drop=(0.1, 8, 2)), E.cast_bf16(pe, drop=(0.1, 8, 2)), 1e-2, "cast pos table")


def test_embed_dropout_fwd_bwd():
    V, d, R, L = 50, 64, 60, 12
    emb, pe = rnd(V, d, seed=1), rnd(L, d, seed=2)
    tok = torch.randint(0, V, (R,), generator=torch.Generator().manual_seed(1)).int()
    drop = (0.1, 3, 9)
    close(K().embed_fwd(emb.cuda(), tok.cuda(), pe.cuda(), L, 8.0, drop=drop), E.embed_fwd(emb, tok, pe, L, 8.0, drop=drop), 1e-5,
          "embed fwd")
    dx = rnd(R, d, seed=3)
    de, wde = torch.zeros(V, d, device="cuda"), torch.zeros(V, d)
    K().embed_bwd(dx.cuda(), tok.cuda(), 8.0, de, drop=drop)
    E.embed_bwd(dx, tok, 8.0, wde, drop=drop)
    close(de, wde, 1e-5, "embed bwd")


@pytest.mark.parametrize("mode,B,T,H", [("latest", 3, 61, 2), ("legacy", 2, 50, 4), ("none", 2, 40, 2), ("latest", 2, 300, 2)])
def test_softmax_dropout_fwd_bwd(mode, B, T, H):
    ld = E.round8(T)
    Pn = 2 * T - 1 if mode == "latest" else T
    ac = rnd(H, B, T, ld, seed=1, scale=3.0)
    bd = rnd(H, B, T, E.round8(Pn), seed=2, scale=3.0) if mode != "none" else None
    lens = torch.tensor([T, T - 7, max(T // 2, 1)][:B]).int()
    drop = (0.1, 12345, 77)
    p, pd = K().softmax_fwd(ac.cuda(), bd.cuda() if bd is not None else None, lens.cuda(), T, mode, False, 0.125, drop=drop)
    wp, wpd = E.softmax_fwd(ac, bd, lens, T, mode, False, 0.125, drop=drop)
    close(p[..., :T], wp[..., :T], 1e-2, "p")
    close(pd[..., :T], wpd[..., :T], 1e-2, "dropout(p)")
    keep = torch.from_numpy(dropout_ref.keep_mask(*drop, H * B * T * ld).reshape(H, B, T, ld)).bool()
    assert (pd.float().cpu()[~keep][..., :] == 0).all()
    dpd = rnd(H, B, T, ld, seed=3)
    dac, dbd = K().softmax_bwd(dpd.cuda(), p, T, mode, 0.125, ld_bd=E.round8(Pn), drop=drop)
    wdac, wdbd = E.softmax_bwd(dpd, p.float().cpu(), T, mode, 0.125, ld_bd=E.round8(Pn), drop=drop)
    close(dac[..., :T], wdac[..., :T], 1.5e-2, "d_ac")
    if dbd is not None:
        close(dbd[..., :Pn], wdbd[..., :Pn], 1.5e-2, "d_bd")


@pytest.mark.parametrize("mode,B,T,H,causal", [("latest", 3, 124, 8, False), ("latest", 2, 155, 8, False), ("none", 3, 17, 8, True),
                                               ("none", 2, 20, 8, False)])
def test_fused_attention_dropout_matches_exported_decisions(mode, B, T, H, causal):
    if not K().relattn_supported(T, T, 64, mode):
        pytest.skip("shape not on the fused kernel")
    d = H * 64
    qu, qv = rnd(B * T, d, seed=1, scale=0.5).to(BF), rnd(B * T, d, seed=2, scale=0.5).to(BF)
    kv = rnd(B * T, 2 * d, seed=3, scale=0.5).to(BF)
    pos = rnd(2 * T - 1, d, seed=4, scale=0.5).to(BF) if mode == "latest" else None
    lens = torch.tensor([T, T - 9, T // 2][:B]).int()
    drop = (0.1, 987, 55)
    kc = kv.cuda()
    p, pd, ctx = K().relattn_fwd(qu.cuda(), qv.cuda() if mode == "latest" else None, kc[:, :d], kc[:, d:],
                                 pos.cuda() if pos is not None else None, lens.cuda(), B, T, T, H, mode, causal, 0.125, drop=drop)
    wp, wpd, wctx = E.relattn_fwd(qu, qv, kv[:, :d], kv[:, d:], pos, lens, B, T, T, H, mode, causal, 0.125, drop=drop)
    close(p[..., :T], wp[..., :T], 1e-2, "p")
    close(pd[..., :T], wpd[..., :T], 1.5e-2, "dropout(p)")
    close(ctx, wctx, 2e-2, "context")


# ---------------------------------------------------------------------------------------------- whole step, dropout ON
def _cfgs(V, blocks, dblocks, rel, p):
    from espnet_slurp_b200 import step as bstep
    enc = bstep.EncCfg(80, 512, 8, 2048, blocks, 31, rel, p, p, p)
    dec = bstep.DecCfg(V, 512, 8, 2048, dblocks, p, p, p, p)
    return bstep.ModelCfg(enc=enc, dec=dec, vocab=V), om.ModelCfg(enc=om.EncCfg(80, 512, 8, 2048, blocks, 31, rel),
                                                                   dec=om.DecCfg(V, 512, 8, 2048, dblocks), vocab=V)


def _batch(B, Tin, V, Lmax, seed=5):
    g = torch.Generator().manual_seed(seed)
    feats = torch.randn(B, Tin, 80, generator=g)
    flens = torch.tensor([max(Tin - 19 * b, Tin * 3 // 5) for b in range(B)])
    for b in range(B):
        feats[b, flens[b]:] = 0.3
    tl = torch.randint(2, Lmax + 1, (B,), generator=g)
    tl[0] = Lmax
    text = torch.full((B, Lmax), -1, dtype=torch.long)
    for b in range(B):
        text[b, : tl[b]] = torch.randint(1, V - 1, (int(tl[b]),), generator=g)
    return feats, flens, text, tl


def _block_error_growth(cfg, ocfg, sd, pack, feats, flens, seed, replay):
    """bf16 error of the residual stream after every encoder block (same dropout decisions on both sides), relative to the
    oracle's max-abs at that block; written to gpurun_out/ for DESIGN.md section 2."""
    import json
    import os
    from espnet_slurp_b200 import step as bstep
    B = feats.shape[0]
    ctx = bstep.StepCtx(seed, True)  # the encoder's sites come first in a step: same site ids as in the full step
    enc32, _, _, saved = bstep.encoder_fwd(feats.cuda(), flens.int().cuda(), pack.P, pack.W, cfg.enc, ctx)
    layers, x_last = saved[5], saved[6]
    outs = [layers[n + 1][0][0] for n in range(cfg.enc.blocks - 1)] + [x_last]
    with torch.no_grad():
        oenc, _, oblocks = om.conformer_encoder(feats, flens, {k: v for k, v in sd.items()}, ocfg.enc, return_blocks=True,
                                                drop=replay)
    growth = []
    for n, (got, want) in enumerate(zip(outs, oblocks)):
        got = got.view(B, -1, cfg.enc.d).float().cpu()
        growth.append({"block": n + 1, "max_abs_err_rel": float((got - want).abs().max() / want.abs().max()),
                       "rms_err_rel": float((got - want).pow(2).mean().sqrt() / want.pow(2).mean().sqrt())})
    e = enc32.view(B, -1, cfg.enc.d).float().cpu()
    final = float((e - oenc).abs().max() / oenc.abs().max())
    print("per-block bf16 error growth:", growth, "after_norm:", final)
    assert final < 0.03, final  # the bar of tests/test_step_gpu.py, now at full depth with dropout on
    assert growth[-1]["rms_err_rel"] < 0.02, growth
    try:
        os.makedirs("gpurun_out", exist_ok=True)
        with open("gpurun_out/r2_block_error_growth.json", "w") as f:
            json.dump({"config": "12 enc blocks d=512, B=8, T=124, dropout 0.1 replayed", "blocks": growth,
                       "encoder_out_max_abs_err_rel": final}, f, indent=1)
    except OSError:
        pass


@pytest.mark.parametrize("rel,V,Tin,blocks,dblocks,B", [("latest", 600, 150, 2, 1, 3), ("legacy", 77, 150, 2, 1, 3),
                                                        ("latest", 600, 830, 2, 1, 3), ("latest", 600, 1876, 2, 1, 2),
                                                        ("legacy", 600, 1876, 1, 1, 2), ("latest", 5000, 150, 2, 1, 3),
                                                        ("latest", 600, 501, 12, 6, 8)])
def test_step_with_dropout_vs_oracle_replaying_the_decisions(rel, V, Tin, blocks, dblocks, B):
    """The real layer sizes with dropout 0.1 at EVERY site (decoder attention dropouts too), against the CPU oracle that
    replays the exported decisions.  Tin = 830 -> T = 206 and Tin = 1876 -> T = 468 (BASELINE config 3, 15 s): more than
    192 keys, i.e. the attention path beyond the one-tile fused kernel and the conv module's multi-tile depthwise kernels;
    V = 5000: the LibriSpeech-scale vocabulary (config 4); the last case is the BENCHMARKED model -- 12 encoder + 6 decoder
    blocks, 4 s utterances (Tin = 501 -> T = 124), V = 600 -- at 8 utterances."""
    from espnet_slurp_b200 import params as bparams
    from espnet_slurp_b200 import step as bstep
    from espnet_slurp_b200.testing import make_sd
    from test_step_gpu import grad_report
    cfg, ocfg = _cfgs(V, blocks, dblocks, rel, 0.1)
    sd = make_sd(cfg.enc, cfg.dec, V, seed=3)
    feats, flens, text, tl = _batch(B, Tin, V, 9)
    pack = bparams.ParamPack({k: v.cuda() for k, v in sd.items()})
    pack.refresh()
    pack.zero_grad()
    seed = 20261105
    ctx = bstep.StepCtx(seed, True)
    out = bstep.train_step_from_feats(feats.cuda(), flens.int().cuda(), text.cuda(), tl.cuda(), pack.P, pack.W, pack.G, cfg, ctx)
    pack.finalize_grads()
    torch.cuda.synchronize()
    # the oracle replays the decisions the DEVICE exports (not the CPU restatement: one less link in the chain)
    replay = dropout_ref.DropReplay(ctx.sites, seed, mask_fn=lambda p, s, site, n: K().dropout_mask((p, s, site), n).cpu().numpy())
    osd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    oloss, ostats = om.asr_forward_from_feats(feats, flens, text, tl, osd, ocfg, drop=replay)
    oloss.backward()
    assert replay.used == set(ctx.sites)
    print("loss", out["loss"].item(), oloss.item(), "ctc", out["loss_ctc"].item(), ostats["loss_ctc"].item(), "att",
          out["loss_att"].item(), ostats["loss_att"].item())
    np.testing.assert_allclose(out["loss_ctc"].item(), ostats["loss_ctc"].item(), rtol=2e-2)
    np.testing.assert_allclose(out["loss_att"].item(), ostats["loss_att"].item(), rtol=2e-2)
    want = {k: (v.grad.numpy() if v.grad is not None else np.zeros(tuple(v.shape), np.float32)) for k, v in osd.items()}
    rows = grad_report(pack.grads(), want)
    print("worst grads:", rows[:8])
    assert rows[0][0] < 0.08, rows[:8]
    if blocks >= 12:
        _block_error_growth(cfg, ocfg, sd, pack, feats, flens, seed, replay)
        return
    # and the no-dropout oracle is far away: the comparison above is not vacuous
    osd0 = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    l0, _ = om.asr_forward_from_feats(feats, flens, text, tl, osd0, ocfg)
    l0.backward()
    rows0 = grad_report(pack.grads(), {k: (v.grad.numpy() if v.grad is not None else np.zeros(tuple(v.shape), np.float32))
                                       for k, v in osd0.items()})
    assert rows0[0][0] > 0.2, rows0[:3]
