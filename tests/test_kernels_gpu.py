"""Every CUDA kernel behind the C ABI against its plain-torch statement (tests/emul_ops.py) on the same inputs (GPU)."""
import math

import numpy as np
import pytest
import torch

import emul_ops as E

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


def K():
    from espnet_slurp_b200 import ops
    return ops


@pytest.fixture(autouse=True)
def _exact():
    old = E.exact
    E.exact = True  # emulate in fp32; the tolerance below covers the kernels' bf16 storage
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


# ---------------------------------------------------------------------------------------------- GEMM variants
@pytest.mark.parametrize("ta,tb", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K_", [(304, 200, 136), (128, 64, 64), (1000, 512, 520), (72, 40, 72)])
def test_gemm_majorness(ta, tb, M, N, K_):
    a = rnd(K_, M, seed=1).to(BF) if ta else rnd(M, K_, seed=1).to(BF)
    b = rnd(K_, N, seed=2).to(BF) if tb else rnd(N, K_, seed=2).to(BF)
    out = K().gemm(a.cuda(), b.cuda(), trans_a=ta, trans_b=tb, out_dtype=torch.float32)
    close(out, E.gemm(a, b, trans_a=ta, trans_b=tb, out_dtype=torch.float32), 1e-4, "gemm")


@pytest.mark.parametrize("B,Tx,Ty,H,dk", [(3, 50, 50, 2, 8), (4, 124, 124, 8, 64), (2, 17, 130, 4, 64), (2, 300, 300, 8, 64)])
def test_attention_contractions(B, Tx, Ty, H, dk):
    d = H * dk
    x = rnd(B * Tx, 3 * d, seed=3).to(BF)
    y = rnd(B * Ty, 2 * d, seed=4).to(BF)
    xc, yc = x.cuda(), y.cuda()
    xs, ys = x[:, d:2 * d], y[:, d:]
    s = K().attn_f1(xc[:, d:2 * d], yc[:, d:], B, Tx, Ty, H)
    want = E.attn_f1(xs, ys, B, Tx, Ty, H)
    close(s[..., :Ty], want[..., :Ty], 1e-4, "f1")
    yb = rnd(Ty, d, seed=5).to(BF)
    sb = K().attn_f1(xc[:, d:2 * d], yb.cuda(), B, Tx, Ty, H, y_bcast=True)
    close(sb[..., :Ty], E.attn_f1(xs, yb, B, Tx, Ty, H, y_bcast=True)[..., :Ty], 1e-4, "f1 bcast")
    p = torch.zeros(H, B, Tx, E.round8(Ty))
    p[..., :Ty] = torch.softmax(want[..., :Ty] * 0.05, -1)
    p = p.to(BF)
    pc = p.cuda()
    pc[..., Ty:] = float("nan")  # padding columns must never be read
    close(K().attn_f2(pc, yc[:, d:], B, Tx, Ty, H), E.attn_f2(p, ys, B, Tx, Ty, H), 1e-2, "f2")
    close(K().attn_f2(pc, yb.cuda(), B, Tx, Ty, H, y_bcast=True), E.attn_f2(p, yb, B, Tx, Ty, H, y_bcast=True), 1e-2, "f2 bcast")
    out = torch.zeros(B * Ty, 3 * d, device="cuda", dtype=BF)
    K().attn_f3(pc, xc[:, d:2 * d], B, Tx, Ty, H, out=out[:, d:2 * d])
    close(out[:, d:2 * d], E.attn_f3(p, xs, B, Tx, Ty, H), 1e-2, "f3")
    assert out[:, :d].abs().max().item() == 0 and out[:, 2 * d:].abs().max().item() == 0
    if Tx == Ty:
        acc = torch.zeros(Ty, d, device="cuda")
        pz = pc.clone()
        pz[..., Ty:] = 0  # f4 reads [B*T, ld] rows as one matrix: padding columns are outputs >= P, discarded
        K().attn_f4(pz, xc[:, d:2 * d], B, Tx, Ty, H, acc)
        close(acc, E.attn_f4(p, xs, B, Tx, Ty, H, torch.zeros(Ty, d)), 1e-3, "f4")


# ---------------------------------------------------------------------------------------------- LayerNorm
@pytest.mark.parametrize("M,D", [(1000, 512), (37, 16), (513, 256), (64, 2048)])
def test_layernorm(M, D):
    x, ga, be = rnd(M, D, seed=1) * 3 + 1, rnd(D, seed=2) + 1, rnd(D, seed=3)
    y16, y32, st = K().layernorm_fwd(x.cuda(), ga.cuda(), be.cuda(), want_bf16=True, want_f32=True)
    _, w32, wst = E.layernorm_fwd(x, ga, be, want_bf16=False, want_f32=True)
    close(y32, w32, 1e-5, "ln y32")
    close(y16, w32, 1e-2, "ln y16")
    close(st, wst, 1e-5, "ln stats")
    for dy in (rnd(M, D, seed=4), rnd(M, D, seed=4).to(BF)):
        add = rnd(M, D, seed=5)
        dg, db = torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda")
        dx, g16 = K().layernorm_bwd(dy.cuda(), x.cuda(), ga.cuda(), st, dg, db, dx_add=add.cuda(), want_g=True, g_alpha=0.5)
        wg, wb = torch.zeros(D), torch.zeros(D)
        wdx, wg16 = E.layernorm_bwd(dy, x, ga, wst, wg, wb, dx_add=add, want_g=True, g_alpha=0.5)
        close(dx, wdx, 2e-5, "ln dx")
        close(g16, wg16, 1e-2, "ln g16")
        close(dg, wg, 1e-4, "ln dgamma")
        close(db, wb, 1e-4, "ln dbeta")
        # fused branch head: g16 with dropout + its column sums (accumulated) must equal the separate cast_colsum pass
        drop = (0.1, 77, 5)
        cs, dg2, db2 = torch.ones(D, device="cuda"), torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda")
        dx2, g2 = K().layernorm_bwd(dy.cuda(), x.cuda(), ga.cuda(), st, dg2, db2, dx_add=add.cuda(), want_g=True, g_alpha=0.5,
                                    g_drop=drop, g_colsum_out=cs)
        assert torch.equal(dx2, dx)
        cs_ref = torch.ones(D, device="cuda")
        g_ref = K().cast_bf16(dx, alpha=0.5, drop=drop, colsum_out=cs_ref)
        assert torch.equal(g2, g_ref)
        close(cs, cs_ref, 1e-4, "ln g colsum")
        close(dg2, wg, 1e-4, "ln dgamma (fused head)")


# ---------------------------------------------------------------------------------------------- small helpers
def test_cast_colsum_posbias_embed():
    x = rnd(1003, seed=1)
    close(K().cast_bf16(x.cuda(), alpha=0.5), x * 0.5, 1e-2, "cast")
    dy = rnd(777, 520, seed=2).to(BF)
    acc = torch.ones(512, device="cuda")
    K().colsum(dy.cuda(), acc, n=512)
    close(acc, 1 + dy.float()[:, :512].sum(0), 1e-4, "colsum")
    acc2 = torch.zeros(20, device="cuda")
    K().colsum(dy.cuda()[:, :24], acc2, n=20)
    close(acc2, dy.float()[:, :20].sum(0), 1e-4, "colsum ragged")
    # fused cast + column sums (branch gradient + bias gradient in one pass), with and without dropout
    xg = rnd(1237, 520, seed=6)
    accf = torch.full((520,), 2.0, device="cuda")
    yg = K().cast_bf16(xg.cuda(), alpha=0.5, colsum_out=accf)
    close(yg, xg * 0.5, 1e-2, "cast_colsum y")
    close(accf, 2 + (xg * 0.5).sum(0), 1e-4, "cast_colsum sums")
    accd = torch.zeros(520, device="cuda")
    yd = K().cast_bf16(xg.cuda(), alpha=1.0, drop=(0.2, 77, 5), colsum_out=accd)
    yd2 = K().cast_bf16(xg.cuda(), alpha=1.0, drop=(0.2, 77, 5))     # same hash stream as the plain cast
    assert torch.equal(yd, yd2)
    close(accd, yd.float().sum(0), 2e-2, "cast_colsum dropout sums")
    M, d = 333, 64
    qkv = rnd(M, 3 * d, seed=3).to(BF)
    u, v = rnd(d, seed=4), rnd(d, seed=5)
    qu, qv = K().pos_bias_fwd(qkv.cuda()[:, :d], u.cuda(), v.cuda())
    wu, wv = E.pos_bias_fwd(qkv[:, :d], u, v)
    close(qu, wu, 1e-2, "qu")
    close(qv, wv, 1e-2, "qv")
    dq = torch.zeros(M, 3 * d, device="cuda", dtype=BF)
    du, dv = torch.zeros(d, device="cuda"), torch.zeros(d, device="cuda")
    a, b = rnd(M, d, seed=6).to(BF), rnd(M, d, seed=7).to(BF)
    K().pos_bias_bwd(a.cuda(), b.cuda(), dq[:, :d], du, dv)
    close(dq[:, :d], a.float() + b.float(), 1e-2, "dq")
    close(du, a.float().sum(0), 1e-4, "du")
    close(dv, b.float().sum(0), 1e-4, "dv")
    assert dq[:, d:].abs().max().item() == 0
    V, L, R = 50, 7, 28
    emb, pe = rnd(V, d, seed=8), rnd(L, d, seed=9)
    tok = torch.randint(0, V, (R,), generator=torch.Generator().manual_seed(0)).to(torch.int32)
    close(K().embed_fwd(emb.cuda(), tok.cuda(), pe.cuda(), L, 8.0), E.embed_fwd(emb, tok, pe, L, 8.0), 1e-6, "embed")
    dx = rnd(R, d, seed=10)
    de = torch.zeros(V, d, device="cuda")
    K().embed_bwd(dx.cuda(), tok.cuda(), 8.0, de)
    we = torch.zeros(V, d)
    E.embed_bwd(dx, tok, 8.0, we)
    close(de, we, 1e-5, "demb")


# ---------------------------------------------------------------------------------------------- softmax + rel_shift
@pytest.mark.parametrize("mode", ["latest", "legacy", "none"])
@pytest.mark.parametrize("T", [7, 33, 124, 300])
def test_softmax_relshift(mode, T):
    H, B = 2, 3
    ld = E.round8(T)
    Pn = 2 * T - 1 if mode == "latest" else T
    ac = torch.full((H, B, T, ld), float("nan"))
    ac[..., :T] = rnd(H, B, T, T, seed=1) * 3
    bd = None
    if mode != "none":
        bd = torch.full((H, B, T, E.round8(Pn)), float("nan"))
        bd[..., :Pn] = rnd(H, B, T, Pn, seed=2) * 3
    lens = torch.tensor([T, max(1, T - 3), max(1, T // 2)], dtype=torch.int32)
    causal = mode == "none"
    p, pd = K().softmax_fwd(ac.cuda(), bd.cuda() if bd is not None else None, lens.cuda(), T, mode, causal, 0.125)
    ac0, bd0 = torch.nan_to_num(ac), (torch.nan_to_num(bd) if bd is not None else None)
    wp, _ = E.softmax_fwd(ac0, bd0, lens, T, mode, causal, 0.125)
    close(p, wp, 1e-2, "softmax p")
    assert p.data_ptr() == pd.data_ptr()
    assert p[..., T:].abs().max().item() == 0 if ld > T else True
    dpd = torch.full((H, B, T, ld), float("nan"))
    dpd[..., :T] = rnd(H, B, T, T, seed=3)
    dac, dbd = K().softmax_bwd(dpd.cuda(), p, T, mode, 0.125, ld_bd=E.round8(Pn))
    wdac, wdbd = E.softmax_bwd(torch.nan_to_num(dpd), p.cpu(), T, mode, 0.125, ld_bd=E.round8(Pn))
    close(dac, wdac, 1e-2, "softmax dac")
    if mode != "none":
        close(dbd, wdbd, 1e-2, "softmax dbd")


@pytest.mark.parametrize("mode,causal,Tq,Tk", [("latest", False, 155, 155), ("latest", False, 124, 124), ("latest", False, 7, 7),
                                               ("latest", False, 192, 192), ("latest", False, 129, 129), ("latest", False, 33, 33),
                                               ("none", True, 21, 21), ("none", False, 21, 155), ("none", False, 140, 99),
                                               ("none", True, 150, 150)])
def test_fused_attention_fwd(mode, causal, Tq, Tk):
    """the fused tcgen05 kernel against the statement of the unfused chain, and against the unfused CUDA chain"""
    H, B, dk = 4, 3, 64
    d = H * dk
    assert K().relattn_supported(Tq, Tk, dk, mode)
    qkv = (rnd(B * Tq, 3 * d, seed=1) * 0.7).to(BF)
    kv = (rnd(B * Tk, 2 * d, seed=2) * 0.7).to(BF)
    qv = (rnd(B * Tq, d, seed=3) * 0.7).to(BF)
    Pn = 2 * Tq - 1
    pos = (rnd(Pn, d, seed=4) * 0.7).to(BF) if mode == "latest" else None
    lens = torch.tensor([Tk, max(1, Tk - 3), max(1, Tk // 2)], dtype=torch.int32)
    qu = qkv[:, :d]
    k, v = kv[:, :d], kv[:, d:]
    qkv_c, kv_c = qkv.cuda(), kv.cuda()
    args_c = (qkv_c[:, :d].contiguous() if mode == "latest" else qkv_c[:, :d], qv.cuda() if mode == "latest" else None, kv_c[:, :d], kv_c[:, d:],
              pos.cuda() if pos is not None else None, lens.cuda(), B, Tq, Tk, H, mode, causal, 0.125)
    p, pd, ctx = K().relattn_fwd(*args_c)
    wp, _, wctx = E.relattn_fwd(qu, qv if mode == "latest" else None, k, v, pos, lens, B, Tq, Tk, H, mode, causal, 0.125)
    assert p.data_ptr() == pd.data_ptr()
    close(p[..., :Tk], wp[..., :Tk], 1e-2, "fused p")
    close(ctx, wctx, 1e-2, "fused ctx")
    for bi in range(B):  # masked keys get probability exactly 0
        assert p[:, bi, :, int(lens[bi]):Tk].abs().max().item() == 0 if int(lens[bi]) < Tk else True
    # dropout: the decisions are those of the unfused softmax kernel (the unfused backward regenerates them)
    drop = (0.1, 1234, 7)
    p2, pd2, ctx2 = K().relattn_fwd(*args_c, drop=drop)
    ac = K().attn_f1(args_c[0], args_c[2], B, Tq, Tk, H)
    bd = K().attn_f1(args_c[1], args_c[4], B, Tq, Pn, H, y_bcast=True) if mode == "latest" else None
    up, upd = K().softmax_fwd(ac, bd, args_c[5], Tk, mode, causal, 0.125, drop=drop)
    close(p2[..., :Tk], up[..., :Tk], 1e-2, "fused p vs unfused")
    assert ((pd2[..., :Tk] == 0) == (upd[..., :Tk] == 0)).float().mean().item() > 0.999
    close(pd2[..., :Tk], upd[..., :Tk], 2e-2, "fused p_drop vs unfused")
    close(ctx2, K().attn_f2(upd, args_c[3], B, Tq, Tk, H), 2e-2, "fused ctx (dropout) vs unfused")


@pytest.mark.parametrize("drop", [None, (0.1, 99, 5)])
@pytest.mark.parametrize("mode,Tq,Tk", [("latest", 300, 300), ("latest", 193, 193), ("latest", 468, 468), ("none", 37, 430),
                                        ("none", 130, 385), ("latest", 384, 384), ("latest", 155, 155), ("none", 21, 99)])
def test_fused_attention_key_chunks(mode, Tq, Tk, drop, monkeypatch):
    """More than 192 keys: one launch of the fused forward / backward per 192-key chunk + the combine pass, against the
    statement of the unfused chain over the whole key axis (same dropout decisions: the chunk offset enters the counter)."""
    H, B, dk = 2, 3, 64
    d = H * dk
    rel = mode == "latest"
    monkeypatch.setattr(K(), "TAIL_MIN_UNITS", 0 if drop is None else 10 ** 6)  # both forms of the backward's last chunk
    assert K().relattn_supported(Tq, Tk, dk, mode) and K().relattn_bwd_supported(Tq, Tk, dk, mode)
    qkv = (rnd(B * Tq, 3 * d, seed=1) * 0.7).to(BF)
    kv = (rnd(B * Tk, 2 * d, seed=2) * 0.7).to(BF)
    qv = (rnd(B * Tq, d, seed=3) * 0.7).to(BF)
    Pn = 2 * Tq - 1
    pos = (rnd(Pn, d, seed=4) * 0.7).to(BF) if rel else None
    lens = torch.tensor([Tk, max(1, Tk - 200), 150], dtype=torch.int32)   # whole chunks masked for the shorter utterances
    dctx = rnd(B * Tq, d, seed=5).to(BF)
    qkv_c, kv_c = qkv.cuda(), kv.cuda()
    qu_c = qkv_c[:, :d].contiguous() if rel else qkv_c[:, :d]
    p, _, ctx = K().relattn_fwd(qu_c, qv.cuda() if rel else None, kv_c[:, :d], kv_c[:, d:], pos.cuda() if rel else None, lens.cuda(),
                                B, Tq, Tk, H, mode, False, 0.125, drop=drop, want_pd=False)
    wp, wpd, wctx = E.relattn_fwd(qkv[:, :d], qv if rel else None, kv[:, :d], kv[:, d:], pos, lens, B, Tq, Tk, H, mode, False, 0.125,
                                  drop=drop)
    close(p[..., :Tk], wp[..., :Tk], 1.5e-2, "p over key chunks")
    close(ctx, wctx, 2e-2, "ctx over key chunks")
    assert torch.isfinite(p.float()).all() and (p.shape[-1] == Tk or p.float()[..., Tk:].abs().max().item() == 0)
    for bi in range(B):
        if int(lens[bi]) < Tk:
            assert p[:, bi, :, int(lens[bi]):Tk].abs().max().item() == 0
    dq = torch.full((B * Tq, 3 * d), 7.0, device="cuda", dtype=BF)
    dkv = torch.full((B * Tk, 2 * d), 7.0, device="cuda", dtype=BF)
    dqv = torch.full((B * Tq, 3 * d), 7.0, device="cuda", dtype=BF) if rel else None
    dpos = torch.ones(Pn, d, device="cuda") if rel else None
    cks, cvs = torch.ones(d, device="cuda"), torch.ones(d, device="cuda")   # bias gradients of the K / V projections, accumulated
    K().relattn_bwd(dctx.cuda(), qu_c, kv_c[:, :d], kv_c[:, d:], p, ctx, B, Tq, Tk, H, mode, 0.125, dq[:, d:2 * d], dkv[:, :d], dkv[:, d:],
                    drop=drop, qv=qv.cuda() if rel else None, pos=pos.cuda() if rel else None, dqv=dqv[:, d:2 * d] if rel else None,
                    dpos=dpos, dk_colsum=cks, dv_colsum=cvs)
    torch.cuda.synchronize()
    wdq, wdk, wdv = torch.zeros(B * Tq, d), torch.zeros(B * Tk, d), torch.zeros(B * Tk, d)
    wdqv, wdpos = (torch.zeros(B * Tq, d), torch.ones(Pn, d)) if rel else (None, None)
    E.relattn_bwd(dctx, qkv[:, :d], kv[:, :d], kv[:, d:], p.float().cpu(), ctx.float().cpu(), B, Tq, Tk, H, mode, 0.125, wdq, wdk, wdv,
                  drop=drop, qv=qv if rel else None, pos=pos, dqv=wdqv, dpos=wdpos)
    close(dq[:, d:2 * d], wdq, 2e-2, "dq_u")
    close(dkv[:, :d], wdk, 2e-2, "dk")
    close(dkv[:, d:], wdv, 2e-2, "dv")
    assert (dq[:, :d] == 7.0).all() and (dq[:, 2 * d:] == 7.0).all()
    close(cks, 1.0 + wdk.sum(0), 2e-2, "colsum(dk)")
    close(cvs, 1.0 + wdv.sum(0), 2e-2, "colsum(dv)")
    if rel:
        close(dqv[:, d:2 * d], wdqv, 2e-2, "dq_v")
        close(dpos, wdpos, 2e-2, "d(projected positions)")


@pytest.mark.parametrize("drop", [None, (0.1, 4321, 9)])
@pytest.mark.parametrize("mode,Tq,Tk", [("latest", 124, 124), ("latest", 155, 155), ("latest", 192, 192), ("latest", 129, 129),
                                        ("latest", 33, 33), ("latest", 64, 64), ("none", 21, 21), ("none", 21, 155),
                                        ("none", 140, 99), ("none", 5, 190), ("none", 150, 150)])
def test_fused_attention_bwd(mode, Tq, Tk, drop):
    """the fused tcgen05 backward (dP, dS, dropout(p) on the SM) against the statement of the unfused chain, fed the
    probabilities and context the fused forward produced"""
    H, B, dk = 4, 3, 64
    d = H * dk
    if not K().relattn_bwd_supported(Tq, Tk, dk, mode):
        pytest.skip("shape not on the fused backward")
    qkv = (rnd(B * Tq, 3 * d, seed=1) * 0.7).to(BF)
    kv = (rnd(B * Tk, 2 * d, seed=2) * 0.7).to(BF)
    qv = (rnd(B * Tq, d, seed=3) * 0.7).to(BF)
    pos = (rnd(2 * Tq - 1, d, seed=4) * 0.7).to(BF) if mode == "latest" else None
    lens = torch.tensor([Tk, max(1, Tk - 3), max(1, Tk // 2)], dtype=torch.int32)
    dctx = (rnd(B * Tq, d, seed=5)).to(BF)
    qkv_c, kv_c = qkv.cuda(), kv.cuda()
    qu_c = qkv_c[:, :d].contiguous() if mode == "latest" else qkv_c[:, :d]
    p, _, ctx = K().relattn_fwd(qu_c, qv.cuda() if mode == "latest" else None, kv_c[:, :d], kv_c[:, d:],
                                pos.cuda() if pos is not None else None, lens.cuda(), B, Tq, Tk, H, mode, False, 0.125, drop=drop,
                                want_pd=False)
    dq = torch.full((B * Tq, 3 * d), 7.0, device="cuda", dtype=BF)   # written through strided views, like dqkv in the step
    dkv = torch.full((B * Tk, 2 * d), 7.0, device="cuda", dtype=BF)
    rel = mode == "latest"
    Pn = 2 * Tq - 1
    dqv = torch.full((B * Tq, 3 * d), 7.0, device="cuda", dtype=BF) if rel else None
    dpos = torch.ones(Pn, d, device="cuda") if rel else None   # accumulated into
    dac = K().relattn_bwd(dctx.cuda(), qu_c, kv_c[:, :d], kv_c[:, d:], p, ctx, B, Tq, Tk, H, mode, 0.125, dq[:, d:2 * d], dkv[:, :d],
                          dkv[:, d:], drop=drop, want_dac=True, qv=qv.cuda() if rel else None, pos=pos.cuda() if rel else None,
                          dqv=dqv[:, d:2 * d] if rel else None, dpos=dpos)
    torch.cuda.synchronize()
    wdq, wdk, wdv = torch.zeros(B * Tq, d), torch.zeros(B * Tk, d), torch.zeros(B * Tk, d)
    wdqv, wdpos = (torch.zeros(B * Tq, d), torch.ones(Pn, d)) if rel else (None, None)
    wdac = E.relattn_bwd(dctx, qkv[:, :d], kv[:, :d], kv[:, d:], p.float().cpu(), ctx.float().cpu(), B, Tq, Tk, H, mode, 0.125, wdq,
                         wdk, wdv, drop=drop, want_dac=True, qv=qv if rel else None, pos=pos, dqv=wdqv, dpos=wdpos)
    close(dac[..., :Tk], wdac[..., :Tk], 2e-2, "dS")
    close(dq[:, d:2 * d], wdq, 2e-2, "dq_u")
    close(dkv[:, :d], wdk, 2e-2, "dk")
    close(dkv[:, d:], wdv, 2e-2, "dv")
    assert (dq[:, :d] == 7.0).all() and (dq[:, 2 * d:] == 7.0).all()  # neighbours of the strided outputs untouched
    if rel:
        close(dqv[:, d:2 * d], wdqv, 2e-2, "dq_v")
        close(dpos, wdpos, 2e-2, "d(projected positions)")
        assert (dqv[:, :d] == 7.0).all() and (dqv[:, 2 * d:] == 7.0).all()
        dbd = K().rel_unshift(dac, Tq, mode, E.round8(Pn))
        wdbd = E.rel_unshift(dac.float().cpu(), Tq, mode, E.round8(Pn))
        assert torch.equal(dbd.float().cpu()[..., :Pn], wdbd[..., :Pn])


@pytest.mark.parametrize("T", [1, 2, 3, 7, 16, 33, 64])
def test_rel_shift_index_bit_exact(T, golden_dir):
    import os
    z = np.load(os.path.join(golden_dir, "rel_shift.npz"))
    for mode in ("latest", "legacy"):
        code = z[f"{mode}_{T}"]  # reference: row*10000 + col + 1, 0 = literal zero
        for i in range(T):
            for j in range(T):
                v = K().rel_shift_index(i, j, T, mode)
                mine = 0 if v == 0 else (v >> 16) * 10000 + (v & 0xFFFF)
                assert mine == code[i, j], (mode, T, i, j)


# ---------------------------------------------------------------------------------------------- conv module
@pytest.mark.parametrize("B,T,C,Kk", [(3, 50, 16, 7), (4, 124, 512, 31), (2, 31, 64, 15)])
def test_conv_module_middle(B, T, C, Kk):
    pw1 = rnd(B * T, 2 * C, seed=1).to(BF)
    w, bias = rnd(C, Kk, seed=2) * 0.3, rnd(C, seed=3)
    out, sums = K().glu_dwconv_fwd(pw1.cuda(), w.cuda(), bias.cuda(), B, T, Kk)
    wout, wsums = E.glu_dwconv_fwd(pw1, w, bias, B, T, Kk)
    close(out, wout, 1e-2, "dw out")
    close(sums, wsums, 1e-3, "bn sums")
    rm, rv, nb = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda"), torch.zeros(1, device="cuda", dtype=torch.long)
    mr = K().bn_finalize(sums, B * T, 1e-5, 0.1, rm, rv, nb)
    wrm, wrv, wnb = torch.zeros(C), torch.ones(C), torch.zeros(1, dtype=torch.long)
    wmr = E.bn_finalize(wsums, B * T, 1e-5, 0.1, wrm, wrv, wnb)
    close(mr, wmr, 1e-3, "mean_rstd")
    close(rm, wrm, 1e-3, "running_mean")
    close(rv, wrv, 1e-3, "running_var")
    assert nb.item() == 1
    ga, be = rnd(C, seed=4) + 1, rnd(C, seed=5)
    y = K().bn_swish_fwd(out, mr, ga.cuda(), be.cuda())
    close(y, E.bn_swish_fwd(out.cpu(), mr.cpu(), ga, be), 1e-2, "bn_swish")
    dy = rnd(B * T, C, seed=6).to(BF)
    dg, db = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    dx = K().bn_swish_bwd(dy.cuda(), out, mr, ga.cuda(), be.cuda(), dg, db)
    wg, wb = torch.zeros(C), torch.zeros(C)
    wdx = E.bn_swish_bwd(dy, out.cpu(), mr.cpu(), ga, be, wg, wb)
    close(dx, wdx, 1e-2, "bn dx")
    close(dg, wg, 1e-3, "bn dgamma")
    close(db, wb, 1e-3, "bn dbeta")
    dwg, dbg = torch.zeros(C, Kk, device="cuda"), torch.zeros(C, device="cuda")
    dps = torch.ones(2 * C, device="cuda")   # pointwise_conv1's bias gradient, accumulated
    dpw1 = K().glu_dwconv_bwd(dy.cuda(), pw1.cuda(), w.cuda(), B, T, Kk, dwg, dbg, dpw1_colsum=dps)
    wdw, wdb, wps = torch.zeros(C, Kk), torch.zeros(C), torch.ones(2 * C)
    wdpw1 = E.glu_dwconv_bwd(dy, pw1, w, B, T, Kk, wdw, wdb, dpw1_colsum=wps)
    close(dpw1, wdpw1, 1e-2, "d_pw1")
    close(dps, wps, 1e-3, "colsum(d_pw1)")
    close(dwg, wdw, 1e-3, "d_dw_weight")
    close(dbg, wdb, 1e-3, "d_dw_bias")


# ---------------------------------------------------------------------------------------------- subsampling
@pytest.mark.parametrize("B,T1,F1,C", [(2, 23, 39, 256), (3, 24, 40, 256), (1, 7, 5, 256), (2, 3, 3, 256), (4, 50, 39, 512),
                                       (5, 312, 39, 256)])
def test_conv2_implicit_gemms_bit_exact(B, T1, F1, C):
    """conv.2 forward / wgrad / dgrad as implicit GEMMs (TMA im2col operand loads, parity-class gather for dgrad) on
    small-integer data, where every sum is exact in fp32 and representable in bf16: the results must EQUAL the explicit
    im2col statement -- any wrong pixel, tap, channel or parity class shows as a changed integer."""
    g = torch.Generator().manual_seed(B * 1000 + T1 * 10 + F1 + C)
    T2, F2 = (T1 - 3) // 2 + 1, (F1 - 3) // 2 + 1
    def sparse(shape, vals, p):
        idx = torch.randint(0, len(vals), shape, generator=g)
        keep = torch.rand(shape, generator=g) < p
        return (torch.tensor(vals, dtype=torch.float32)[idx] * keep).to(BF)
    h1 = sparse((B, T1, F1, C), [1.0, 2.0], 0.15)
    w = sparse((C, 9 * C), [-1.0, 1.0], 0.3)
    bias = torch.randint(-3, 4, (C,), generator=g).float()
    dh2 = sparse((B * T2 * F2, C), [-1.0, 1.0, 2.0], 0.1)
    col = E.im2col(h1).float()
    want_h2 = torch.relu(col @ w.float().t() + bias)
    assert want_h2.max() < 256
    h2 = K().conv2_fwd(h1.cuda(), w.cuda(), bias.cuda())
    assert torch.equal(h2.float().cpu(), want_h2), ("fwd", (h2.float().cpu() - want_h2).abs().max())
    dw = torch.full((C, 9 * C), 3.0, device="cuda")
    K().conv2_wgrad(dh2.cuda(), h1.cuda(), dw)
    want_dw = 3.0 + dh2.float().t() @ col
    assert torch.equal(dw.cpu(), want_dw), ("wgrad", (dw.cpu() - want_dw).abs().max())
    want_dh1 = E.conv2_dgrad_relu(dh2, w, h1).float()
    assert want_dh1.abs().max() < 256
    dh1 = K().conv2_dgrad_relu(dh2.cuda(), w.cuda(), h1.cuda())
    assert torch.equal(dh1.float().cpu(), want_dh1), ("dgrad", (dh1.float().cpu() - want_dh1).abs().max())


@pytest.mark.parametrize("B,Tin,F,C", [(2, 47, 80, 16), (3, 101, 80, 512), (2, 30, 20, 64)])
def test_subsampling(B, Tin, F, C):
    x, w, b = rnd(B, Tin, F, seed=1), rnd(C, 9, seed=2) * 0.3, rnd(C, seed=3) * 0.1
    h1 = K().conv1_fwd(x.cuda(), w.cuda(), b.cuda())
    close(h1, E.conv1_fwd(x, w, b), 1e-2, "conv1")
    if C % 256 == 0:  # the (h1 > 0) bit rows conv.2's dgrad reads: next to h1 from the same kernel, or derived from h1
        h1b, bits = K().conv1_fwd(x.cuda(), w.cuda(), b.cuda(), want_bits=True)
        assert torch.equal(h1b, h1)
        assert torch.equal(bits.cpu(), E.relu_bits(h1.cpu()))
        assert torch.equal(K().relu_bits(h1), bits)
    col = K().im2col(h1)
    assert torch.equal(col.cpu(), E.im2col(h1.cpu()))
    dcol = rnd(*col.shape, seed=4).to(BF)
    dh1 = K().col2im_relu(dcol.cuda(), h1)
    close(dh1, E.col2im_relu(dcol, h1.cpu()), 1e-2, "col2im")
    dw, db = torch.zeros(C, 9, device="cuda"), torch.zeros(C, device="cuda")
    K().conv1_bwd(dh1, x.cuda(), dw, db)
    wdw, wdb = torch.zeros(C, 9), torch.zeros(C)
    E.conv1_bwd(dh1.cpu(), x, wdw, wdb)
    close(dw, wdw, 1e-3, "conv1 dw")
    close(db, wdb, 1e-3, "conv1 db")


def test_vocab_softmax_fwd_bwd():
    """Posterior for self-conditioned intermediate CTC and its backward, against the emulation (V = 600 and a ragged V)."""
    for R, V in ((777, 600), (65, 24), (33, 50)):
        ld = ((V + 7) // 8) * 8
        x = rnd(R, ld, seed=R) * 3
        p = K().vocab_softmax_fwd(x.cuda(), V)
        want = E.vocab_softmax_fwd(x, V)
        assert p.shape == (R, ld) and float(p[:, V:].float().abs().sum()) == 0.0
        close(p, want, 1e-2, "vocab softmax")
        dp, add = rnd(R, ld, seed=R + 1).to(BF), rnd(R, ld, seed=R + 2).to(BF)
        for a in (None, add):
            d = K().vocab_softmax_bwd(p, dp.cuda(), V, add=a.cuda() if a is not None else None)
            wd = E.vocab_softmax_bwd(p.cpu(), dp, V, add=a)
            assert float(d[:, V:].float().abs().sum()) == 0.0
            close(d, wd, 2e-2, "vocab softmax bwd")
