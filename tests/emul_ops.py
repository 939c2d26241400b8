"""Plain-torch statement of what every function of ``espnet_slurp_b200.ops`` computes (TEST CODE ONLY).

Two uses: (1) on the GPU each CUDA kernel is compared with its function here on the same inputs;
(2) on the CPU the step orchestration (``espnet_slurp_b200.step``: forward + hand-written backward) runs
over these functions instead of the kernels and is compared with autograd over ``oracle/model.py``, which
checks every backward formula and every buffer layout without a GPU.  Dropout is emulated with the library's own
counter-based decisions (tests/dropout_ref.py restates the hash; element index conventions as in the kernels).
``exact=True`` keeps everything in fp32 (no bf16 rounding) for tight checks of the formulas.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

exact = False
BF = torch.bfloat16
F32 = torch.float32


def _r(x):  # bf16 storage rounding
    return x.float() if exact else x.to(BF)


def _lowp():
    return F32 if exact else BF


def round8(n):
    return (n + 7) // 8 * 8


def _mul(drop, shape):
    """keep / (1 - p) multipliers of a dropout site over a row-major tensor of ``shape`` (None when inactive)."""
    if drop is None or drop[0] <= 0.0:
        return None
    import dropout_ref
    return dropout_ref.multiplier(drop, tuple(shape))


def _act(y, act):
    if act == "swish":
        return y * torch.sigmoid(y)
    if act == "relu":
        return torch.relu(y)
    return y


def _act_grad(x, act):
    if act == "mul":
        return x
    if act == "none" or act is None:
        return torch.ones_like(x)
    if act == "swish":
        s = torch.sigmoid(x)
        return s * (1 + x * (1 - s))
    return (x > 0).float()


def gemm(a, b, *, trans=False, trans_a=None, trans_b=None, bias=None, act=None, alpha=1.0, residual=None,
         act_grad_input=None, act_grad=None, act_grad_out=False, out_dtype=None, out=None, accumulate=False, drop=None,
         dropout_p=0.0, seed=0, site=0, colsum_out=None):
    if drop is None and dropout_p > 0.0:
        drop = (dropout_p, seed, site)
    ta = trans if trans_a is None else trans_a
    tb = trans if trans_b is None else trans_b
    A = a.float().t() if ta else a.float()
    Bm = b.float() if tb else b.float().t()
    y = A @ Bm
    N = y.shape[1]
    if bias is not None:
        y = y + bias[:N]
    pre_act_out = act_grad_out
    m = _mul(drop, y.shape)  # element index row * N + column (gemm_sm100.cu epilogue)
    pre = None
    if pre_act_out is not False and pre_act_out is not None:
        pre = _act_grad(y, act)
        pre = _r(pre * m if m is not None else pre)  # the saved derivative carries the element's dropout multiplier
    y = _act(y, act)
    if m is not None:
        y = y * m
    y = y * alpha
    if act_grad_input is not None:
        y = y * _act_grad(act_grad_input.float(), act_grad)
    if residual is not None:
        y = y + residual[:, :N]
    if out is not None:
        if accumulate:
            out[:, :N] += y.to(out.dtype)
        else:
            out[:, :N] = y.to(out.dtype)
        res = out
    else:
        dt = out_dtype if out_dtype is not None else _lowp()
        if dt == BF and exact:
            dt = F32
        res = y.to(dt)
    if colsum_out is not None:  # of the output as stored
        colsum_out += res.float()[:, :N].sum(0)
    return (res, pre) if pre_act_out is True else res


def attn_f1(x, y, B, Tx, Ty, H, y_bcast=False):
    dk = x.shape[1] // H
    xs = x.float().reshape(B, Tx, H, dk)
    ys = y.float().reshape(1 if y_bcast else B, Ty, H, dk)
    out = torch.zeros(H, B, Tx, round8(Ty))
    out[..., :Ty] = torch.einsum("bihc,bjhc->hbij", xs, ys.expand(B, Ty, H, dk))
    return out


def attn_f2(s, y, B, Tq, Tk, H, out=None, y_bcast=False):
    dk = y.shape[1] // H
    ys = y.float().reshape(1 if y_bcast else B, Tk, H, dk).expand(B, Tk, H, dk)
    r = torch.einsum("hbij,bjhc->bihc", s.float()[..., :Tk], ys).reshape(B * Tq, H * dk)
    if out is None:
        return _r(r)
    out.copy_(r.to(out.dtype))
    return out


def attn_f3(s, x, B, Tq, Tk, H, out=None):
    dk = x.shape[1] // H
    xs = x.float().reshape(B, Tq, H, dk)
    r = torch.einsum("hbij,bihc->bjhc", s.float()[..., :Tk], xs).reshape(B * Tk, H * dk)
    if out is None:
        return _r(r)
    out.copy_(r.to(out.dtype))
    return out


def attn_f4(s, x, B, T, P, H, out):
    dk = x.shape[1] // H
    xs = x.float().reshape(B, T, H, dk)
    out += torch.einsum("hbin,bihc->nhc", s.float()[..., :P], xs).reshape(P, H * dk)
    return out


def layernorm_fwd(x, gamma, beta, eps=1e-12, want_bf16=True, want_f32=False):
    mean = x.mean(-1, keepdim=True)
    var = ((x - mean) ** 2).mean(-1, keepdim=True)
    rstd = torch.rsqrt(var + eps)
    y = (x - mean) * rstd * gamma + beta
    return (_r(y) if want_bf16 else None), (y if want_f32 else None), torch.cat([mean, rstd], 1)


def layernorm_bwd(dy, x, gamma, stats, dgamma, dbeta, dx_add=None, want_g=False, g_alpha=1.0, g_drop=None, g_colsum_out=None):
    dy = dy.float()
    mean, rstd = stats[:, 0:1], stats[:, 1:2]
    xh = (x - mean) * rstd
    dyg = dy * gamma
    dx = rstd * (dyg - dyg.mean(-1, keepdim=True) - xh * (dyg * xh).mean(-1, keepdim=True))
    if dx_add is not None:
        dx = dx + dx_add
    dgamma += (dy * xh).sum(0)
    dbeta += dy.sum(0)
    m = _mul(g_drop, dx.shape)
    gd = dx * g_alpha if m is None else dx * m * g_alpha
    if g_colsum_out is not None:
        g_colsum_out += gd.float().sum(0)
    return dx, (_r(gd) if want_g else None)


def cast_bf16(x, alpha=1.0, drop=None, out=None, colsum_out=None):
    m = _mul(drop, x.shape)
    xd = x * alpha if m is None else x * m * alpha
    y = _r(xd)
    if colsum_out is not None:
        colsum_out += xd.float().sum(0)
    if out is not None:
        out.copy_(y.to(out.dtype))
        return out
    return y


def colsum(dy, out_accum, n=None):
    N = dy.shape[1] if n is None else n
    out_accum[:N] += dy.float()[:, :N].sum(0)
    return out_accum


def pos_bias_fwd(q, u, v):
    return _r(q.float() + u), _r(q.float() + v)


def pos_bias_bwd(dqu, dqv, dq_out, du, dv, dq_colsum=None):
    dq_out.copy_((dqu.float() + dqv.float()).to(dq_out.dtype))
    du += dqu.float().sum(0)
    dv += dqv.float().sum(0)
    if dq_colsum is not None:
        dq_colsum += dqu.float().sum(0) + dqv.float().sum(0)


def embed_fwd(emb, tokens, pe, L, scale, drop=None):
    R = tokens.numel()
    pos = torch.arange(R) % L
    x = emb[tokens.long()] * scale + pe[pos]
    m = _mul(drop, x.shape)
    return x if m is None else x * m


def embed_bwd(dx, tokens, scale, demb, drop=None):
    m = _mul(drop, dx.shape)
    demb.index_add_(0, tokens.long(), (dx if m is None else dx * m) * scale)


def _shift(bd, T, mode):
    """un-shifted bd [H,B,T,ld] -> shifted [H,B,T,T]"""
    i = torch.arange(T)[:, None]
    j = torch.arange(T)[None, :]
    if mode == "latest":
        return torch.gather(bd, 3, (T - 1 - i + j).expand(bd.shape[0], bd.shape[1], T, T))
    row = torch.where(j <= i, i, (i + 1).clamp(max=T - 1)).expand(T, T)
    col = torch.where(j <= i, T - 1 - i + j, (j - i - 2).clamp(min=0))
    out = bd[:, :, row, col]
    return out.masked_fill((j == i + 1)[None, None], 0.0)


def softmax_fwd(ac, bd, key_lens, Tk, mode, causal, scale, drop=None):
    H, B, Tq, ld = ac.shape
    s = ac[..., :Tk].clone()
    if mode in ("latest", "legacy"):
        s = s + _shift(bd, Tq, mode)
    s = s * scale
    valid = torch.arange(Tk)[None, :] < key_lens[:, None].long()  # [B,Tk]
    valid = valid[None, :, None, :].expand(H, B, Tq, Tk)
    if causal:
        valid = valid & torch.tril(torch.ones(Tq, Tk, dtype=torch.bool))[None, None]
    s = s.masked_fill(~valid, float("-inf"))
    p = torch.softmax(s, -1)
    p = torch.where(valid, p, torch.zeros_like(p))
    p = torch.nan_to_num(p, nan=0.0)
    out = torch.zeros(H, B, Tq, ld, dtype=_lowp())
    out[..., :Tk] = p.to(out.dtype)
    m = _mul(drop, out.shape)  # element index over the padded head-major tensor (softmax.cu / attention.cu)
    if m is None:
        return out, out
    return out, _r(out.float() * m)


def softmax_bwd(dpd, p, Tk, mode, scale, ld_bd=0, drop=None):
    H, B, Tq, ld = p.shape
    pf = p.float()[..., :Tk]
    m = _mul(drop, p.shape)
    dp = dpd[..., :Tk] if m is None else dpd[..., :Tk] * m[..., :Tk]
    ds = pf * (dp - (dp * pf).sum(-1, keepdim=True)) * scale
    dac = torch.zeros(H, B, Tq, ld, dtype=_lowp())
    dac[..., :Tk] = ds.to(dac.dtype)
    dbd = None
    if mode in ("latest", "legacy"):
        T = Tq
        bd = torch.zeros(H, B, T, ld_bd, requires_grad=True)
        (_shift(bd, T, mode) * ds).sum().backward()
        dbd = _r(bd.grad)
    return dac, dbd


def relattn_supported(Tq, Tk, dk, mode, causal=False):
    if dk != 64 or not ((mode == "latest" and Tq == Tk) or mode in (None, "none")):
        return False
    return Tk <= 192 or (not causal and Tk <= 16 * 192)   # longer key axes: key chunks + combine, same arithmetic


def relattn_fwd(qu, qv, k, v, pos, key_lens, B, Tq, Tk, H, mode, causal, scale, drop=None, want_pd=True):
    """the fused kernel = the unfused chain (scores stay fp32 on chip, as in attn_f1's fp32 outputs)"""
    ac = attn_f1(qu, k, B, Tq, Tk, H)
    bd = attn_f1(qv, pos, B, Tq, pos.shape[0], H, y_bcast=True) if mode == "latest" else None
    p, pd = softmax_fwd(ac, bd, key_lens, Tk, mode, causal, scale, drop=drop)
    ctx = attn_f2(pd, v, B, Tq, Tk, H)
    if not want_pd and pd is not p:
        pd = None
    return p, pd, ctx


def relattn_bwd_supported(Tq, Tk, dk, mode):
    return relattn_supported(Tq, Tk, dk, mode)


def relattn_bwd(dctx, qu, k, v, p, ctx, B, Tq, Tk, H, mode, scale, dq, dk, dv, drop=None, want_dac=False, qv=None, pos=None,
                dqv=None, dpos=None, dk_colsum=None, dv_colsum=None):
    """the fused backward = the unfused chain: dPd = dO v^T, dS = p (m dPd - <dO, O>) scale, dq = dS k, dk = dS^T q, dv = pd^T dO,
    and in rel-pos mode d(bd) = unshift(dS), dq_v = d(bd) pos, dpos += d(bd)^T q_v"""
    dpd = attn_f1(dctx, v, B, Tq, Tk, H)
    m = _mul(drop, p.shape)
    pf = p.float()
    pd = pf if m is None else pf * m
    Hd = dctx.shape[1]
    dkh = Hd // H
    delta = (dctx.float() * ctx.float()).reshape(B, Tq, H, dkh).sum(-1).permute(2, 0, 1)[..., None]  # [H,B,Tq,1]
    dP = dpd if m is None else dpd * m
    ds = torch.zeros_like(pf)
    ds[..., :Tk] = (pf * (dP - delta) * scale)[..., :Tk]
    ds = _r(ds)
    attn_f2(ds, k, B, Tq, Tk, H, out=dq)
    attn_f3(ds, qu, B, Tq, Tk, H, out=dk)
    attn_f3(_r(pd), dctx, B, Tq, Tk, H, out=dv)
    if dk_colsum is not None:  # from the fp32 accumulators in the kernel; here from the rounded rows (bf16-level difference)
        dk_colsum += dk.float().sum(0)
        dv_colsum += dv.float().sum(0)
    if mode == "latest":
        Pn = pos.shape[0]
        dbd = rel_unshift(ds, Tq, mode, round8(Pn))
        attn_f2(dbd, pos, B, Tq, Pn, H, out=dqv, y_bcast=True)
        attn_f4(dbd, qv, B, Tq, Pn, H, dpos)
    return ds if want_dac else None


def rel_unshift(dac, T, mode, ld_bd):
    H, B, Tq, ld = dac.shape
    bd = torch.zeros(H, B, T, ld_bd, requires_grad=True)
    (_shift(bd, T, mode) * dac.float()[..., :T]).sum().backward()
    return _r(bd.grad)


def glu_dwconv_fwd(pw1, w, bias, B, T, K, sums=None):
    C = pw1.shape[1] // 2
    x = pw1.float().reshape(B, T, 2 * C)
    g = x[..., :C] * torch.sigmoid(x[..., C:])
    y = F.conv1d(g.transpose(1, 2), w.reshape(C, 1, K), bias, padding=(K - 1) // 2, groups=C).transpose(1, 2).reshape(B * T, C)
    return _r(y), torch.cat([y.sum(0), (y * y).sum(0)])


def glu_dwconv_bwd(d_out, pw1, w, B, T, K, dw, dbias, dpw1_colsum=None):
    C = pw1.shape[1] // 2
    x = pw1.float().reshape(B, T, 2 * C).requires_grad_(True)
    wl = w.detach().clone().reshape(C, 1, K).requires_grad_(True)
    bl = torch.zeros(C, requires_grad=True)
    g = x[..., :C] * torch.sigmoid(x[..., C:])
    y = F.conv1d(g.transpose(1, 2), wl, bl, padding=(K - 1) // 2, groups=C).transpose(1, 2).reshape(B * T, C)
    y.backward(d_out.float())
    dw += wl.grad.reshape(dw.shape)
    dbias += bl.grad
    if dpw1_colsum is not None:
        dpw1_colsum += x.grad.reshape(B * T, 2 * C).sum(0)
    return _r(x.grad.reshape(B * T, 2 * C))


def bn_finalize(sums, n, eps, momentum, running_mean=None, running_var=None, num_batches=None):
    C = sums.numel() // 2
    mean = sums[:C] / n
    var = (sums[C:] / n - mean * mean).clamp(min=0)
    if running_mean is not None:
        running_mean.mul_(1 - momentum).add_(momentum * mean)
        running_var.mul_(1 - momentum).add_(momentum * var * n / max(n - 1, 1))
        num_batches += 1
    return torch.cat([mean, torch.rsqrt(var + eps)])


def bn_swish_fwd(x, mean_rstd, gamma, beta):
    C = x.shape[1]
    z = (x.float() - mean_rstd[:C]) * mean_rstd[C:] * gamma + beta
    return _r(z * torch.sigmoid(z))


def bn_swish_bwd(dy, x, mean_rstd, gamma, beta, dgamma, dbeta, scratch=None):
    C = x.shape[1]
    xh = (x.float() - mean_rstd[:C]) * mean_rstd[C:]
    z = xh * gamma + beta
    dz = dy.float() * _act_grad(z, "swish")
    dgamma += (dz * xh).sum(0)
    dbeta += dz.sum(0)
    return _r(gamma * mean_rstd[C:] * (dz - dz.mean(0) - xh * (dz * xh).mean(0)))


def conv1_fwd(x, w, bias, want_bits=False):
    C = w.shape[0]
    y = torch.relu(F.conv2d(x[:, None], w.reshape(C, 1, 3, 3), bias, stride=2))  # [B,C,T1,F1]
    h1 = _r(y.permute(0, 2, 3, 1).contiguous())
    return (h1, relu_bits(h1)) if want_bits else h1


def relu_bits(h1):
    """int32 [pixels, C/32]: bit (c % 32) of word c // 32 = (h1[pixel, c] > 0)."""
    C = h1.shape[-1]
    b = (h1.float().reshape(-1, C // 32, 32) > 0).to(torch.int64)
    v = (b << torch.arange(32)).sum(-1)
    return torch.where(v >= 2 ** 31, v - 2 ** 32, v).to(torch.int32)


def conv1_bwd(dh1, x, dw, db):
    C = dh1.shape[-1]
    wl = torch.zeros(C, 1, 3, 3, requires_grad=True)
    bl = torch.zeros(C, requires_grad=True)
    y = F.conv2d(x.bfloat16().float()[:, None], wl, bl, stride=2)   # the wgrad GEMM reads bf16 input patches
    y.backward(dh1.float().permute(0, 3, 1, 2))
    dw += wl.grad.reshape(dw.shape)
    db += bl.grad


def im2col(h1):
    B, T1, F1, C = h1.shape
    T2, F2 = (T1 - 3) // 2 + 1, (F1 - 3) // 2 + 1
    cols = []
    for kh in range(3):
        for kw in range(3):
            cols.append(h1[:, kh:kh + 2 * T2 - 1:2, kw:kw + 2 * F2 - 1:2, :])
    return torch.stack(cols, 3).reshape(B * T2 * F2, 9 * C)


def col2im_relu(dcol, h1):
    B, T1, F1, C = h1.shape
    T2, F2 = (T1 - 3) // 2 + 1, (F1 - 3) // 2 + 1
    d = dcol.float().reshape(B, T2, F2, 3, 3, C)
    out = torch.zeros(B, T1, F1, C)
    for kh in range(3):
        for kw in range(3):
            out[:, kh:kh + 2 * T2 - 1:2, kw:kw + 2 * F2 - 1:2, :] += d[:, :, :, kh, kw, :]
    return _r(out * (h1.float() > 0))


def conv2_implicit_supported(B, T1, F1, C):
    return C % 256 == 0


def conv2_fwd(h1, w16, bias):
    return gemm(im2col(h1), w16, bias=bias, act="relu")


def conv2_wgrad(dh2, h1, dw_accum):
    gemm(dh2, im2col(h1), trans=True, out=dw_accum, accumulate=True)


def conv2_dgrad_relu(dh2, w16, h1, bits=None):
    B, T1, F1, C = h1.shape
    T2, F2 = (T1 - 3) // 2 + 1, (F1 - 3) // 2 + 1
    d = (dh2.float() @ w16.float()).reshape(B, T2, F2, 3, 3, C)   # one rounding, after the sum over taps
    out = torch.zeros(B, T1, F1, C)
    for kh in range(3):
        for kw in range(3):
            out[:, kh:kh + 2 * T2 - 1:2, kw:kw + 2 * F2 - 1:2, :] += d[:, :, :, kh, kw, :]
    return _r(out * (h1.float() > 0))


def _ctc_inputs(logits, V, hlens, labels, label_lens):
    B, T, _ = logits.shape
    lp = logits[..., :V].detach().clone().requires_grad_(True)
    tgt = torch.cat([labels[b, : int(label_lens[b])] for b in range(B)]).long()
    return lp, tgt


def ctc_fwd(logits, V, hlens, labels, label_lens, blank=0):
    lp, tgt = _ctc_inputs(logits, V, hlens, labels, label_lens)
    loss = F.ctc_loss(lp.log_softmax(-1).transpose(0, 1), tgt, hlens.long(), label_lens.long(), blank=blank,
                      reduction="none", zero_infinity=True)
    return loss.detach(), loss.detach().sum().reshape(1), (lp, loss)


def ctc_bwd(logits, V, hlens, labels, label_lens, ws, scale, blank=0, want_f32=False, want_bf16=True):
    lp, loss = ws
    (g,) = torch.autograd.grad(loss.sum() * scale, lp)
    B, T, _ = logits.shape
    g32 = torch.zeros(B, T, round8(V))
    g32[..., :V] = g
    return (g32 if want_f32 else None), (_r(g32) if want_bf16 else None)


def lsm_loss(logits, V, targets, smoothing, grad_scale, ignore=-1, want_f32=False, want_bf16=True):
    x = logits[:, :V].detach().clone().requires_grad_(True)
    t = targets.long()
    ign = t == ignore
    tt = t.masked_fill(ign, 0)
    true = torch.full_like(x, smoothing / (V - 1))
    true.scatter_(1, tt[:, None], 1.0 - smoothing)
    kl = (true * (torch.log(true) - torch.log_softmax(x, 1))).masked_fill(ign[:, None], 0.0)
    loss = kl.sum()
    (g,) = torch.autograd.grad(loss * grad_scale, x)
    correct = ((x.argmax(1) == t) & ~ign).sum().float()
    stats = torch.stack([loss.detach(), correct, (~ign).sum().float()])
    g32 = torch.zeros(x.shape[0], round8(V))
    g32[:, :V] = g
    return stats, (g32 if want_f32 else None), (_r(g32) if want_bf16 else None)


def logmel(wav, lens, melmat, n_fft=512, hop=128):
    from oracle import model as om
    feats, flens = om.frontend_logmel(wav, lens.long(), melmat, n_fft, hop)
    return feats, flens.to(torch.int32)


def specaug_mvn(x, lens, center=None, warped=None, fmask=None, tmask=None, do_mvn=True):
    from oracle import model as om
    B = x.shape[0]
    y = x
    if center is not None or fmask is not None or tmask is not None:
        z = torch.zeros(B, 0, dtype=torch.long)
        draws = {
            "center": center.long() if center is not None else None,
            "warped": warped.long() if warped is not None else None,
            "f_pos": fmask[..., 0].long() if fmask is not None else z, "f_len": fmask[..., 1].long() if fmask is not None else z,
            "t_pos": tmask[..., 0].long() if tmask is not None else z, "t_len": tmask[..., 1].long() if tmask is not None else z,
        }
        y = om.spec_augment(x, lens.long(), draws)
    return om.utterance_mvn(y, lens.long()) if do_mvn else y


def global_mvn(x, lens, mean=None, inv_std=None):
    valid = (torch.arange(x.shape[1])[None, :] < lens[:, None].long())[:, :, None]
    y = x - mean if mean is not None else x.clone()
    y = torch.where(valid, y, torch.zeros_like(y))
    return y * inv_std if inv_std is not None else y


def vocab_softmax_fwd(logits, V):
    """softmax over the first V columns of f32 [M, ld]; bf16-rounded [M, ld] with zero padding columns."""
    out = torch.zeros_like(logits)
    out[:, :V] = torch.softmax(logits[:, :V].float(), dim=-1)
    return _r(out)


def vocab_log_softmax(logits, V):
    return torch.log_softmax(logits[:, :V].float(), dim=-1)


def vocab_softmax_bwd(sm, dsm, V, add=None):
    """dlogits = sm * (dsm - sum(sm * dsm)) (+ add) over the first V columns; padding columns 0."""
    s, d = sm[:, :V].float(), dsm[:, :V].float()
    out = torch.zeros(sm.shape[0], sm.shape[1])
    out[:, :V] = s * (d - (s * d).sum(-1, keepdim=True))
    if add is not None:
        out[:, :V] += add[:, :V].float()
    return _r(out)


def adam_step_sched(params, grads, exp_avg, exp_avg_sq, base_lr, warmup_steps, state4, betas=(0.9, 0.999), eps=1e-8,
                    weight_decay=0.0, grad_inv_scale=1.0, max_norm=5.0):
    t = float(state4[3]) + 1.0
    w = float(warmup_steps)
    lr = base_lr * w ** 0.5 * min(t ** -0.5, t * w ** -1.5)
    adam_step(params, grads, exp_avg, exp_avg_sq, lr, t, state4, betas, eps, weight_decay, grad_inv_scale, max_norm)
    if float(state4[2]) == 1.0:
        state4[3] += 1.0


def adam_step(params, grads, exp_avg, exp_avg_sq, lr, step, state3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0,
              grad_inv_scale=1.0, max_norm=5.0):
    g = grads * grad_inv_scale
    norm = g.double().pow(2).sum().sqrt()
    state3[0], state3[1], state3[2] = norm ** 2, norm, float(torch.isfinite(norm))
    if not torch.isfinite(norm):
        return
    if max_norm > 0:
        g = g * min(1.0, max_norm / (norm.item() + 1e-6))
    if weight_decay:
        g = g + weight_decay * params
    exp_avg.mul_(betas[0]).add_(g, alpha=1 - betas[0])
    exp_avg_sq.mul_(betas[1]).addcmul_(g, g, value=1 - betas[1])
    bc1, bc2 = 1 - betas[0] ** step, 1 - betas[1] ** step
    params.addcdiv_(exp_avg, exp_avg_sq.sqrt() / math.sqrt(bc2) + eps, value=-lr / bc1)
