"""tcgen05 GEMM vs torch fp32 matmul on the same bf16-rounded operands (GPU)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ops():
    from espnet_slurp_b200 import ops
    return ops


def ref_mm(a, b, trans):
    a32, b32 = a.float(), b.float()
    return (a32.t() @ b32) if trans else (a32 @ b32.t())


SHAPES = [(128, 128, 64), (128, 128, 512), (256, 512, 512), (3968, 512, 512), (3968, 2048, 512), (1000, 600, 512),
          (77, 40, 72), (4100, 1536, 512), (300, 512, 9728), (513, 1024, 2048)]


@pytest.mark.parametrize("M,N,K", SHAPES)
def test_gemm_tn(M, N, K):
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + K)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = torch.randn(N, K, device="cuda", generator=g).bfloat16()
    out = ops.gemm(a, b, out_dtype=torch.float32)
    torch.cuda.synchronize()
    want = ref_mm(a, b, False)
    err = (out - want).abs().max().item() / (want.abs().max().item() + 1e-6)
    assert err < 1e-4, err


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (512, 512, 3968), (2048, 512, 4000), (512, 9728, 777), (600, 512, 1344),
                                   (72, 40, 200)])
def test_gemm_nt_wgrad(M, N, K):
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(K, M, device="cuda", generator=g).bfloat16()
    b = torch.randn(K, N, device="cuda", generator=g).bfloat16()
    out = torch.zeros(M, N, device="cuda")
    ops.gemm(a, b, trans=True, out=out, accumulate=True)
    ops.gemm(a, b, trans=True, out=out, accumulate=True)
    torch.cuda.synchronize()
    want = 2 * ref_mm(a, b, True)
    err = (out - want).abs().max().item() / (want.abs().max().item() + 1e-6)
    assert err < 1e-4, err


def test_gemm_epilogue_bias_swish_residual():
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(5)
    M, N, K = 1000, 512, 256
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) * 0.1).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g)
    out = ops.gemm(a, b, bias=bias, act="swish", alpha=0.5, residual=res, out_dtype=torch.float32)
    y = ref_mm(a, b, False) + bias
    want = res + 0.5 * y * torch.sigmoid(y)
    assert (out - want).abs().max().item() < 3e-3
    outb, dact = ops.gemm(a, b, bias=bias, act="swish", act_grad_out=True)   # the saved derivative rides with a bf16 output
    assert (outb.float() - y * torch.sigmoid(y)).abs().max().item() < 0.03
    sg = torch.sigmoid(y)
    want_d = sg * (1 + y * (1 - sg))   # swish'(y): what the forward saves for the dgrad epilogue
    assert (dact.float() - want_d).abs().max().item() < 0.01
    # saved-derivative epilogue: dH = (g @ W) * dact
    gy = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    out1 = ops.gemm(gy, b, act_grad_input=dact, act_grad="mul")
    want1 = ref_mm(gy, b, False) * dact.float()
    assert (out1.float() - want1).abs().max().item() < 0.02 * want1.abs().max().item()
    # activation-gradient epilogue from a pre-activation tensor: dH = (g @ W) * swish'(pre)
    pre = y.bfloat16()
    out2 = ops.gemm(gy, b, act_grad_input=pre, act_grad="swish")
    s = torch.sigmoid(pre.float())
    want2 = ref_mm(gy, b, False) * (s * (1 + pre.float() * (1 - s)))
    assert (out2.float() - want2).abs().max().item() < 0.05 * want2.abs().max().item()
    out3 = ops.gemm(gy, b, act_grad_input=pre, act_grad="relu", bias=bias, act="relu")
    want3 = torch.relu(ref_mm(gy, b, False) + bias) * (pre.float() > 0)
    assert (out3.float() - want3).abs().max().item() < 0.02 * want3.abs().max().item()
    # ReLU forward saves a 0/1 mask; with dropout the saved tensor carries the element's multiplier
    o4, d4 = ops.gemm(a, b, bias=bias, act="relu", act_grad_out=True, drop=(0.25, 11, 3))
    o4f, d4f = o4.float(), d4.float()
    yb = torch.relu(y)
    kept = d4f > 0
    assert ((d4f == 0) | ((d4f - 1 / 0.75).abs() < 0.01)).all()
    assert ((o4f - yb / 0.75).abs()[kept] < 0.02 * yb.abs().max()).all()
    assert (o4f[(y > 0.05) & ~kept] == 0).all()
    frac = kept[y > 0].float().mean().item()
    assert abs(frac - 0.75) < 0.01


@pytest.mark.parametrize("M,N,K", [(1000, 512, 2048), (333, 600, 1024), (4100, 768, 4608)])
def test_gemm_cta_pair_epilogues(M, N, K):
    """K >= 1024, unbatched: the host picks the CTA-pair kernel (cta_group::2, 256-row tiles).  Ragged M / N edges,
    every majorness, bias + residual (fp32), saved-derivative product, split-K accumulation."""
    ops = _ops()
    g = torch.Generator(device="cuda").manual_seed(M + 3 * N + K)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g)
    y = ref_mm(a, w, False)
    scale = y.abs().max().item()
    out = ops.gemm(a, w, bias=bias, residual=res, alpha=0.5, out_dtype=torch.float32)
    assert (out - (res + 0.5 * (y + bias))).abs().max().item() < 2e-4 * scale + 1e-4
    outb = ops.gemm(a, w, bias=bias, act="relu")
    assert (outb.float() - torch.relu(y + bias)).abs().max().item() < 0.01 * scale
    # dgrad: dX[M,K'] = dY[M,N'] W[N',K'] with N' playing K (long), mul epilogue
    wt = (torch.randn(K, N, device="cuda", generator=g) * 0.05).bfloat16()       # [K_contract, N_out]
    d = torch.rand(M, N, device="cuda", generator=g).bfloat16()
    o2 = ops.gemm(a, wt, trans_b=True, act_grad_input=d, act_grad="mul")
    want2 = (a.float() @ wt.float()) * d.float()
    assert (o2.float() - want2).abs().max().item() < 0.02 * want2.abs().max().item()
    # wgrad: dW[N, K2] += dY^T X over M rows (contraction = M)
    x2 = torch.randn(M, 512, device="cuda", generator=g).bfloat16()
    dy = torch.randn(M, N, device="cuda", generator=g).bfloat16()
    acc = torch.ones(N, 512, device="cuda")
    ops.gemm(dy, x2, trans=True, out=acc, accumulate=True)
    want3 = 1 + dy.float().t() @ x2.float()
    assert (acc - want3).abs().max().item() < 1e-4 * want3.abs().max().item() + 1e-3


def test_gemm_dropout_statistics_and_determinism():
    ops = _ops()
    M, N, K = 2048, 512, 64
    a = torch.ones(M, K, device="cuda").bfloat16()
    b = torch.ones(N, K, device="cuda").bfloat16()
    o1 = ops.gemm(a, b, dropout_p=0.1, seed=1234, site=7, out_dtype=torch.float32)
    o2 = ops.gemm(a, b, dropout_p=0.1, seed=1234, site=7, out_dtype=torch.float32)
    o3 = ops.gemm(a, b, dropout_p=0.1, seed=1234, site=8, out_dtype=torch.float32)
    assert torch.equal(o1, o2) and not torch.equal(o1, o3)
    keep = (o1 != 0).float().mean().item()
    assert abs(keep - 0.9) < 0.005, keep
    assert abs(o1.max().item() - 64 / 0.9) < 1e-3


@pytest.mark.parametrize("M,N,K_", [(1000, 512, 264), (19, 64, 64), (4100, 2048, 512)])
def test_gemm_output_colsum(M, N, K_):
    """out_colsum_accum: the column sums of the bf16 output, taken from the staged tiles in the epilogue, equal the sums of
    the output that was stored (fp32 accumulation of bf16 values; accumulated onto the buffer's contents)."""
    from espnet_slurp_b200 import ops
    g = torch.Generator().manual_seed(M + N)
    a = torch.randn(M, K_, generator=g).bfloat16()
    w = (torch.randn(K_, N, generator=g) * 0.1).bfloat16()
    aux = torch.rand(M, N, generator=g).bfloat16()
    for kw in (dict(), dict(act_grad_input=aux.cuda(), act_grad="mul")):
        cs = torch.full((N,), 2.0, device="cuda")
        out = ops.gemm(a.cuda(), w.cuda(), trans_b=True, colsum_out=cs, **kw)
        want = 2.0 + out.float().sum(0)
        torch.testing.assert_close(cs, want, rtol=2e-4, atol=2e-3 * float(out.float().abs().max()))
        ref = ops.gemm(a.cuda(), w.cuda(), trans_b=True, **kw)
        assert torch.equal(out, ref)
