"""espnet_slurp_b200.trainer.TrainStep on the GPU: CUDA graphs cached per batch shape must give the same
losses and parameters as eager launches, the optimizer must follow torch.optim.Adam + clip_grad_norm_ + WarmupLR
(trainer.py:640-690, warmup_lr.py:43-50) on the oracle's gradients, and the shape cache must evict least-recently-used."""
import numpy as np
import pytest
import torch

from oracle import model as om
from test_step_gpu import make_sd

pytestmark = pytest.mark.gpu


def _cfg():
    from espnet_slurp_b200 import step as bstep
    enc = bstep.EncCfg(80, 64, 4, 128, 2, 7, "latest", 0.0, 0.0, 0.0)
    dec = bstep.DecCfg(30, 64, 4, 128, 1, 0.0, 0.0, 0.0, 0.0)
    return bstep.ModelCfg(enc=enc, dec=dec, vocab=30), om.ModelCfg(enc=om.EncCfg(80, 64, 4, 128, 2, 7, "latest"),
                                                                 dec=om.DecCfg(30, 64, 4, 128, 1), vocab=30)


def _batch(B, N, L, seed):
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(N * 3 // 5, N + 1, (B,), generator=g)
    lens[0] = N
    wav = torch.randn(B, N, generator=g) * 0.1
    for b in range(B):
        wav[b, lens[b]:] = 0
    tl = torch.randint(2, L + 1, (B,), generator=g)
    text = torch.randint(2, 28, (B, L), generator=g)
    text = torch.where(torch.arange(L)[None] < tl[:, None], text, torch.full_like(text, -1))
    return {"wav": wav, "lens": lens.int(), "text": text.int(), "tl": tl.int()}


def test_graph_buckets_match_eager_and_evict():
    from espnet_slurp_b200.trainer import TrainStep
    cfg, ocfg = _cfg()
    sd = make_sd(ocfg.enc, ocfg.dec, 30, seed=3)
    eager = TrainStep(cfg, sd, use_graph=False, lr=2e-3, warmup_steps=4)
    graph = TrainStep(cfg, sd, use_graph=True, lr=2e-3, warmup_steps=4, max_buckets=2)
    shapes = [(3, 8000, 5), (2, 12000, 7)]
    batches = [_batch(*shapes[i % 2], seed=10 + i) for i in range(8)]
    for i, b in enumerate(batches):
        le = eager({k: v.cuda() for k, v in b.items()}).clone()
        lg = graph({k: v.pin_memory() for k, v in b.items()}, fetch_loss=True).clone()  # host inputs, like the e2e path
        torch.cuda.synchronize()
        # same kernels, same inputs; only the order of the floating-point atomics differs from run to run
        # (losses; the accuracy is an argmax count over ~10 tokens: one near-tie flips it by 1 / #tokens)
        np.testing.assert_allclose(lg.cpu().numpy()[:3], le.cpu().numpy()[:3], rtol=1e-2, atol=1e-4, err_msg=f"step {i}")
        assert abs(float(lg[3]) - float(le[3])) <= 0.15
        np.testing.assert_allclose(graph.loss_host.numpy(), lg.cpu().numpy(), rtol=0, atol=0)
    assert len(graph.buckets) == 2 and all(b.graph is not None for b in graph.buckets.values())
    assert all(b.eager_steps == 1 for b in graph.buckets.values())  # one eager step per shape, replays afterwards
    pe, pg = eager.state_dict(), graph.state_dict()
    # both took the same 8 Adam steps; the summation order of the atomics differs between runs and Adam amplifies noise on
    # near-zero gradients (a mathematically zero one -- the softmax-shift-invariant linear_k.bias, the depthwise bias in
    # front of BatchNorm -- becomes +-lr noise): compare the total update, those two families left out
    ue, ug = [], []
    for k in pe:
        if k.endswith(("linear_k.bias", "depthwise_conv.bias")) or "running_" in k or "num_batches" in k:
            continue
        ue.append((pe[k].cpu().float() - sd[k]).flatten())
        ug.append((pg[k].cpu().float() - sd[k]).flatten())
    ue, ug = torch.cat(ue), torch.cat(ug)
    cos = float(torch.dot(ue, ug) / (ue.norm() * ug.norm()))
    assert cos > 0.97 and abs(float(ug.norm() / ue.norm()) - 1) < 0.05, (cos, float(ug.norm() / ue.norm()))
    assert float(eager.bn["encoder.encoders.0.conv_module.norm.num_batches_tracked"]) == 8
    assert eager.step_no == graph.step_no == 8 and float(graph.state[2]) == 1.0  # last update applied (finite norm)
    # a third shape evicts the least recently used one
    first_key = next(iter(graph.buckets))
    graph({k: v.cuda() for k, v in _batch(2, 9000, 4, seed=99).items()})
    assert len(graph.buckets) == 2 and first_key not in graph.buckets


def test_optimizer_follows_reference_recipe():
    """Two steps from the same start: clip_grad_norm_(5.0) + Adam(lr from WarmupLR) applied to the oracle's fp32 gradients."""
    from espnet_slurp_b200.trainer import TrainStep
    cfg, ocfg = _cfg()
    sd = make_sd(ocfg.enc, ocfg.dec, 30, seed=4)
    ts = TrainStep(cfg, sd, use_graph=False, lr=1e-3, warmup_steps=10)
    osd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    opt = torch.optim.Adam(list(osd.values()), lr=1e-3)
    mel = torch.from_numpy(om.slaney_mel_matrix())
    for i in range(2):
        b = _batch(3, 8000, 5, seed=20 + i)
        loss4 = ts({k: v.cuda() for k, v in b.items()})
        step = i + 1
        for gp in opt.param_groups:
            gp["lr"] = 1e-3 * 10 ** 0.5 * min(step ** -0.5, step * 10 ** -1.5)
        opt.zero_grad()
        oloss, _ = om.asr_forward(b["wav"], b["lens"].long(), b["text"].long(), b["tl"].long(), osd, ocfg, mel)
        oloss.backward()
        torch.nn.utils.clip_grad_norm_(list(osd.values()), 5.0)
        opt.step()
        np.testing.assert_allclose(loss4[0].item(), oloss.item(), rtol=2e-2)
    got = ts.state_dict()
    # Adam's first steps move every weight by ~lr whatever the gradient's scale: compare sign and size of the UPDATES where
    # the reference's update is not negligible (parameters with a mathematically zero gradient get +-lr noise here, from
    # bf16 rounding, and ~0 in the fp32 oracle, where eps dominates)
    agree = total = 0
    ur, ug = [], []
    for k, v in osd.items():
        if k.endswith(("linear_k.bias", "depthwise_conv.bias")):
            continue
        du_ref = (v.detach() - sd[k]).flatten()
        du_got = (got[k].cpu() - sd[k]).flatten()
        big = du_ref.abs() > 0.5 * du_ref.abs().max()
        agree += int((torch.sign(du_ref[big]) == torch.sign(du_got[big])).sum())
        total += int(big.sum())
        ur.append(du_ref)
        ug.append(du_got)
    ur, ug = torch.cat(ur), torch.cat(ug)
    assert agree / total > 0.97, (agree, total)
    assert float(ug.abs().max()) <= 1.05 * float(ur.abs().max()), (float(ug.abs().max()), float(ur.abs().max()))
    cos = float(torch.dot(ur, ug) / (ur.norm() * ug.norm()))
    assert cos > 0.9, cos


def test_chained_graph_segments_replay_like_one_graph():
    """dp_mode "graph" with more than one rank cuts the captured step into a chain of graphs (shared memory pool) so that
    gradient buckets can be all-reduced between replays.  On one GPU the same chain (cuts forced after the heads and after
    encoder block 1) must reproduce the eager step: same losses, same parameters after several optimizer steps."""
    from espnet_slurp_b200.trainer import TrainStep
    cfg, ocfg = _cfg()
    sd = make_sd(ocfg.enc, ocfg.dec, 30, seed=5)
    eager = TrainStep(cfg, sd, use_graph=False, lr=2e-3, warmup_steps=4)
    chain = TrainStep(cfg, sd, use_graph=True, lr=2e-3, warmup_steps=4)
    chain.force_cuts = ("heads", "enc1")
    batches = [_batch(3, 8000, 5, seed=20 + i) for i in range(6)]
    for i, b in enumerate(batches):
        le = eager({k: v.cuda() for k, v in b.items()}).clone()
        lc = chain({k: v.cuda() for k, v in b.items()}).clone()
        torch.cuda.synchronize()
        np.testing.assert_allclose(lc.cpu().numpy()[:3], le.cpu().numpy()[:3], rtol=1e-2, atol=1e-4, err_msg=f"step {i}")
        assert abs(float(lc[3]) - float(le[3])) <= 0.15
    bk = next(iter(chain.buckets.values()))
    assert [st for _, st in bk.graph] == ["heads", "enc1", None]  # three segments, replayed in capture order
    pe, pc = eager.state_dict(), chain.state_dict()
    ue, uc = [], []
    for k in pe:
        if k.endswith(("linear_k.bias", "depthwise_conv.bias")) or "running_" in k or "num_batches" in k:
            continue
        ue.append((pe[k].cpu().float() - sd[k]).flatten())
        uc.append((pc[k].cpu().float() - sd[k]).flatten())
    ue, uc = torch.cat(ue), torch.cat(uc)
    cos = float(torch.dot(ue, uc) / (ue.norm() * uc.norm()))
    assert cos > 0.97 and abs(float(uc.norm() / ue.norm()) - 1) < 0.05, (cos, float(uc.norm() / ue.norm()))


def test_close_unregisters_the_device_record():
    """A TrainStep registers its 24-byte device record with the library; close() / __del__ must unregister it so that later
    users of the library (drop-in modules with dropout on) never read a pointer that went back to the allocator."""
    from espnet_slurp_b200 import ops
    from espnet_slurp_b200.trainer import TrainStep
    import dropout_ref
    cfg, ocfg = _cfg()
    ts = TrainStep(cfg, make_sd(ocfg.enc, ocfg.dec, 30, seed=5), use_graph=False)
    ts.sp_np[0] = (12345, 0.0, 0.0, 0.0, 0.0)
    ts.sp_dev.copy_(ts.sp_host[0])
    got = ops.dropout_mask((0.1, 7, 3), 4096).cpu().numpy()
    assert np.array_equal(got, dropout_ref.keep_mask(0.1, 7, 3, 4096, dev_seed=12345))
    del ts
    import gc
    gc.collect()
    got = ops.dropout_mask((0.1, 7, 3), 4096).cpu().numpy()
    assert np.array_equal(got, dropout_ref.keep_mask(0.1, 7, 3, 4096))


def test_skipped_step_does_not_advance_schedule_and_state_exports():
    """b200slu_adam_step_sched vs torch.optim.Adam + WarmupLR driven the way trainer.py:651-690 drives them: a step whose
    gradient norm is not finite skips optimizer.step() AND scheduler.step(), so the next applied update still uses t = k+1."""
    from espnet_slurp_b200 import ops
    g = torch.Generator().manual_seed(0)
    n, base_lr, warm = 4096, 2e-3, 3.0
    p0 = torch.randn(n, generator=g)
    ref = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref], lr=base_lr)
    p = p0.cuda()
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    st = torch.zeros(4, device="cuda", dtype=torch.float64)
    applied = 0
    for i in range(6):
        grad = torch.randn(n, generator=g) * (10.0 if i == 1 else 0.01)   # step 1 gets clipped
        if i in (2, 4):
            grad[17] = float("inf") if i == 2 else float("nan")
        before = p.clone()
        ops.adam_step_sched(p, grad.cuda(), m, v, base_lr, warm, st, max_norm=5.0)
        if i in (2, 4):
            assert st[2].item() == 0.0 and st[3].item() == applied and torch.equal(p, before)
            continue
        applied += 1
        assert st[2].item() == 1.0 and st[3].item() == applied
        for gp in opt.param_groups:
            gp["lr"] = base_lr * warm ** 0.5 * min(applied ** -0.5, applied * warm ** -1.5)
        ref.grad = grad.clone()
        torch.nn.utils.clip_grad_norm_([ref], 5.0)
        opt.step()
        torch.testing.assert_close(p.cpu(), ref.detach(), rtol=2e-5, atol=2e-6)
    sto = opt.state[ref]
    assert float(sto["step"]) == applied
    torch.testing.assert_close(m.cpu(), sto["exp_avg"], rtol=1e-4, atol=1e-8)
    torch.testing.assert_close(v.cpu(), sto["exp_avg_sq"], rtol=1e-4, atol=1e-10)

    # the trainer exports that state per parameter name in torch.optim.Adam's vocabulary, and takes it back
    from espnet_slurp_b200.trainer import TrainStep
    cfg, ocfg = _cfg()
    sd = make_sd(ocfg.enc, ocfg.dec, 30, seed=6)
    a = TrainStep(cfg, sd, use_graph=False, lr=1e-3, warmup_steps=10)
    for i in range(2):
        a({k: t.cuda() for k, t in _batch(2, 6000, 4, seed=40 + i).items()})
    osd = a.optimizer_state_dict()
    assert set(osd) == set(sd) and all(float(s["step"]) == 2 and s["exp_avg"].shape == sd[k].shape for k, s in osd.items())
    b = TrainStep(cfg, a.state_dict(), use_graph=False, lr=1e-3, warmup_steps=10)
    b.load_optimizer_state_dict(osd)
    nb = {k: t.cuda() for k, t in _batch(2, 6000, 4, seed=50).items()}
    b.step_no = a.step_no            # same dropout / SpecAug stream position (all rates are 0 here)
    la, lb = a(nb), b(nb)
    torch.cuda.synchronize()
    torch.testing.assert_close(la, lb, rtol=1e-5, atol=1e-6)
    sa, sb = a.state_dict(), b.state_dict()
    # split-K wgrad sums arrive in a different order from run to run, so the two updates agree to rounding, not bitwise;
    # parameters whose gradient is mathematically zero move by +-lr on that noise and are left out
    worst = max(float((sa[k] - sb[k]).abs().max()) for k in sd if not k.endswith(("linear_k.bias", "depthwise_conv.bias")))
    assert worst < 2e-5, worst


def test_prefetched_batches_give_the_same_steps():
    """TrainStep.prefetch: the next batch's H2D copy on the copy stream + a device-to-device pass into the graph's inputs
    must be indistinguishable from the inline copy, over alternating shapes and reused pinned buffers."""
    from espnet_slurp_b200.trainer import TrainStep
    cfg, ocfg = _cfg()
    sd = make_sd(ocfg.enc, ocfg.dec, 30, seed=8)
    a = TrainStep(cfg, sd, use_graph=True, lr=2e-3, warmup_steps=4)
    b = TrainStep(cfg, sd, use_graph=True, lr=2e-3, warmup_steps=4)
    shapes = [(3, 8000, 5), (2, 6000, 4)]
    pinned = [{k: v.pin_memory() for k, v in _batch(*shapes[i % 2], seed=60 + i).items()} for i in range(2)]
    ready = [None, None]
    nxt = None
    for i in range(6):
        fresh = _batch(*shapes[i % 2], seed=70 + i)
        la = a({k: v.cuda() for k, v in fresh.items()}).clone()
        hb = pinned[i % 2]
        if nxt is None:  # first step: inline copy from the pinned buffers
            for k, v in fresh.items():
                hb[k].copy_(v)
        lb = b(hb).clone()
        if i + 1 < 6:  # prepare and prefetch the next batch while this step runs
            j = (i + 1) % 2
            if ready[j] is not None:
                ready[j].synchronize()
            for k, v in _batch(*shapes[j], seed=70 + i + 1).items():
                pinned[j][k].copy_(v)
            ready[j] = b.prefetch(pinned[j])
            nxt = i + 1
        torch.cuda.synchronize()
        # two runs of the same steps differ by the order of the split-K reduce-adds (1e-4 relative after a few Adam steps);
        # a stale or half-copied batch would be a different utterance set: losses apart by several percent
        np.testing.assert_allclose(la[:3].cpu().numpy(), lb[:3].cpu().numpy(), rtol=3e-3, atol=1e-4)
