"""Data-parallel path on CPU: world_size-2 ``gloo`` processes run the step orchestration (over the torch emulation of
the kernels), shard the batch the reference's way, reduce the flat gradient buffer through GradReducer's overlapped
buckets, and must reproduce the reference's DDP semantics: the batch-weighted mean of the per-shard gradients with
per-rank BatchNorm statistics (abs_task.py:1534-1542, trainer.py:594-606)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _problem(cond=False):
    from espnet_slurp_b200.testing import make_sd
    from oracle import model as om
    V = 24 if cond else 20  # the conditioning Linear wants V % 8 == 0
    enc, dec = om.EncCfg(80, 16, 2, 32, 2, 7, "latest"), om.DecCfg(V, 16, 2, 32, 1)
    sd = make_sd(enc, dec, V, seed=2)
    g = torch.Generator().manual_seed(4)
    if cond:
        sd["encoder.conditioning_layer.weight"] = torch.randn(16, V, generator=g) / V ** 0.5
        sd["encoder.conditioning_layer.bias"] = torch.randn(16, generator=g) * 0.1
    B, Tin = 5, 64
    feats = torch.randn(B, Tin, 80, generator=g)
    flens = torch.tensor([64, 60, 55, 48, 40])
    tl = torch.tensor([5, 3, 4, 2, 3])
    text = torch.full((B, 5), -1, dtype=torch.long)
    for b in range(B):
        text[b, : tl[b]] = torch.randint(1, V - 1, (int(tl[b]),), generator=g)
    ocfg = om.ModelCfg(enc=enc, dec=dec, vocab=V)
    if cond:
        ocfg.interctc_weight, ocfg.interctc_layer_idx, ocfg.interctc_use_conditioning = 0.4, (1,), True
    return ocfg, sd, feats, flens, text, tl


def _worker(rank, world, port, overlap, out_dir, cond=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import emul_ops
        from espnet_slurp_b200 import dist as bdist
        from espnet_slurp_b200 import params as bparams
        from espnet_slurp_b200 import step as bstep
        bstep.K = emul_ops
        emul_ops.exact = True
        ocfg, sd, feats, flens, text, tl = _problem(cond)
        V = ocfg.vocab
        cfg = bstep.ModelCfg(enc=bstep.EncCfg(80, 16, 2, 32, 2, 7, "latest", 0.0, 0.0, 0.0,
                                              interctc_layer_idx=(1,) if cond else (), interctc_use_conditioning=cond, vocab=V),
                             dec=bstep.DecCfg(V, 16, 2, 32, 1, 0.0, 0.0, 0.0, 0.0), vocab=V,
                             interctc_weight=0.4 if cond else 0.0)
        idx = bdist.shard_batch(list(range(feats.shape[0])), rank, world)
        pack = bparams.ParamPack(sd, lowp=torch.float32)
        pack.refresh()
        pack.zero_grad()
        # self-conditioning: the encoder backward still adds to ctc.ctc_lo's gradient after "heads" is reported final, so
        # those parameters must ride in the last bucket (what trainer.TrainStep asks for) -- an all-reduce issued at "heads"
        # would race with / miss those contributions
        red = bdist.GradReducer(pack, world, bucket_bytes=4096, overlap=overlap, late_prefixes=("ctc.",) if cond else ())
        stages = [s for s, _, _ in red.stages]
        assert stages[0] == "heads" and stages[-1] == "front" and any(s.startswith("enc") for s in stages)
        cover = torch.zeros(pack.total)
        for _, lo, hi in red.stages:
            cover[lo:hi] += 1
        assert bool((cover == 1).all()), "every gradient element belongs to exactly one bucket"
        w = bdist.loss_weight(len(idx), "cpu")
        called = []
        red.begin()

        def on_ready(stage):
            called.append(stage)
            red.ready(stage)

        out = bstep.train_step_from_feats(feats[idx], flens[idx].int(), text[idx], tl[idx], pack.P, pack.W, pack.G, cfg,
                                          bstep.StepCtx(0, True), grad_scale=w, on_ready=on_ready)
        pack.finalize_grads()
        red.finish()
        assert called == ["heads", "enc1", "enc0"], called
        g = {k: (v / world).numpy().copy() for k, v in pack.grads().items()}
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), loss=out["loss"].numpy(), w=np.float32(w), **g)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("overlap,cond", [(True, False), (False, False), (True, True)])
def test_two_rank_gradients_follow_reference_ddp(tmp_path, overlap, cond):
    from oracle import model as om
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, overlap, str(tmp_path), cond), nprocs=world, join=True)
    r0, r1 = np.load(tmp_path / "rank0.npz"), np.load(tmp_path / "rank1.npz")
    ocfg, sd, feats, flens, text, tl = _problem(cond)
    # reference semantics: every rank runs its shard (BatchNorm over the shard only), loss weights B_r * world / sum B,
    # DDP averages the gradients
    want = {k: np.zeros(v.shape, np.float32) for k, v in sd.items()}
    for r in range(world):
        idx = list(range(5))[r::world]
        osd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        loss, _ = om.asr_forward_from_feats(feats[idx], flens[idx], text[idx], tl[idx], osd, ocfg)
        wr = len(idx) * world / 5.0
        assert abs(float([r0, r1][r]["w"]) - wr) < 1e-6
        np.testing.assert_allclose(float([r0, r1][r]["loss"].reshape(-1)[0]), loss.item(), rtol=1e-4)
        (loss * wr).backward()
        for k, v in osd.items():
            if v.grad is not None:
                want[k] += v.grad.numpy() / world
    for k, v in want.items():
        scale = max(np.abs(v).max(), 1e-3)
        assert np.abs(r0[k] - v).max() / scale < 3e-3, k
        assert np.array_equal(r0[k], r1[k]), k      # both ranks hold the same reduced gradients


def test_shard_batch_matches_reference_rule():
    from espnet_slurp_b200 import dist as bdist
    keys = [f"utt{i}" for i in range(7)]
    assert bdist.shard_batch(keys, 1, 3) == ["utt1", "utt4"]
    assert sorted(sum((bdist.shard_batch(keys, r, 3) for r in range(3)), [])) == sorted(keys)
    with pytest.raises(RuntimeError):
        bdist.shard_batch(keys[:2], 0, 3)


def test_peer_reducer_slices_partition_each_bucket():
    """dist.shard_range: the ranks' slices of a bucket are disjoint, ordered, 8-element aligned and cover it exactly
    (also when the bucket is shorter than world * 8 elements and some slices are empty)."""
    from espnet_slurp_b200 import dist as bdist
    for lo, hi in ((0, 8), (16, 16 + 7 * 8), (1024, 1024 + 109_000_008), (8, 8 + 24)):
        for world in (2, 3, 8):
            pos, sizes = lo, []
            for r in range(world):
                a, b = bdist.shard_range(lo, hi, world, r)
                assert a == pos and a <= b <= hi and (a - lo) % 8 == 0
                pos = b
                sizes.append(b - a)
            assert pos == hi
            assert max(sizes) == bdist.shard_range(lo, hi, world, 0)[1] - lo  # rank 0's slice is the staging size
