"""dist.PeerReducer (gradient buckets exchanged by the copy engines over peer memory) against ncclAllReduce on the same
buffers: one process per visible GPU, at least two (skipped on a one-GPU box; `gpurun --gpus 2 -- python -m pytest tests/test_peer_reduce_gpu.py`).
The slice arithmetic is covered on the CPU in test_dist_gloo.py."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


class _Pack:
    """The part of params.ParamPack the reducers read: names, offsets, sizes, the flat gradient buffer."""

    def __init__(self, sizes, alloc):
        self.names = list(sizes)
        self.offs, off = {}, 0
        for n in self.names:
            self.offs[n] = off
            off += (sizes[n] + 7) // 8 * 8
        self.total, self._sizes = off, sizes
        self.gflat = alloc(off).zero_()

    def shapes_numel(self, n):
        return self._sizes[n]


def _worker(rank, world, port, scale, out):
    import torch.distributed as dist
    from espnet_slurp_b200 import dist as bdist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        sizes = {"encoder.embed.conv.0.weight": 2304 * scale + 5, "encoder.after_norm.weight": 512}
        for b in range(4):
            sizes["encoder.encoders.%d.feed_forward.w_1.weight" % b] = 100003 * scale
            sizes["encoder.encoders.%d.norm_ff.bias" % b] = 512
        sizes["decoder.output_layer.weight"] = 77777 * scale
        sizes["ctc.ctc_lo.bias"] = 600
        pack = _Pack(sizes, bdist.peer_grad_alloc(dev))
        red = bdist.PeerReducer(pack, world, bucket_bytes=(100003 * scale * 4 * 3) // 2)
        assert len(red.stages) >= 3, red.stages
        gen = torch.Generator(device=dev).manual_seed(1234 + rank)
        worst = 0.0
        for it in range(3):
            pack.gflat.copy_(torch.randn(pack.total, device=dev, generator=gen))
            want = pack.gflat.clone()
            dist.all_reduce(want)
            torch.cuda.synchronize()
            dist.barrier()
            red.begin()
            for stage in ["heads"] + ["enc%d" % b for b in (3, 2, 1, 0)]:
                red.ready(stage)
            red.finish()
            torch.cuda.synchronize()
            worst = max(worst, float((pack.gflat - want).abs().max()))
            # every replica holds the same bits
            mine = pack.gflat.view(torch.int32).to(torch.int64).sum().reshape(1)
            both = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(both, mine)
            assert all(int(b) == int(both[0]) for b in both), both
        assert worst < 1e-5, worst
        if out is not None and rank == 0:
            out.put(worst)
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
def test_peer_reducer_matches_nccl_all_reduce():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    world = min(torch.cuda.device_count(), 8)  # more than two ranks also deal a phase's copies over several streams
    procs = [ctx.Process(target=_worker, args=(r, world, 29611, 1, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
    for p in procs:
        if p.is_alive():
            p.terminate()
            pytest.fail("peer reducer worker hung")
        assert p.exitcode == 0
    assert out.get(timeout=5) < 1e-5


def _train_worker(rank, world, port, out):
    import numpy as np
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_step_gpu import make_sd
    from test_trainer_gpu import _batch, _cfg
    from espnet_slurp_b200.trainer import TrainStep

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        cfg, ocfg = _cfg()
        sd = make_sd(ocfg.enc, ocfg.dec, 30, seed=3)
        peer = TrainStep(cfg, sd, rank, world, dp_mode="graph", dp_comm="peer", lr=2e-3, warmup_steps=4)
        nccl = TrainStep(cfg, sd, rank, world, dp_mode="graph", dp_comm="nccl", lr=2e-3, warmup_steps=4)
        assert peer.dp_comm == "peer" and peer.peer_one_graph, getattr(peer, "dp_comm_note", None)
        for i in range(5):  # step 0 eager, step 1 captures, the rest replay
            b = {k: v.cuda() for k, v in _batch(3, 8000, 5, seed=100 + 7 * i + rank).items()}
            lp = peer(b).clone()
            ln = nccl(b).clone()
            torch.cuda.synchronize()
            np.testing.assert_allclose(lp.cpu().numpy()[:3], ln.cpu().numpy()[:3], rtol=1e-2, atol=1e-4, err_msg=f"step {i}")
            # the reduced gradients of the two exchanges (same sums in another order).  The parameters themselves may drift
            # apart by +-lr per step where a gradient is mathematically zero (Adam turns rounding noise into a sign), which
            # by definition does not move the loss or the other gradients.
            gp, gn = peer.pack.gflat, nccl.pack.gflat
            assert float((gp - gn).abs().max()) <= 1e-2 * float(gn.abs().max()), (i, float((gp - gn).abs().max()))
        assert peer.last.graph is not None and len(peer.last.graph) == 1   # one graph, the exchange inside it
        # the replicas stay bit-identical: same reduced gradients everywhere -> same parameters
        mine = peer.pack.flat32.view(torch.int32).to(torch.int64).sum().reshape(1)
        alls = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(alls, mine)
        assert all(int(a) == int(alls[0]) for a in alls), alls
        if rank == 0:
            out.put(float((peer.pack.flat32 - nccl.pack.flat32).abs().max()))
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
def test_train_step_peer_exchange_in_one_graph_matches_nccl_buckets():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    world = 2
    procs = [ctx.Process(target=_train_worker, args=(r, world, 29621, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
    for p in procs:
        if p.is_alive():
            p.terminate()
            pytest.fail("train-step worker hung")
        assert p.exitcode == 0
    assert out.get(timeout=5) < 5 * 2e-3 + 1e-3  # at most lr per step on the zero-gradient parameters
