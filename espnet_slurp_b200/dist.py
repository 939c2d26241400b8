"""Data-parallel plumbing of the train step: one process per GPU, torch.distributed for the rendezvous; the gradient
exchange runs over NVLink peer memory on the copy engines (PeerReducer, below) or as ncclAllReduce buckets (GradReducer).

Mirrors what the reference does with DistributedDataParallel:
  * each global mini-batch is dealt ``batch[rank::world_size]`` (espnet2/tasks/abs_task.py:1534-1542),
  * each rank runs the full model on its shard; BatchNorm statistics stay per rank (convolution.py:45),
  * the loss is the batch-weighted mean over ranks: the trainer multiplies each rank's loss by
    ``B_rank * world / sum(B)`` before DDP's mean (espnet2/train/trainer.py:594-606),
  * gradients are averaged with one bucketed all-reduce overlapped with the backward.

The path has exactly one exchange step, the gradient all-reduce; no other collective touches the data path.  The
gradients already sit in ONE flat fp32 buffer (params.ParamPack.gflat), so a bucket is a contiguous slice and needs no
copy; buckets are issued on a side stream as soon as the backward has finished the parameters they cover, in the order
the backward finishes them: decoder + CTC head first, then the encoder blocks last to first, then the subsampling
front.  NCCL's NVLS / in-switch reduction is used when the fabric offers it; nothing here depends on link count.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_batch(batch: Sequence, rank: int, world: int):
    """The reference's sharding of one mini-batch of utterance keys (abs_task.py:1534-1542): ``batch[rank::world]``."""
    if len(batch) < world:
        raise RuntimeError(f"The batch-size must be equal or more than world_size: {len(batch)} < {world}")  # :1528-1532
    return batch[rank::world]


def loss_weight(batch_size: int, device, group=None) -> float:
    """``B_rank * world / sum_r B_r``: the factor that turns DDP's plain mean into the batch-weighted mean
    (trainer.py:594-606).  One tiny all-reduce; 1.0 when every rank holds the same number of utterances."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return 1.0
    t = torch.tensor([float(batch_size)], device=device)
    dist.all_reduce(t, group=group)
    return float(batch_size) * dist.get_world_size(group) / float(t.item())


class GradReducer:
    """Bucketed all-reduce of the flat gradient buffer, overlapped with the backward."""

    def __init__(self, pack, world: Optional[int] = None, group=None, bucket_bytes: int = 64 << 20, overlap: bool = True,
                 late_prefixes: Sequence[str] = ()):
        """late_prefixes: parameter-name prefixes whose gradients keep changing after their usual stage (self-conditioned
        intermediate CTC: the encoder backward adds to ctc.ctc_lo after "heads"); they go into the last bucket."""
        self.pack, self.group = pack, group
        self.late_prefixes = tuple(late_prefixes)
        self.world = world if world is not None else (dist.get_world_size(group) if dist.is_initialized() else 1)
        self.overlap = overlap
        self.cuda = pack.gflat.is_cuda
        self.comm = torch.cuda.Stream(device=pack.gflat.device) if (self.cuda and self.world > 1) else None
        self.stages = self._plan(bucket_bytes)
        self._issued = 0
        self.handles = []

    # -- bucket plan ---------------------------------------------------------------------------------------------------
    def _plan(self, bucket_bytes: int) -> List[Tuple[str, int, int]]:
        """[(stage, lo, hi)] slices of gflat in backward-completion order.  Stages: "heads" (decoder + ctc, final once
        the decoder backward is done), "enc<n>" (final once encoder block n is done; small neighbours are merged up to
        bucket_bytes and keyed by the LAST block of the merge to finish), "front" (embed + anything else, final after
        ParamPack.finalize_grads)."""
        pk = self.pack
        end = {n: pk.offs[n] + (pk.shapes_numel(n) + 7) // 8 * 8 for n in pk.names}
        spans: Dict[str, List[int]] = {}

        def add(key, n):
            lo, hi = pk.offs[n], end[n]
            s = spans.setdefault(key, [lo, hi])
            s[0], s[1] = min(s[0], lo), max(s[1], hi)

        for n in pk.names:
            if any(n.startswith(lp) for lp in self.late_prefixes):
                continue  # left uncovered here -> part of the "front" remainder below
            if n.startswith("decoder.") or n.startswith("ctc."):
                add("heads", n)
            elif n.startswith("encoder.encoders."):
                add("enc%d" % int(n.split(".")[2]), n)
            else:
                add("front", n)
        out: List[Tuple[str, int, int]] = []
        if "heads" in spans:
            out.append(("heads", spans["heads"][0], spans["heads"][1]))
        blocks = sorted((int(k[3:]) for k in spans if k.startswith("enc")), reverse=True)
        cur = None
        for b in blocks:  # merge neighbouring blocks (contiguous in gflat) until the bucket is large enough
            lo, hi = spans["enc%d" % b]
            if cur is None:
                cur = [b, lo, hi]
            elif lo == cur[2] or hi == cur[1]:
                cur = [b, min(cur[1], lo), max(cur[2], hi)]
            else:
                out.append(("enc%d" % cur[0], cur[1], cur[2]))
                cur = [b, lo, hi]
            if (cur[2] - cur[1]) * 4 >= bucket_bytes:
                out.append(("enc%d" % cur[0], cur[1], cur[2]))
                cur = None
        if cur is not None:
            out.append(("enc%d" % cur[0], cur[1], cur[2]))
        # whatever is left (front of the encoder, after_norm, gaps) goes last
        covered = sorted((lo, hi) for _, lo, hi in out)
        pos = 0
        for lo, hi in covered + [(pk.total, pk.total)]:
            if lo > pos:
                out.append(("front", pos, lo))
            pos = max(pos, hi)
        return out

    # -- per step ------------------------------------------------------------------------------------------------------
    def begin(self):
        self._issued = 0
        self.handles = []

    def ready(self, stage: str):
        """Called by the step when ``stage`` ("heads", "enc<n>", "front") has its final gradients: issue, in plan order,
        every bucket up to and including that stage."""
        if self.world == 1:
            return
        if not self.overlap and stage != "front":
            return
        upto = max((i for i, (s, _, _) in enumerate(self.stages) if s == stage), default=-1)
        if stage == "front":
            upto = len(self.stages) - 1
        if upto < self._issued:
            return
        if self.comm is not None:
            self.comm.wait_stream(torch.cuda.current_stream())
        for i in range(self._issued, upto + 1):
            _, lo, hi = self.stages[i]
            view = self.pack.gflat[lo:hi]
            if self.comm is not None:
                with torch.cuda.stream(self.comm):
                    dist.all_reduce(view, group=self.group)
            else:
                self.handles.append(dist.all_reduce(view, group=self.group, async_op=True))
        self._issued = upto + 1

    def finish(self):
        """All buckets reduced (SUM) and visible to the current stream; the caller divides by world in the optimizer."""
        if self.world == 1:
            return
        self.ready("front")
        if self.comm is not None:
            torch.cuda.current_stream().wait_stream(self.comm)
        for h in self.handles:
            h.wait()
        self.handles = []


# ----------------------------------------------------------------------------------------------------------------------
# The same exchange without a single SM held by communication: peer memory + copy engines
# ----------------------------------------------------------------------------------------------------------------------
def shard_range(lo: int, hi: int, world: int, rank: int, align: int = 8) -> Tuple[int, int]:
    """Rank ``rank``'s slice of the bucket [lo, hi): ``world`` consecutive pieces of ceil((hi - lo) / world) elements rounded
    up to ``align`` (the last ones may be short or empty)."""
    m = -(-(hi - lo) // world)
    m = -(-m // align) * align
    a = min(lo + rank * m, hi)
    return a, min(a + m, hi)


def peer_grad_alloc(device):
    """Allocator of the flat gradient buffer for ``ParamPack(grad_alloc=...)``: memory every rank of the node can map
    (CUDA virtual-memory handles exchanged by torch's symmetric-memory rendezvous)."""
    import torch.distributed._symmetric_memory as symm

    return lambda n: symm.empty(n, dtype=torch.float32, device=device)


class PeerReducer(GradReducer):
    """GradReducer's bucket plan and call protocol, exchanged over NVLink PEER MEMORY by the COPY ENGINES.

    Why: the step's persistent kernels size their grids for all 148 SMs and fill the shared memory of each (the GEMM's
    operand ring takes the whole 227 KB), so a communication kernel cannot share an SM with them -- while an NCCL kernel
    holds some SMs, the CTAs assigned to those SMs start late and the whole statically partitioned kernel ends late
    (~1.0-1.5 ms per step whatever N is).  Here the flat gradient buffer lives in memory all ranks
    map (``peer_grad_alloc``) and a bucket [lo, hi) is reduced as

        barrier                      every rank's bucket is final (a one-CTA flag kernel, torch symmetric-memory barrier)
        pull   (N-1 DMA copies)      this rank's slice of the bucket, from every other rank's buffer into local staging
        add    (one small kernel)    own slice += staged slices (ops.add_shards_f32, fixed order)
        push   (N-1 DMA copies)      the reduced slice into every other rank's buffer

    on the communication stream; one more barrier at the end makes all pushes visible before clip + Adam.  Rank r alone
    computes slice r and every rank receives those very bits, so the replicas stay bit-identical.  No slice is written
    while anyone may still read it: only rank r reads slice r of the others (its pulls precede its pushes in stream
    order), and the closing barrier orders everything before the next step clears the buffer."""

    def __init__(self, pack, world: Optional[int] = None, group=None, bucket_bytes: int = 64 << 20, overlap: bool = True,
                 late_prefixes: Sequence[str] = ()):
        super().__init__(pack, world, group, bucket_bytes, overlap, late_prefixes)
        if self.world == 1:
            return
        import torch.distributed._symmetric_memory as symm

        self.hdl = symm.rendezvous(pack.gflat, group if group is not None else dist.group.WORLD)
        self.rank = self.hdl.rank
        assert self.hdl.world_size == self.world
        self.peer = [pack.gflat if r == self.rank else self.hdl.get_buffer(r, (pack.total,), torch.float32)
                     for r in range(self.world)]
        self.shard_max = max(shard_range(lo, hi, self.world, 0)[1] - lo for _, lo, hi in self.stages)
        self.staging = torch.empty((self.world - 1) * self.shard_max, device=pack.gflat.device, dtype=torch.float32)
        # above the compute stream: the flag and add kernels are tiny and take the first SM slot that frees up
        self.comm = torch.cuda.Stream(device=pack.gflat.device, priority=-1)
        # the N-1 copies of a phase are independent: dealt over a few streams they run on several copy engines at once
        # (what matters for the last, exposed bucket)
        self.lanes = [self.comm] + [torch.cuda.Stream(device=pack.gflat.device, priority=-1)
                                    for _ in range(min(self.world - 2, 3))]

    def _fan(self, copies):
        """Run the (dst, src) copies of one phase after everything queued on the communication stream, on all lanes;
        the communication stream continues after the last of them."""
        if len(self.lanes) > 1:
            start = torch.cuda.Event()
            start.record(self.comm)
        for k, (dst, src) in enumerate(copies):
            lane = self.lanes[k % len(self.lanes)]
            if lane is not self.comm and k < len(self.lanes):
                lane.wait_event(start)
            with torch.cuda.stream(lane):
                dst.copy_(src, non_blocking=True)
        for lane in self.lanes[1:min(len(self.lanes), len(copies))]:
            self.comm.wait_stream(lane)

    def ready(self, stage: str):
        if self.world == 1:
            return
        if not self.overlap and stage != "front":
            return
        upto = max((i for i, (s, _, _) in enumerate(self.stages) if s == stage), default=-1)
        if stage == "front":
            upto = len(self.stages) - 1
        if upto < self._issued:
            return
        from . import ops

        self.comm.wait_stream(torch.cuda.current_stream())
        g, W, r, sm = self.pack.gflat, self.world, self.rank, self.shard_max
        with torch.cuda.stream(self.comm):
            for i in range(self._issued, upto + 1):
                _, lo, hi = self.stages[i]
                a, b = shard_range(lo, hi, W, r)
                self.hdl.barrier(channel=0)
                if b > a:
                    others = [(r + s) % W for s in range(1, W)]
                    self._fan([(self.staging[k * sm: k * sm + (b - a)], self.peer[p][a:b]) for k, p in enumerate(others)])
                    ops.add_shards_f32(g[a:b], self.staging, sm, W - 1)
                    self._fan([(self.peer[p][a:b], g[a:b]) for p in others])
        self._issued = upto + 1

    def finish(self):
        if self.world == 1:
            return
        self.ready("front")
        with torch.cuda.stream(self.comm):
            self.hdl.barrier(channel=1)
        torch.cuda.current_stream().wait_stream(self.comm)
