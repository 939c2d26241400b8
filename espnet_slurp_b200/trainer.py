"""One rank of the data-parallel train step, replayed from CUDA graphs kept per batch shape.

Mirrors one iteration of the reference's training loop -- espnet2/train/trainer.py:464-720: forward, backward, gradient
all-reduce (DDP averages over ranks), grad-norm clipping at 5.0, the non-finite-norm skip, Adam, WarmupLR
(espnet2/schedulers/warmup_lr.py:43-50) -- over the kernels of this package:

    step = TrainStep(cfg, state_dict, rank, world)
    loss4 = step(batch)        # batch: dict of tensors, pinned host or device:
                               #   wav f32 [B, N], lens i32 [B], text i32 [B, Lmax], tl i32 [B]
                               #   and optionally the SpecAug draws center / warped i32 [B], fmask / tmask i32 [B, n, 2]
                               #   (modules.specaug_draws makes them in the reference's RNG order)

Nothing in a step depends on a host value: lengths, targets and SpecAug draws are device tensors, the dropout seed, the
learning rate and Adam's bias corrections are read by the kernels from a 24-byte device record refreshed per step
(ops.set_step_params).  A step is therefore capturable as ONE CUDA graph per input shape.  Batches of the reference's
samplers come in a handful of padded shapes (length-bucketed batching: `batch_type: folded / numel`, sorted by length), so
the graphs are cached per shape ("length buckets") with least-recently-used eviction: the first step of a new shape runs
eagerly (one-time kernel attribute setup), the next one is captured, later ones replay.

More than one rank (the reference's DDP overlaps its bucketed all-reduce with the backward, trainer.py:221-236).  The
exchange itself is dp_comm="peer" (default; dist.PeerReducer: the buckets travel over NVLink peer memory on the copy
engines, no SM is held by communication) or dp_comm="nccl" (dist.GradReducer: ncclAllReduce per bucket):
  dp_mode="graph" (default)   dp_comm="peer": ONE graph per batch shape as on a single GPU -- the exchange of each bucket
                              (flag kernel, DMA copies, one add kernel) is captured as a parallel branch that forks where
                              step.py reports the bucket final and joins before clip + Adam.
                              dp_comm="nccl": the step is captured as a CHAIN of graphs that share one memory pool, cut where a gradient
                              bucket of dist.GradReducer becomes final (decoder + CTC head, then every few encoder blocks,
                              last to first).  Between two replays the host issues that bucket's NCCL all-reduce on the
                              communication stream, so it runs under the next segment's kernels; only the small front
                              bucket (subsampling convs) is reduced after the last segment.  NCCL calls stay outside any
                              capture (capturing them hung on this stack).
  dp_mode="graph_serial"      one graph, then ONE all-reduce of the whole 437 MB gradient buffer, then clip + Adam
                              (round 1's measured mode: 0.94 ms exposed at 8 GPUs).
  dp_mode="eager_overlap"     eager launches with the bucketed all-reduce overlapping the backward.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Dict, Optional

import numpy as np
import torch

from . import ops, params as bparams, step as bstep
from .dist import GradReducer, PeerReducer, peer_grad_alloc
from .melbank import mel_matrix

_DRAW_KEYS = ("center", "warped", "fmask", "tmask")


class _Bucket:
    """Static input buffers, result buffer and (once captured) the CUDA graph of one batch shape."""

    def __init__(self, batch: Dict[str, torch.Tensor], dev):
        self.inp = {k: torch.empty(v.shape, dtype=v.dtype, device=dev) for k, v in batch.items()}
        self.res: Optional[torch.Tensor] = None
        self.graph = None  # once captured: list of (CUDAGraph, stage whose gradient bucket is final after it or None)
        self.eager_steps = 0
        # prefetch(): two device staging sets filled on the copy stream, alternating; `free[i]` = the D2D that last read set i
        self.staging = [None, None]
        self.free = [None, None]
        self.next_set = 0


class TrainStep:
    def __init__(self, cfg: bstep.ModelCfg, state_dict: Dict[str, torch.Tensor], rank: int = 0, world: int = 1,
                 dp_mode: str = "graph", use_graph: bool = True, lr: float = 1e-3, warmup_steps: int = 25000,
                 max_norm: float = 5.0, max_buckets: int = 4, eager_steps_before_capture: int = 1,
                 dp_comm: Optional[str] = None):
        """dp_comm (more than one rank): "peer" = gradient buckets exchanged by the copy engines over peer memory
        (dist.PeerReducer), "nccl" = ncclAllReduce per bucket (dist.GradReducer); default: env B200SLU_DP_COMM or "peer",
        falling back to "nccl" (recorded in ``self.dp_comm``) when the ranks cannot map each other's memory."""
        assert dp_mode in ("graph", "graph_serial", "eager_overlap"), dp_mode
        dev = torch.device("cuda", torch.cuda.current_device())
        self.cfg, self.rank, self.world, self.dev, self.dp_mode = cfg, rank, world, dev, dp_mode
        self.use_graph = use_graph and not (world > 1 and dp_mode == "eager_overlap")
        self.base_lr, self.warmup_steps, self.max_norm = lr, warmup_steps, max_norm
        self.max_buckets, self.eager_before = max_buckets, eager_steps_before_capture
        self.dp_comm = (dp_comm or os.environ.get("B200SLU_DP_COMM", "peer")) if world > 1 else None
        assert self.dp_comm in (None, "peer", "nccl"), self.dp_comm
        galloc = None
        if self.dp_comm == "peer":
            try:
                galloc = peer_grad_alloc(dev)
                galloc(8)
            except Exception as e:  # no symmetric-memory support in this build / on this node
                self.dp_comm, self.dp_comm_note, galloc = "nccl", "peer memory unavailable: %r" % (e,), None
        self.pack = bparams.ParamPack({k: v.to(dev) for k, v in state_dict.items() if v.dtype == torch.float32
                                       and "running_" not in k and not k.endswith("melmat")}, grad_alloc=galloc)
        n = self.pack.total
        self.m = torch.zeros(n, device=dev)       # Adam moments over the flat fp32 parameter buffer
        self.v = torch.zeros(n, device=dev)
        # (sum of squares, grad norm, update-applied flag, number of applied updates): the schedule's step lives on the device
        self.state = torch.zeros(4, device=dev, dtype=torch.float64)
        self.mel = mel_matrix(device=dev)
        self.bn: Dict[str, torch.Tensor] = {}
        for nblk in range(cfg.enc.blocks):
            p = f"encoder.encoders.{nblk}.conv_module.norm."
            rm, rv, nb = state_dict.get(p + "running_mean"), state_dict.get(p + "running_var"), state_dict.get(p + "num_batches_tracked")
            self.bn[p + "running_mean"] = rm.to(dev).float().clone() if rm is not None else torch.zeros(cfg.enc.d, device=dev)
            self.bn[p + "running_var"] = rv.to(dev).float().clone() if rv is not None else torch.ones(cfg.enc.d, device=dev)
            self.bn[p + "num_batches_tracked"] = nb.to(dev).long().reshape(1).clone() if nb is not None else \
                torch.zeros(1, device=dev, dtype=torch.long)
        self.step_no = 0
        self.pack.refresh()  # the bf16 operand copies; from here on the optimizer kernel keeps them current
        # with self-conditioning the encoder backward still adds to ctc.ctc_lo's gradient after the heads are done
        # (step.encoder_bwd): those parameters then belong to the last bucket, not to "heads"
        late = ("ctc.",) if cfg.enc.interctc_use_conditioning else ()
        self.reducer = None
        if self.dp_comm == "peer":
            try:
                self.reducer = PeerReducer(self.pack, world, overlap=(dp_mode != "graph_serial"), late_prefixes=late)
            except Exception as e:  # the rendezvous is collective and fails on every rank alike (no peer access, no fabric)
                self.dp_comm, self.dp_comm_note = "nccl", "peer memory unavailable: %r" % (e,)
        if self.reducer is None:
            self.reducer = GradReducer(self.pack, world, overlap=(dp_mode != "graph_serial"), late_prefixes=late)
        # The peer-memory exchange is made of plain stream work (flag kernels, DMA copies, one add kernel), so it is captured
        # WITH the step: one graph per batch shape as on a single GPU, the exchange a parallel branch of it, clip + Adam at
        # its end.  (B200SLU_DP_PEER_CHAIN=1: the chain of graphs with the exchange issued between replays, as with NCCL.)
        self.peer_one_graph = (self.dp_comm == "peer" and dp_mode == "graph"
                               and os.environ.get("B200SLU_DP_PEER_CHAIN", "0") != "1")
        # SMs left to the compute kernels while a gradient bucket is being all-reduced (0 = all; tuning: B200SLU_DP_SM_BUDGET)
        self.sm_budget = int(os.environ.get("B200SLU_DP_SM_BUDGET", "0"))
        self.buckets: "OrderedDict[tuple, _Bucket]" = OrderedDict()
        # per-step device record {u64 dropout seed; f32 lr, bias correction 1, bias correction 2, pad}: a ring of pinned
        # host slots so that the asynchronous copy of step s is never overwritten by the host preparing step s + 1
        self.sp_host = torch.zeros(64, 24, dtype=torch.uint8).pin_memory()
        self.sp_np = self.sp_host.numpy().view(np.dtype([("seed", "<u8"), ("lr", "<f4"), ("bc1", "<f4"), ("bc2", "<f4"),
                                                         ("pad", "<f4")]))
        self.sp_events: list = [None] * 64  # per slot: "the H2D copy out of this slot has executed"
        self.sp_dev = torch.zeros(24, dtype=torch.uint8, device=dev)
        ops.set_step_params(self.sp_dev)
        self._closed = False
        self.loss_host = torch.zeros(4, dtype=torch.float32).pin_memory()
        self._prefetched, self._h2d = None, None
        self.last: Optional[_Bucket] = None

    # warmup_lr.py:43-50
    def lr(self) -> float:
        s, w = self.step_no, self.warmup_steps
        return self.base_lr * w ** 0.5 * min(s ** -0.5, s * w ** -1.5)

    def close(self):
        """Unregister this step's device record from the library (the kernels of OTHER users of the library -- the drop-in
        modules with dropout on -- would otherwise keep reading a pointer that goes back to the allocator with this
        object).  Called by __del__; idempotent."""
        if not self._closed:
            self._closed = True
            try:
                with torch.cuda.device(self.dev):
                    ops.set_step_params(None)
            except Exception:  # interpreter shutdown: the library may be gone already
                pass

    def __del__(self):
        self.close()

    # ---- the step body: only device-resident inputs (the bucket's static buffers, the 24-byte record) -> capturable
    def _body(self, b: _Bucket, overlap: bool, on_ready=None):
        pack, cfg, inp = self.pack, self.cfg, b.inp
        # flat16 is kept current by the optimizer kernel (_tail); only the two re-laid-out weights are rebuilt per step
        pack.refresh(cast=False)
        pack.zero_grad()
        draws = {k: inp[k] for k in _DRAW_KEYS if k in inp}
        feats, flens = bstep.frontend_fwd(inp["wav"], inp["lens"], self.mel, cfg, draws)
        # trainer.py:594-606: the loss is the batch-weighted mean over ranks and DDP averages the gradients: the summed
        # gradients are scaled by 1 / world inside the optimizer kernel
        out = bstep.train_step_from_feats(feats, flens, inp["text"], inp["tl"], pack.P, pack.W, pack.G, cfg,
                                          bstep.StepCtx(1000003 * (self.rank + 1), True), bn_buffers=self.bn,
                                          on_ready=on_ready if on_ready is not None else
                                          (self.reducer.ready if (overlap and self.world > 1) else None))
        pack.finalize_grads()
        b.res = torch.cat([out["loss"], out["loss_ctc"], out["loss_att"], out["acc"]])

    def _tail(self):
        self.reducer.finish()  # issues whatever gradient buckets are still pending, then joins the side stream
        # WarmupLR and Adam's bias corrections are evaluated in the kernel from the device-side count of APPLIED updates: a
        # step skipped for a non-finite gradient norm advances neither (trainer.py:651-690), and nothing here is a host value
        ops.adam_step_sched(self.pack.flat32, self.pack.gflat, self.m, self.v, self.base_lr, self.warmup_steps, self.state,
                            grad_inv_scale=1.0 / self.world, max_norm=self.max_norm, params_bf16=self.pack.flat16)

    def capture(self, b: Optional[_Bucket] = None):
        """Record the whole step of this bucket (weight cast -> front-end -> forward -> backward [-> clip + Adam when
        single-GPU or when the gradient exchange is captured with it]) as ONE CUDA graph.  Called automatically by __call__; public for benchmarks that want control."""
        b = b if b is not None else self.last
        ops.set_step_params(self.sp_dev)
        torch.cuda.synchronize()
        self.graph_has_tail = self.world == 1 or self.peer_one_graph
        cuts = self._cut_stages()
        if not cuts:
            g = torch.cuda.CUDAGraph()
            self.reducer.begin()
            with torch.cuda.graph(g):
                self._body(b, overlap=self.peer_one_graph)
                if self.graph_has_tail:
                    self._tail()
            b.graph = [(g, None)]
            return
        # a chain of graphs sharing one memory pool, replayed in capture order; a cut sits where step.py reports a
        # gradient bucket final (on_ready), so the host can issue that bucket's all-reduce between two replays
        import gc
        self.graph_has_tail = False  # clip + Adam follow the last segment eagerly (after the last bucket's all-reduce)
        gc.collect()
        torch.cuda.empty_cache()
        chain, cur = [], [torch.cuda.CUDAGraph()]
        side = torch.cuda.Stream(device=self.dev)
        side.wait_stream(torch.cuda.current_stream())

        def cut(stage):
            if stage in cuts:
                cur[0].capture_end()
                chain.append((cur[0], stage))
                cur[0] = torch.cuda.CUDAGraph()
                cur[0].capture_begin(pool=chain[0][0].pool())
                # from here on a bucket's all-reduce runs beside the kernels: size the persistent grids for the SMs NCCL
                # leaves (a grid of all SMs would wait for NCCL's CTAs and, statically partitioned, end that much later)
                ops.set_sm_budget(self.sm_budget)

        with torch.cuda.stream(side):
            cur[0].capture_begin()
            self._body(b, overlap=False, on_ready=cut)
            cur[0].capture_end()
            chain.append((cur[0], None))
        ops.set_sm_budget(0)
        torch.cuda.current_stream().wait_stream(side)
        b.graph = chain

    def _cut_stages(self):
        """Stages after which the captured step is cut (dp_mode "graph", more than one rank): every bucket of the reducer's
        plan except the last one, which is reduced after the last segment anyway."""
        if self.world == 1 or self.dp_mode != "graph" or self.peer_one_graph:
            return set(getattr(self, "force_cuts", ()))  # tests cut a single-GPU step to check the chain replays identically
        return {s for s, _, _ in self.reducer.stages if s != "front"}

    def drop_graphs(self):
        """Back to eager launches (profiling passes bracket every kernel with events)."""
        for b in self.buckets.values():
            b.graph = None
        self.use_graph = False

    def _bucket(self, batch: Dict[str, torch.Tensor]) -> _Bucket:
        key = tuple((k, tuple(v.shape), v.dtype) for k, v in sorted(batch.items()))
        b = self.buckets.get(key)
        if b is None:
            while len(self.buckets) >= self.max_buckets:  # least recently used shape goes, with its graph's memory pool
                self.buckets.popitem(last=False)
            b = self.buckets[key] = _Bucket(batch, self.dev)
        else:
            self.buckets.move_to_end(key)
        return b

    def __call__(self, batch: Dict[str, torch.Tensor], fetch_loss: bool = False) -> torch.Tensor:
        """One optimizer step on ``batch``.  Returns the device tensor (loss, loss_ctc, loss_att, acc); with fetch_loss it
        is also copied (asynchronously) into the pinned ``self.loss_host``.  No host synchronisation."""
        for k in ("wav", "lens", "text", "tl"):
            assert k in batch, f"TrainStep: batch lacks {k!r}"
        b = self.last = self._bucket(batch)
        ops.set_step_params(self.sp_dev)  # host-side registration, handed to the kernels as an argument (baked into a graph
        self.step_no += 1                 # at capture): re-registered per call so that several TrainSteps can coexist
        slot = self.step_no % 64
        if self.sp_events[slot] is not None:
            self.sp_events[slot].synchronize()  # the host never runs more than 64 steps ahead of the copies out of the ring
        self.sp_np[slot] = (self.step_no * 0x9E3779B97F4A7C15 % (1 << 64), self.lr(), 1 - 0.9 ** self.step_no,
                            1 - 0.999 ** self.step_no, 0.0)
        self.sp_dev.copy_(self.sp_host[slot], non_blocking=True)
        if self.sp_events[slot] is None:
            self.sp_events[slot] = torch.cuda.Event()
        self.sp_events[slot].record()
        pf = self._prefetched
        if pf is not None and pf[0] is batch and pf[1] is b:
            # the copy stream brought this batch over while the previous step ran: one device-to-device pass into the
            # graph's input buffers (41 MB: ~12 us) instead of the H2D copy on the critical path
            _, _, i, ready = pf
            torch.cuda.current_stream().wait_event(ready)
            for k in batch:
                b.inp[k].copy_(b.staging[i][k], non_blocking=True)
            if b.free[i] is None:
                b.free[i] = torch.cuda.Event()
            b.free[i].record()
        else:
            for k, v in batch.items():
                b.inp[k].copy_(v, non_blocking=True)
        self._prefetched = None
        self.reducer.begin()
        if b.graph is not None:
            for g, stage in b.graph:
                g.replay()
                if stage is not None:
                    self.reducer.ready(stage)  # this bucket's all-reduce runs under the next segment
            if not self.graph_has_tail:
                self._tail()
            res = b.res
        else:
            self._body(b, overlap=self.dp_mode != "graph")
            self._tail()
            res = b.res  # this step's result; a capture below re-points b.res at the graph's own output buffer
            b.eager_steps += 1
            if self.use_graph and b.eager_steps >= self.eager_before:
                self.capture(b)
        if fetch_loss:
            self.loss_host.copy_(res, non_blocking=True)
        return res

    def prefetch(self, batch: Dict[str, torch.Tensor]) -> torch.cuda.Event:
        """Start the host-to-device copy of the NEXT step's batch (pinned host tensors) on a copy stream, so that it runs
        under the current step's kernels; the next __call__ with this same dict picks the staged copy up.  Returns the
        event after which the host buffers may be rewritten.  What a data loader's pinned-memory prefetch does in the
        reference's loop (`trainer.py:464-475` iterates a DataLoader with `pin_memory`)."""
        b = self._bucket(batch)
        if self._h2d is None:
            self._h2d = torch.cuda.Stream(device=self.dev)
        i = b.next_set
        b.next_set ^= 1
        if b.staging[i] is None:
            b.staging[i] = {k: torch.empty(v.shape, dtype=v.dtype, device=self.dev) for k, v in batch.items()}
        if b.free[i] is not None:
            self._h2d.wait_event(b.free[i])  # the step that last used this staging set has copied it out
        with torch.cuda.stream(self._h2d):
            for k, v in batch.items():
                b.staging[i][k].copy_(v, non_blocking=True)
            ready = torch.cuda.Event()
            ready.record()
        self._prefetched = (batch, b, i, ready)
        return ready

    def optimizer_state_dict(self) -> Dict[str, Dict[str, torch.Tensor]]:
        """Adam's state per parameter NAME in torch.optim.Adam's vocabulary (step, exp_avg, exp_avg_sq, reference shapes): what
        the reference checkpoints under `optimizers` once mapped onto its parameter order (`names` below gives ours)."""
        step = self.state[3:4].to(torch.float32).clone()
        out = {}
        for n in self.pack.names:
            o, num = self.pack.offs[n], int(np.prod(self.pack.shapes[n], dtype=np.int64))
            out[n] = {"step": step.clone(), "exp_avg": self.m[o:o + num].view(self.pack.shapes[n]).clone(),
                      "exp_avg_sq": self.v[o:o + num].view(self.pack.shapes[n]).clone()}
        return out

    def load_optimizer_state_dict(self, sd: Dict[str, Dict[str, torch.Tensor]]):
        steps = set()
        for n, st in sd.items():
            o, num = self.pack.offs[n], int(np.prod(self.pack.shapes[n], dtype=np.int64))
            self.m[o:o + num].copy_(st["exp_avg"].reshape(-1))
            self.v[o:o + num].copy_(st["exp_avg_sq"].reshape(-1))
            steps.add(float(st["step"]))
        assert len(steps) == 1, "one optimizer step count for all parameters"
        self.state[3] = steps.pop()

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """Parameters in the reference's layout and names, plus the BatchNorm buffers (loadable by the reference model)."""
        out = {n: self.pack.P[n].detach().clone() for n in self.pack.names}
        out.update({k: v.clone() for k, v in self.bn.items()})
        return out
