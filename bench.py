#!/usr/bin/env python
"""Train-step throughput of the SLURP Conformer CTC/attention model on B200 (BASELINE.json metric:
audio-sec/sec train throughput; config[1] = egs2/slurp Conformer, 12 blocks, CTC 0.3 + attention decoder).

    python bench.py --gpus 1 --steps K --warmup W                 # one process (default workload: slurp)
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...                          # the reference's CPU path (oracle port) on the host cores
    python bench.py --workload {slurp,slurp_legacy,long15s,stress1500,libri17,ctc_sweep}   # BASELINE configs 2-5

One "step" = waveform batch -> log-mel front-end -> SpecAug -> utterance MVN -> 12-block Conformer encoder -> CTC +
6-block attention decoder losses -> full backward -> (N > 1: NCCL gradient all-reduce overlapped with backward) ->
fused clip + Adam.  Dropout 0.1 and SpecAug are ON (training configuration).  Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FS = 16000

# BASELINE.json configs.  "slurp" is the line the metric is quoted on (config 2, rel-pos "latest" = the
# egs2/slurp_entity YAML); "slurp_legacy" = the egs2/slurp YAML, which leaves rel_pos_type at its default "legacy".
WORKLOADS = {
    "slurp": dict(batch=128, min_sec=3.0, max_sec=5.0, vocab=600, rel_pos="latest", enc_blocks=12, dec_blocks=6, hop=128,
                  tok=(6, 20), label="egs2/slurp Conformer SLU (config 2)"),
    "slurp_legacy": dict(batch=128, min_sec=3.0, max_sec=5.0, vocab=600, rel_pos="legacy", enc_blocks=12, dec_blocks=6, hop=128,
                         tok=(6, 20), label="egs2/slurp Conformer SLU, legacy rel-pos as the egs2/slurp YAML leaves it (config 2')"),
    "long15s": dict(batch=32, min_sec=12.0, max_sec=15.0, vocab=600, rel_pos="latest", enc_blocks=12, dec_blocks=6, hop=128,
                    tok=(20, 60), label="long utterances U[12,15] s, T = 374..468 (config 3)"),
    "stress1500": dict(batch=4, min_sec=47.9, max_sec=48.0, vocab=600, rel_pos="latest", enc_blocks=12, dec_blocks=6, hop=128,
                       tok=(60, 160), label="T = 1500 after subsampling (48 s of audio; config 3 stress point, synthetic only)"),
    "libri17": dict(batch=80, min_sec=10.0, max_sec=16.0, vocab=5000, rel_pos="latest", enc_blocks=17, dec_blocks=6, hop=160,
                    tok=(20, 80), label="LibriSpeech-scale Conformer, 17 blocks, V = 5000, hop 160 (config 4)"),
    "ctc_sweep": dict(label="standalone CTC fwd+bwd sweep, B = 64 (config 5)"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="slurp", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=None, help="utterances per GPU per step (SURVEY 8d sweep: 8, 32, 128)")
    ap.add_argument("--vocab", type=int, default=None)
    ap.add_argument("--rel-pos", default=None, choices=["latest", "legacy"])
    ap.add_argument("--min-sec", type=float, default=None)
    ap.add_argument("--max-sec", type=float, default=None)
    ap.add_argument("--enc-blocks", type=int, default=None)
    ap.add_argument("--dec-blocks", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-eager-baseline", action="store_true")
    ap.add_argument("--no-ctc-record", action="store_true")
    ap.add_argument("--no-prof", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel eagerly instead of replaying a CUDA graph")
    ap.add_argument("--cpu-batch", type=int, default=8)
    ap.add_argument("--dp-mode", default="graph", choices=["graph", "graph_serial", "eager_overlap"],
                    help="N > 1: how the gradient all-reduce meets the CUDA graph (see espnet_slurp_b200/trainer.py)")
    ap.add_argument("--dp-comm", default=None, choices=["peer", "nccl"],
                    help="N > 1: gradient buckets over peer memory by the copy engines (default) or ncclAllReduce")
    a = ap.parse_args()
    w = WORKLOADS[a.workload]
    for k in ("batch", "vocab", "rel_pos", "min_sec", "max_sec", "enc_blocks", "dec_blocks"):
        if getattr(a, k) is None:
            setattr(a, k, w.get(k))
    a.hop = w.get("hop", 128)
    a.tok = w.get("tok", (6, 20))
    a.label = w["label"]
    return a


# ----------------------------------------------------------------------------------------------------- synthetic workload
def make_batch(B, V, seed, min_sec, max_sec, tok=(6, 20)):
    """Waveforms ~N(0, 0.1^2), durations U[min,max] s, padded to the batch maximum; tok[0]..tok[1] tokens in 2..V-2."""
    g = torch.Generator().manual_seed(seed)
    lens = (torch.rand(B, generator=g) * (max_sec - min_sec) + min_sec).mul(FS).long()
    lens[0] = int(max_sec * FS)  # fixed padded width -> static shapes from step to step
    N = int(lens.max())
    wav = torch.randn(B, N, generator=g) * 0.1
    wav *= (torch.arange(N)[None, :] < lens[:, None]).float()
    tl = torch.randint(tok[0], tok[1] + 1, (B,), generator=g)
    Lmax = tok[1]
    text = torch.randint(2, V - 1, (B, Lmax), generator=g)
    text = torch.where(torch.arange(Lmax)[None, :] < tl[:, None], text, torch.full_like(text, -1))
    return wav, lens, text, tl


def specaug_draws(flens, T_in, n_mels=80, W=5, F=30, Tm=40, n_f=2, n_t=2, gen=None):
    """The RNG draws SpecAug makes, in the reference's order (time_warp.py:28-29 per utterance, then
    mask_along_axis.py:33-43 for freq and time), as small int32 arrays for b200slu_specaug_mvn_fwd."""
    B = flens.numel()
    center = torch.full((B,), -1, dtype=torch.int32)
    warped = torch.zeros(B, dtype=torch.int32)
    fl = flens.tolist()
    for b in range(B):
        t = int(fl[b])
        if t - W > W:
            c = int(torch.randint(W, t - W, (1,), generator=gen))
            center[b] = c
            warped[b] = int(torch.randint(c - W, c + W, (1,), generator=gen)) + 1
    f_len = torch.randint(0, F, (B, n_f), generator=gen)
    f_pos = torch.randint(0, max(1, n_mels - int(f_len.max())), (B, n_f), generator=gen)
    t_len = torch.randint(0, Tm, (B, n_t), generator=gen)
    t_pos = torch.randint(0, max(1, T_in - int(t_len.max())), (B, n_t), generator=gen)
    return {"center": center, "warped": warped, "fmask": torch.stack([f_pos, f_len], -1).int().contiguous(),
            "tmask": torch.stack([t_pos, t_len], -1).int().contiguous()}


def build_cfg(args):
    from espnet_slurp_b200 import step as bstep
    enc = bstep.EncCfg(80, 512, 8, 2048, args.enc_blocks, 31, args.rel_pos, 0.1, 0.1, 0.1)
    dec = bstep.DecCfg(args.vocab, 512, 8, 2048, args.dec_blocks, 0.1, 0.1, 0.0, 0.0)
    return bstep.ModelCfg(enc=enc, dec=dec, vocab=args.vocab, hop=args.hop)


def make_state_dict(cfg, seed=0):
    """Random-init weights of the architecture under the reference's state_dict names (no checkpoints offline)."""
    from espnet_slurp_b200.testing import make_sd
    return make_sd(cfg.enc, cfg.dec, cfg.vocab, seed=seed)


def workload_text(args, cfg):
    return (f"{args.label}: {cfg.enc.blocks} enc blocks d=512 H=8 ffn=2048 K=31 rel-pos {cfg.enc.rel_pos_type}, "
            f"{cfg.dec.blocks} dec blocks, V={cfg.vocab}, ctc 0.3, lsm 0.1, dropout 0.1, SpecAug on, raw 16 kHz waveforms "
            f"U[{args.min_sec},{args.max_sec}] s (hop {args.hop}), {args.batch} utt/GPU")


# ----------------------------------------------------------------------------------------------------- clocks sampler
class Clocks(threading.Thread):
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": max((int(r[1]) for r in self.rows if r[1].isdigit()), default=None), "reasons": reasons,
                "samples": len(self.rows)}


# ----------------------------------------------------------------------------------------------------- CPU reference arm
def _oracle_cfg(cfg):
    from oracle import model as om
    return om.ModelCfg(enc=om.EncCfg(80, 512, 8, 2048, cfg.enc.blocks, 31, cfg.enc.rel_pos_type),
                       dec=om.DecCfg(cfg.vocab, 512, 8, 2048, cfg.dec.blocks), vocab=cfg.vocab, hop=cfg.hop)


def _oracle_draws(draws, dev="cpu"):
    return {"center": draws["center"].long().to(dev), "warped": draws["warped"].long().to(dev),
            "f_pos": draws["fmask"][..., 0].long().to(dev), "f_len": draws["fmask"][..., 1].long().to(dev),
            "t_pos": draws["tmask"][..., 0].long().to(dev), "t_len": draws["tmask"][..., 1].long().to(dev)}


def cpu_step_time(args, cfg, n_steps, warmup, batch, min_seconds=0.0):
    """The reference's CPU path for the same step: oracle/model.py (CPU restatement pinned against the unmodified
    reference's goldens) fwd + autograd bwd in fp32 on all host cores.  Returns (audio-sec/sec, seconds/step, cores)."""
    from oracle import model as om
    torch.set_num_threads(os.cpu_count() or 1)
    ocfg = _oracle_cfg(cfg)
    sd = {k: v.clone().requires_grad_(True) for k, v in make_state_dict(cfg).items()}
    mel = torch.from_numpy(om.slaney_mel_matrix())
    wav, lens, text, tl = make_batch(batch, cfg.vocab, 1234, args.min_sec, args.max_sec, args.tok)
    audio = float(lens.sum()) / FS
    times = []
    i = -1
    while True:
        i += 1
        # n_steps timed steps, and then more until the timed sample is at least min_seconds long (bounded)
        if i >= warmup + n_steps and (sum(times) >= min_seconds or len(times) >= 40):
            break
        t0 = time.perf_counter()
        feats, flens = om.frontend_logmel(wav, lens, mel, cfg.n_fft, cfg.hop)
        feats = om.utterance_mvn(om.spec_augment(feats, flens, _oracle_draws(specaug_draws(flens, feats.shape[1]))), flens)
        loss, _ = om.asr_forward_from_feats(feats, flens, text, tl, sd, ocfg)
        loss.backward()
        for v in sd.values():
            v.grad = None
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    sec = sum(times) / len(times)
    return audio / sec, sec, torch.get_num_threads(), audio, len(times)


# ----------------------------------------------------------------------------------------------------- same-box library bar
def gpu_eager_step_time(args, cfg, mode, steps=6, warmup=3):
    """The reference's own GPU path for this step on this box: eager PyTorch over cuBLAS / cuDNN / cuFFT / ATen
    (espnet2/train/trainer.py:554-556 `autocast` + stock nn.Module ops, :612-690 clip + Adam), stated by
    oracle/model.py with LIBRARY_OPS = True (the torch library op behind every reference module), dropout and SpecAug on,
    fp32 or under torch.autocast(bfloat16), same synthetic batches and batch size as the timed arm.  Returns ms/step."""
    from oracle import model as om
    dev = torch.device("cuda", torch.cuda.current_device())
    old = om.LIBRARY_OPS
    om.LIBRARY_OPS = True
    try:
        ocfg = _oracle_cfg(cfg)
        sd = {k: v.to(dev).requires_grad_(True) for k, v in make_state_dict(cfg).items()}
        opt = torch.optim.Adam(list(sd.values()), lr=1e-4)
        mel = torch.from_numpy(om.slaney_mel_matrix()).to(dev)
        batches = []
        for i in range(2):
            wav, lens, text, tl = make_batch(args.batch, cfg.vocab, 1234 + i, args.min_sec, args.max_sec, args.tok)
            batches.append((wav.to(dev), lens.to(dev), text.to(dev), tl.to(dev), lens // cfg.hop + 1))
        rates = {"att": cfg.enc.att_dropout, "dec_att": cfg.dec.self_att_dropout, "pos": cfg.enc.pos_dropout, "other": cfg.enc.dropout}

        def drop(x, tag):  # torch.nn.Dropout at every site of the reference graph
            if tag.endswith(".att"):
                p = rates["dec_att"] if tag.startswith("decoder.") else rates["att"]
            elif ".embed" in tag:
                p = rates["pos"]
            else:
                p = rates["other"]
            return torch.nn.functional.dropout(x, p, True) if p > 0 else x

        def one(i):
            wav, lens, text, tl, flens_host = batches[i % 2]
            draws = _oracle_draws(specaug_draws(flens_host, wav.shape[1] // cfg.hop + 1), dev)  # host RNG, as the reference
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=(mode == "autocast_bf16")):
                with torch.autocast("cuda", enabled=False):  # espnet_model.py:328-338
                    feats, flens = om.frontend_logmel(wav, lens, mel, cfg.n_fft, cfg.hop)
                    feats = om.utterance_mvn(om.spec_augment(feats, flens, draws), flens)
                loss, _ = om.asr_forward_from_feats(feats, flens, text, tl, sd, ocfg, drop=drop)
            loss.backward()
            torch.nn.utils.clip_grad_norm_(list(sd.values()), 5.0)
            opt.step()
            opt.zero_grad(set_to_none=True)

        for i in range(warmup):
            one(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            one(i)
        e1.record()
        torch.cuda.synchronize()
        audio = sum(float(batches[i % 2][1].sum()) / FS for i in range(steps))
        return e0.elapsed_time(e1) / steps, audio / (e0.elapsed_time(e1) * 1e-3)
    finally:
        om.LIBRARY_OPS = old
        del sd
        torch.cuda.empty_cache()


# ----------------------------------------------------------------------------------------------------- standalone CTC (config 5)
def ctc_point(B, T, V, L, peak_gbs, reps=5):
    """CTC fwd+bwd on logits [B,T,V] (log_softmax fused), L2 flushed between repetitions: CTC GB/s = 12*T*B*V / time."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from ctc_sweep import make
    from espnet_slurp_b200 import ops
    logits, hl, lab, ll = make(B, T, V, L)
    dl, dh, dlab, dll = logits.cuda(), hl.int().cuda(), lab.int().cuda(), ll.int().cuda()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def once():
        _, loss_sum, ws = ops.ctc_fwd(dl, V, dh, dlab, dll)
        ops.ctc_bwd(dl, V, dh, dlab, dll, ws, 1.0, want_f32=True, want_bf16=False)

    for _ in range(3):
        once()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        once()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    gbs = 12.0 * T * B * V / (ms * 1e-3) / 1e9
    return {"B": B, "T": T, "V": V, "L": L, "ms": round(ms, 4), "gbs": round(gbs, 1), "frac_hbm": round(gbs / peak_gbs, 3)}


# ----------------------------------------------------------------------------------------------------- GPU arm
class Trainer:
    """Synthetic batches around espnet_slurp_b200.trainer.TrainStep -- the package's public train-step API (forward,
    backward, gradient all-reduce, clip + Adam; one CUDA graph per batch shape).  See trainer.py for the step itself."""

    def __init__(self, args, cfg, rank, world):
        from espnet_slurp_b200.trainer import TrainStep
        self.cfg, self.args, self.rank, self.world = cfg, args, rank, world
        dev = torch.device("cuda", torch.cuda.current_device())
        # Graph capture is driven explicitly below (eager_steps_before_capture is never reached) so that the launches of
        # one eager step can be counted first.
        self.ts = TrainStep(cfg, make_state_dict(cfg), rank, world, dp_mode=args.dp_mode, use_graph=False,
                            dp_comm=args.dp_comm)
        if os.environ.get("B200SLU_BENCH_FORCE_CUTS"):  # tuning: the chain-of-graphs capture on one GPU (cost of the cuts alone)
            self.ts.force_cuts = tuple(s for s, _, _ in self.ts.reducer.stages if s != "front")
        self.host_batches, self.dev_batches, self.flens, self.done = [], [], [], []
        self.prepared = None
        self.gen = torch.Generator().manual_seed(99 + rank)
        for i in range(4):
            wav, lens, text, tl = make_batch(args.batch, cfg.vocab, 1234 + rank * 100 + i, args.min_sec, args.max_sec, args.tok)
            flens = lens // cfg.hop + 1
            draws = specaug_draws(flens, wav.shape[1] // cfg.hop + 1, gen=self.gen)
            hb = {"wav": wav.pin_memory(), "lens": lens.int().pin_memory(), "text": text.int().pin_memory(),
                  "tl": tl.int().pin_memory(), **{k: v.pin_memory() for k, v in draws.items()}}
            self.host_batches.append(hb)
            self.flens.append(flens)
            self.done.append(None)
            self.dev_batches.append({k: v.to(dev) for k, v in hb.items()})
        self.audio_sec = [float(hb["lens"].sum()) / FS for hb in self.host_batches]
        self.h2d_bytes = sum(v.numel() * v.element_size() for v in self.host_batches[0].values())

    @property
    def loss_host(self):
        return self.ts.loss_host

    def capture(self):
        self.ts.capture()

    def drop_graphs(self):
        self.ts.drop_graphs()

    def _host_prepare(self, j):
        hb = self.host_batches[j]
        if self.done[j] is not None:
            self.done[j].synchronize()  # the previous H2D copy out of these pinned buffers has finished
        for k, v in specaug_draws(self.flens[j], hb["wav"].shape[1] // self.cfg.hop + 1, gen=self.gen).items():
            hb[k].copy_(v)
        return hb

    def step(self, i, e2e=False, more=False):
        # e2e: what a training loop pays per step -- the SpecAug draws are made on the host (128 per-utterance randint pairs,
        # time_warp.py:28-29, + the mask draws) inside the timed region, the inputs come from pinned HOST memory (H2D inside
        # the timed region) and the losses are read back (D2H)
        # (D2H).  Like a DataLoader with pinned memory, the NEXT step's batch is prepared and its H2D copy started
        # (TrainStep.prefetch, copy stream) while this step runs; every step's draws and copies are inside the timed region
        # (the first step of a region copies inline, the last one prefetches nothing).
        if e2e:
            j = i % 4
            if self.prepared != i:
                hb = self._host_prepare(j)
            else:
                hb = self.host_batches[j]
            self.ts(hb, fetch_loss=True)
            if self.prepared != i:
                self.done[j] = torch.cuda.Event()
                self.done[j].record()
            self.prepared = None
            if more:
                jn = (i + 1) % 4
                self.done[jn] = self.ts.prefetch(self._host_prepare(jn))
                self.prepared = i + 1
        else:
            self.ts(self.dev_batches[i % 4], fetch_loss=False)


def run_ours(args):
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    # stdout carries exactly one JSON line: anything libraries write to fd 1 while the job runs (NCCL prints its version
    # banner there from C, whatever NCCL_DEBUG_FILE says) is sent to stderr; fd 1 is restored just before the line
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from espnet_slurp_b200 import _lib
    lib = _lib.lib()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    if args.workload == "ctc_sweep":
        return run_ctc_sweep(args, rank, world, saved_stdout, hbm_peak, peaks)
    cfg = build_cfg(args)
    tr = Trainer(args, cfg, rank, world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(n, e2e):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last = None
        for i in range(n):
            last = tr.step(i, e2e=e2e, more=i + 1 < n)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item(), last

    for i in range(2):
        tr.step(i)  # eager: one-time kernel attribute / driver-entry-point setup happens outside any capture
    torch.cuda.synchronize()
    lib.b200slu_prof_reset()
    tr.step(2)
    tr.launches_per_step = lib.b200slu_prof_total_launches()  # counted on an eager step; a graph replay issues the same kernels
    if not args.no_graph and not (world > 1 and args.dp_mode == "eager_overlap"):
        tr.capture()
    for i in range(args.warmup):
        tr.step(i)
    clocks = Clocks(local)
    clocks.start()
    lib.b200slu_prof_reset()
    ms, last = timed(args.steps, e2e=False)
    launches = lib.b200slu_prof_total_launches() if args.no_graph else tr.launches_per_step * args.steps
    clk = clocks.stop()
    ms_e2e, last2 = timed(args.steps, e2e=True)
    torch.cuda.synchronize()
    loss_val = float(tr.loss_host[0])
    audio = sum(tr.audio_sec[i % 4] for i in range(args.steps))
    tot = torch.tensor([audio], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot)
    total_audio = tot.item()

    # per-kernel-family device time (separate pass: CUDA events around every launch, on the launching stream)
    fams = {}
    if not args.no_prof:
        import ctypes as C
        lib.b200slu_prof_reset()
        lib.b200slu_prof_enable_timing(1)
        nprof = 2
        tr.drop_graphs()  # eager launches so that every kernel can be bracketed by events
        for i in range(nprof):
            tr.step(i)
        torch.cuda.synchronize()
        lib.b200slu_prof_enable_timing(0)
        for f in range(lib.b200slu_prof_num_families()):
            o = (C.c_double * 4)()
            lib.b200slu_prof_get(f, o)
            if o[0] > 0:
                fams[lib.b200slu_prof_family_name(f).decode()] = {"launches_per_step": o[0] / nprof, "ms_per_step": o[1] / nprof,
                                                                 "flops_per_step": o[2] / nprof, "bytes_per_step": o[3] / nprof}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    tf_peak = peaks.get("bf16_tflops_sustained", 1400.0)  # kernels are timed inside a long step -> sustained figure
    roof = None
    traffic = None
    for name in ("r2c_step_ncu_summary.json", "r2b_step_ncu_summary.json", "r2_step_ncu_summary.json", "r1_step_ncu_summary.json"):
        try:  # DRAM bytes per GEMM launch from the committed ncu capture of this command (tools/ncu_step_summary.py)
            tj = json.load(open(os.path.join(ROOT, "profiles", name)))
            if tj.get("batch_per_gpu") == args.batch and args.workload == "slurp":
                traffic = tj["gemm"]["dram_bytes_per_launch"]
                break
        except Exception:
            pass
    if "gemm_tcgen05" in fams and fams["gemm_tcgen05"]["ms_per_step"] > 0:
        g = fams["gemm_tcgen05"]
        ach = g["flops_per_step"] / (g["ms_per_step"] * 1e-3) / 1e12
        roof = {"kernel": "gemm_bf16_tcgen05 (all GEMM launches of one step)", "bound": "tensor", "achieved": round(ach, 1),
                "peak": tf_peak, "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if peaks else "fallback",
                "unit": "TFLOP/s", "frac": round(ach / tf_peak, 4), "traffic": traffic,
                "algorithmic_bytes_per_launch": round(g["bytes_per_step"] / g["launches_per_step"], 1),
                "launches_per_step": g["launches_per_step"], "avg_launch_us": round(1e3 * g["ms_per_step"] / g["launches_per_step"], 2),
                "share_of_step_device_time": round(g["ms_per_step"] / sum(f["ms_per_step"] for f in fams.values()), 3)}
    total_flops = sum(f["flops_per_step"] for f in fams.values())
    line = {
        "metric": "audio-sec/sec train throughput, SLURP Conformer", "value": round(total_audio / (ms * 1e-3), 1),
        "unit": "audio-sec/sec", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_text(args, cfg) + "; fwd+bwd+grad all-reduce+clip+Adam"
                               + ("" if args.no_graph else "; whole step replayed as CUDA graph(s)"),
                   "name": args.workload, "batch_per_gpu": args.batch, "global_batch": args.batch * world,
                   "parallelism": f"dp{world}", "dp_mode": args.dp_mode if world > 1 else None,
                   "dp_comm": ({"peer": "gradient buckets over NVLink peer memory by the copy engines (dist.PeerReducer)",
                                "nccl": "ncclAllReduce per bucket (dist.GradReducer)"}[tr.ts.dp_comm]
                               + ("; " + tr.ts.dp_comm_note if getattr(tr.ts, "dp_comm_note", None) else "")) if world > 1 else None,
                   "l2": "per-step working set (activations + 1.3 GB of weights/grads/moments) >> 126 MB L2; 4 rotating batches"},
        "e2e": {"value": round(total_audio / (ms_e2e * 1e-3), 1), "unit": "audio-sec/sec", "ms_per_step": round(ms_e2e / args.steps, 3),
                "h2d_bytes_per_step": tr.h2d_bytes, "d2h_bytes_per_step": 16,
                "host_work_in_timed_region": "SpecAug RNG draws (per-utterance randint pairs + mask draws), H2D of waveforms / "
                                             "lengths / tokens / draws from pinned memory, D2H of the 4 loss scalars"},
        "gpu_launches": int(launches), "gpu_launches_per_step": round(launches / args.steps, 1),
        "clocks": clk, "loss": loss_val, "roofline": roof,
        "step_tflops": round(total_flops / (ms / args.steps * 1e-3) / 1e12, 1) if total_flops else None,
        "families": fams,
    }
    del tr
    torch.cuda.empty_cache()
    if not args.no_ctc_record and world == 1:
        # the "CTC GB/s" half of the metric (BASELINE config 5): the model's own shape and two sweep corners
        T_enc = ((int(args.max_sec * FS) // cfg.hop + 1 - 1) // 2 - 1) // 2
        try:
            line["ctc"] = {"unit": "GB/s at 12*T*B*V bytes (fwd+bwd, log_softmax fused, fp32 gradient out)", "peak_hbm_gbs": hbm_peak,
                           "points": [ctc_point(args.batch, T_enc, cfg.vocab, args.tok[1], hbm_peak),
                                      ctc_point(64, 1000, 5000, 50, hbm_peak), ctc_point(64, 1000, 1000, 100, hbm_peak)]}
        except Exception as e:  # never lose the headline line to the side record
            line["ctc"] = {"error": repr(e)[:200]}
    if not args.no_gpu_eager_baseline and world == 1:
        # same-box bar (SURVEY 8d "Also reported"): the reference's eager PyTorch step on this GPU, library kernels
        geb = {"what": "reference's eager PyTorch train step on this B200 (oracle/model.py with LIBRARY_OPS: cuBLAS / cuDNN / "
                       "cuFFT / ATen ops, dropout + SpecAug on, fwd + bwd + clip_grad_norm_ + torch.optim.Adam), same batches",
               "batch_per_gpu": args.batch, "unit": "audio-sec/sec"}
        for mode in ("fp32", "autocast_bf16"):
            try:
                ms_eager, v = gpu_eager_step_time(args, cfg, mode)
                geb[mode] = {"value": round(v, 1), "ms_per_step": round(ms_eager, 2)}
            except Exception as e:
                geb[mode] = {"error": repr(e)[:200]}
        line["gpu_eager_baseline"] = geb
    if not args.no_cpu_baseline:
        v, sec, cores, audio_cpu, nt = cpu_step_time(args, cfg, 2, 1, args.cpu_batch, min_seconds=12.0)
        line["cpu_baseline"] = {"value": round(v, 2), "unit": "audio-sec/sec", "cores": cores, "kind": "port",
                                "sample": f"1 warm-up + {nt} timed fwd+bwd steps ({nt * sec:.1f} s of CPU work) of the same model on "
                                          f"{args.cpu_batch} utterances per step ({audio_cpu:.1f} audio-s), oracle/model.py fp32 "
                                          f"on {cores} torch threads, {sec:.2f} s/step"}
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_ctc_sweep(args, rank, world, saved_stdout, hbm_peak, peaks):
    """BASELINE config 5 as a driver-runnable line: every feasible point of the grid; value = median CTC GB/s."""
    if rank != 0:
        return
    from espnet_slurp_b200 import _lib
    lib = _lib.lib()
    lib.b200slu_prof_reset()
    rows = []
    for V in (500, 1000, 2000, 5000):
        for T in (200, 500, 1000, 2000, 3000):
            for L in (10, 50, 100, 300):
                if 2 * L + 1 > T or T * V * 64 * 4 > (6 << 30):
                    continue
                rows.append(ctc_point(64, T, V, L, hbm_peak, reps=max(3, min(args.steps, 7))))
    gbs = sorted(r["gbs"] for r in rows)
    best = max(rows, key=lambda r: r["gbs"])
    line = {"metric": "CTC GB/s (standalone CTC fwd+bwd sweep)", "value": gbs[len(gbs) // 2], "unit": "GB/s", "n_gpus": 1,
            "steps": args.steps, "warmup": 3, "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOADS["ctc_sweep"]["label"] + ": V in {500,1000,2000,5000} x T in {200..3000} x label length "
                                   "{10,50,100,300}; value = median over the grid of 12*T*B*V bytes / (fwd+bwd time), L2 flushed "
                                   "between repetitions", "name": "ctc_sweep"},
            "roofline": {"kernel": "ctc_prep + ctc_lattice + ctc_grad (best grid point)", "bound": "hbm", "achieved": best["gbs"],
                         "peak": hbm_peak, "unit": "GB/s", "frac": round(best["gbs"] / hbm_peak, 3), "traffic": None,
                         "point": best, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback"},
            "gpu_launches": int(lib.b200slu_prof_total_launches()), "points": rows}
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    if args.workload == "ctc_sweep":
        print(json.dumps({"impl": "reference", "unavailable": "the ctc_sweep workload has no reference arm (tools/ctc_sweep.py "
                                                              "times torch's CPU ctc_loss beside the kernel)"}), flush=True)
        return
    cfg = build_cfg(args)
    # a FIXED sample per step (args.cpu_batch = 8 utterances, same seed every run): the line is reproducible from run to
    # run; cpu_step_time's untimed warm-up steps absorb thread-pool start-up and first-touch allocation
    b = args.cpu_batch
    v, sec, cores, audio, _ = cpu_step_time(args, cfg, args.steps, max(1, args.warmup), b)
    line = {"impl": "reference", "metric": "audio-sec/sec train throughput, SLURP Conformer", "value": round(v, 2),
            "unit": "audio-sec/sec", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(sec * 1e3, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"same model and synthetic waveforms as the GPU arm ({workload_text(args, cfg)}); each step = "
                                   f"{b} utterances ({audio:.1f} audio-s) fwd+bwd on the host cores (no optimizer); ONE host "
                                   "whatever --gpus says", "name": args.workload, "batch_per_step": b},
            "cpu_baseline": {"value": round(v, 2), "unit": "audio-sec/sec", "cores": cores, "kind": "port",
                             "sample": f"{args.steps} timed steps of {b} utterances, oracle/model.py (CPU restatement of the "
                                       "reference's PyTorch path, pinned to its goldens), fp32, torch CPU threads = cores"},
            "e2e": {"value": round(v, 2), "unit": "audio-sec/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
