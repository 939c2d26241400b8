"""BASELINE config 5: standalone CTC loss fwd+bwd sweep (B = 64; V 500..5000; T 200..3000; label length 10..300).

For every shape: seeded logits ~N(0,1) (the kernel fuses log_softmax), input lengths U[0.7T, T], labels U{1..V-1} with
20 % forced repeats, fp32 gradient output.  Reports time (CUDA-graph replay, inputs larger than L2 or L2 flushed
between replays), "CTC GB/s" = 12*T*B*V bytes / time (SURVEY 8d's algorithmic traffic: logits read in the forward, read again
and the gradient written in the backward) and its fraction of the measured HBM copy peak; on a few shapes also
torch.nn.functional.ctc_loss on the host cores and the parity of loss / gradient against it."""
import json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from espnet_slurp_b200 import ops

PEAK = 6553.6
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def make(B, T, V, L, seed=7):
    g = torch.Generator().manual_seed(seed)
    logits = torch.randn(B, T, ops.round8(V), generator=g)
    hl = (torch.rand(B, generator=g) * 0.3 + 0.7).mul(T).long().clamp(max=T)
    ll = torch.randint(max(1, L // 2), L + 1, (B,), generator=g)
    lab = torch.randint(1, V, (B, L), generator=g)
    rep = torch.rand(B, L, generator=g) < 0.2
    lab[:, 1:] = torch.where(rep[:, 1:], lab[:, :-1], lab[:, 1:])
    return logits, hl, lab, ll


def run(B, T, V, L, cpu_check=False):
    logits, hl, lab, ll = make(B, T, V, L)
    dl, dh, dlab, dll = logits.cuda(), hl.int().cuda(), lab.int().cuda(), ll.int().cuda()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def once():
        _, loss_sum, ws = ops.ctc_fwd(dl, V, dh, dlab, dll)
        g32, _ = ops.ctc_bwd(dl, V, dh, dlab, dll, ws, 1.0, want_f32=True, want_bf16=False)
        return loss_sum, g32

    for _ in range(3):
        loss_sum, g32 = once()
    torch.cuda.synchronize()
    ts, tf = [], []
    for _ in range(5):
        flush.zero_()                       # 256 MB > 126 MB L2
        e0, e1, em = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _, loss_sum, ws = ops.ctc_fwd(dl, V, dh, dlab, dll)
        em.record()
        g32, _ = ops.ctc_bwd(dl, V, dh, dlab, dll, ws, 1.0, want_f32=True, want_bf16=False)
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1)); tf.append(e0.elapsed_time(em))
    ms = sorted(ts)[len(ts) // 2]
    ms_fwd = sorted(tf)[len(tf) // 2]
    gbs = 12.0 * T * B * V / (ms * 1e-3) / 1e9
    row = {"B": B, "T": T, "V": V, "L": L, "ms": round(ms, 4), "ms_fwd": round(ms_fwd, 4), "ctc_GBps": round(gbs, 1), "frac_of_hbm_peak": round(gbs / PEAK, 3)}
    if cpu_check:
        tgt = torch.cat([lab[b, : ll[b]] for b in range(B)])
        x32 = logits[:, :, :V].clone().requires_grad_(True)      # the reference's own path: fp32 on the host cores (timed)
        t0 = time.perf_counter()
        ref32 = torch.nn.functional.ctc_loss(torch.log_softmax(x32, -1).transpose(0, 1), tgt, hl, ll, reduction="sum",
                                             zero_infinity=True)
        ref32.backward()
        cpu_ms = (time.perf_counter() - t0) * 1e3
        row["cpu_ms"] = round(cpu_ms, 1)
        row["cpu_cores"] = torch.get_num_threads()
        x = logits[:, :, :V].double().requires_grad_(True)        # fp64: the yardstick for the 1e-5 parity bar
        ref = torch.nn.functional.ctc_loss(torch.log_softmax(x, -1).transpose(0, 1), tgt, hl, ll, reduction="sum",
                                           zero_infinity=True)
        ref.backward()
        row["torch_fp32_grad_rel_err"] = float((x32.grad.double() - x.grad).abs().max() / x.grad.abs().max())
        row["loss_rel_err"] = float(abs(loss_sum.item() - ref.item()) / abs(ref.item()))
        gd = g32[:, :, :V].cpu().double()
        row["grad_rel_err"] = float((gd - x.grad).abs().max() / x.grad.abs().max())
    return row


if __name__ == "__main__":
    B = 64
    rows = []
    for V in (500, 1000, 2000, 5000):
        for T in (200, 500, 1000, 2000, 3000):
            for L in (10, 50, 100, 300):
                if 2 * L + 1 > T and not (T == 200 and L == 100):
                    continue  # infeasible for every utterance (one such combo is kept: exercises zero_infinity)
                if T * V * B * 4 > (6 << 30):
                    continue
                chk = (V, T, L) in ((500, 200, 10), (1000, 500, 50), (2000, 1000, 100), (5000, 1000, 300))
                r = run(B, T, V, L, cpu_check=chk)
                rows.append(r)
                print(json.dumps(r), flush=True)
    best = max(rows, key=lambda r: r["ctc_GBps"])
    print("# best", json.dumps(best))
