"""Fused attention kernels (csrc/attention.cu) timed alone against the unfused GEMM + softmax chain they replace.

python tools/attn_bench.py [--B 128] [--T 155] [--H 8] [--drop 0.1]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from espnet_slurp_b200 import ops as K  # noqa: E402


def timeit(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=128)
    ap.add_argument("--T", type=int, default=155)
    ap.add_argument("--H", type=int, default=8)
    ap.add_argument("--drop", type=float, default=0.1)
    ap.add_argument("--mode", default="latest")
    ap.add_argument("--bwd", action="store_true")
    ap.add_argument("--no-pd", action="store_true")
    a = ap.parse_args()
    B, T, H, d = a.B, a.T, a.H, a.H * 64
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    qkv = (torch.randn(B * T, 3 * d, device=dev, generator=g) * 0.5).bfloat16()
    qu = qkv[:, :d].contiguous()
    qv = (qu.float() + 0.1).bfloat16()
    Pn = 2 * T - 1
    pos = (torch.randn(Pn, d, device=dev, generator=g) * 0.5).bfloat16()
    lens = torch.randint(T * 3 // 5, T + 1, (B,), device=dev, dtype=torch.int32)
    k, v = qkv[:, d:2 * d], qkv[:, 2 * d:]
    drop = (a.drop, 1, 2) if a.drop > 0 else None
    rel = a.mode == "latest"

    def fused():
        return K.relattn_fwd(qu, qv if rel else None, k, v, pos if rel else None, lens, B, T, T, H, a.mode, False, 0.125, drop=drop,
                             want_pd=not a.no_pd)

    def unfused():
        ac = K.attn_f1(qu, k, B, T, T, H)
        bd = K.attn_f1(qv, pos, B, T, Pn, H, y_bcast=True) if rel else None
        pr, pd = K.softmax_fwd(ac, bd, lens, T, a.mode, False, 0.125, drop=drop)
        return pr, pd, K.attn_f2(pd, v, B, T, T, H)

    flops = B * H * (4.0 * T * T * 64 + (2.0 * T * Pn * 64 if rel else 0))
    tf, tu = timeit(fused), timeit(unfused)
    print(f"fwd  B={B} T={T} H={H} mode={a.mode} drop={a.drop}: fused {tf:.1f} us ({flops / tf * 1e-6:.0f} TFLOP/s)  unfused chain {tu:.1f} us")
    if a.bwd:
        pr, pd, c = K.relattn_fwd(qu, qv if rel else None, k, v, pos if rel else None, lens, B, T, T, H, a.mode, False, 0.125,
                                  drop=drop, want_pd=True)
        dc = torch.randn(B * T, d, device=dev, generator=g).bfloat16()
        dqkv = torch.empty_like(qkv)
        dqu, dqv = torch.empty(B * T, d, device=dev, dtype=torch.bfloat16), torch.empty(B * T, d, device=dev, dtype=torch.bfloat16)
        dpp = torch.zeros(Pn, d, device=dev)

        def fused_b():
            return K.relattn_bwd(dc, qu, k, v, pr, c, B, T, T, H, a.mode, 0.125, dqu, dqkv[:, d:2 * d], dqkv[:, 2 * d:], drop=drop,
                                 qv=qv if rel else None, pos=pos if rel else None, dqv=dqv if rel else None, dpos=dpp if rel else None)

        def unfused_b():
            dpd = K.attn_f1(dc, v, B, T, T, H)
            K.attn_f3(pd, dc, B, T, T, H, out=dqkv[:, 2 * d:])
            dac, dbd = K.softmax_bwd(dpd, pr, T, a.mode, 0.125, ld_bd=K.round8(Pn), drop=drop)
            K.attn_f2(dac, k, B, T, T, H, out=dqu)
            K.attn_f3(dac, qu, B, T, T, H, out=dqkv[:, d:2 * d])
            if rel:
                K.attn_f2(dbd, pos, B, T, Pn, H, y_bcast=True, out=dqv)
                K.attn_f4(dbd, qv, B, T, Pn, H, dpp)

        bflops = B * H * (8.0 * T * T * 64 + (4.0 * T * Pn * 64 if rel else 0))
        tb, tub = timeit(fused_b), timeit(unfused_b)
        print(f"bwd  B={B} T={T} H={H} mode={a.mode} drop={a.drop}: fused {tb:.1f} us ({bflops / tb * 1e-6:.0f} TFLOP/s)  unfused chain {tub:.1f} us")


if __name__ == "__main__":
    main()
