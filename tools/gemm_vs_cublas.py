"""Every distinct non-batched GEMM call of one train step, timed in isolation twice: the fused tcgen05 kernel of this
library, and the library path eager PyTorch takes for the same arithmetic (cuBLAS GEMM via torch.addmm / matmul in bf16 +
the separate ATen epilogue passes: activation, dropout, scale, saved derivative, residual add, accumulate).

    python tools/gemm_vs_cublas.py [bench.py workload flags] > profiles/r2_gemm_vs_cublas.txt

Both sides: CUDA events around 20 back-to-back launches after 3 warm-ups, operands rotated over enough buffers to exceed
the 126 MB L2.  Sorted by time per step of our kernel.  "win" = cuBLAS path time / our time (> 1: ours is faster).
"""
import collections
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F

from espnet_slurp_b200 import _lib

calls = collections.OrderedDict()
lib = _lib.lib()
orig = lib.b200slu_gemm_bf16_ex


def key_of(g, ep):
    return (g.trans_a, g.trans_b, g.M, g.N, g.K, g.batch_inner, g.batch_outer, int(bool(ep.bias)), ep.act, int(bool(ep.act_grad_out)),
            int(ep.dropout_p > 0), int(bool(ep.act_grad_input)), ep.act_grad, int(bool(ep.residual)), ep.out_f32, ep.accumulate)


class Rec:
    def __call__(self, gp, epp, st):
        g = C.cast(gp, C.POINTER(_lib.GemmDesc)).contents
        ep = C.cast(epp, C.POINTER(_lib.Epilogue)).contents
        k = key_of(g, ep)
        if k not in calls:
            gc, ec = _lib.GemmDesc(), _lib.Epilogue()
            C.memmove(C.addressof(gc), gp, C.sizeof(gc))
            C.memmove(C.addressof(ec), epp, C.sizeof(ec))
            calls[k] = [0, gc, ec]
        calls[k][0] += 1
        return orig(gp, epp, st)


def timed(fn, reps=20):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


def main():
    import bench
    args = bench.parse()
    cfg = bench.build_cfg(args)
    torch.cuda.set_device(0)
    tr = bench.Trainer(args, cfg, 0, 1)
    tr.step(0)
    torch.cuda.synchronize()
    lib.b200slu_gemm_bf16_ex = Rec()
    tr.step(1)
    torch.cuda.synchronize()
    lib.b200slu_gemm_bf16_ex = orig
    del tr
    torch.cuda.empty_cache()
    st = torch.cuda.current_stream().cuda_stream
    rows = []
    for k, (n, g, ep) in calls.items():
        ta, tb, M, N, K, bi, bo = k[:7]
        if bi * bo != 1:
            continue  # the batched attention contractions are compared in tools/attn_bench.py
        bytes_ops = (M * K + N * K) * 2 + M * N * (4 if ep.out_f32 else 2)
        nbuf = max(2, min(12, int(300e6 // bytes_ops) + 1))
        As = [torch.randn((K, M) if ta else (M, K), device="cuda").bfloat16() for _ in range(nbuf)]
        Bs = [(torch.randn((K, N) if tb else (N, K), device="cuda") / K ** 0.5).bfloat16() for _ in range(nbuf)]
        out_dt = torch.float32 if ep.out_f32 else torch.bfloat16
        Cs = [torch.zeros(M, N, device="cuda", dtype=out_dt) for _ in range(nbuf)]
        bias = torch.randn(N, device="cuda") if ep.bias else None
        pre = [torch.zeros(M, N, device="cuda", dtype=torch.bfloat16) for _ in range(nbuf)] if ep.act_grad_out else None
        agi = [torch.randn(M, N, device="cuda").bfloat16() for _ in range(nbuf)] if ep.act_grad_input else None
        res = [torch.randn(M, N, device="cuda") for _ in range(nbuf)] if ep.residual else None
        g.lda, g.ldb, g.ldc = (M if ta else K), (N if tb else K), N

        def ours(i):
            j = i % nbuf
            g.A, g.B, g.C = As[j].data_ptr(), Bs[j].data_ptr(), Cs[j].data_ptr()
            if bias is not None:
                ep.bias = bias.data_ptr()
            if pre is not None:
                ep.act_grad_out = pre[j].data_ptr()
            if agi is not None:
                ep.act_grad_input, ep.ld_act_grad_input = agi[j].data_ptr(), N
            if res is not None:
                ep.residual, ep.ld_residual = res[j].data_ptr(), N
            _lib.check(orig(C.addressof(g), C.addressof(ep), st))

        bias16 = bias.bfloat16() if bias is not None else None

        def cublas(i):
            j = i % nbuf
            a = As[j].t() if ta else As[j]
            b = Bs[j] if tb else Bs[j].t()
            y = torch.addmm(bias16, a, b) if bias16 is not None else a @ b
            if ep.act == 1:
                if pre is not None:
                    s = torch.sigmoid(y)
                    pre[j].copy_(s * (1 + y * (1 - s)))
                y = F.silu(y)
            elif ep.act == 2:
                if pre is not None:
                    pre[j].copy_(y > 0)
                y = F.relu(y)
            if ep.dropout_p > 0:
                y = F.dropout(y, float(ep.dropout_p), True)
            if ep.alpha != 1.0:
                y = y * ep.alpha
            if agi is not None:
                if ep.act_grad == 3:
                    y = y * agi[j]
                elif ep.act_grad == 2:
                    y = y * (agi[j] > 0)
                else:
                    s = torch.sigmoid(agi[j])
                    y = y * (s * (1 + agi[j] * (1 - s)))
            if res is not None:
                Cs[j].copy_(res[j] + y) if not ep.accumulate else Cs[j].add_(res[j] + y)
            elif ep.accumulate:
                Cs[j].add_(y)
            else:
                Cs[j].copy_(y)

        us, us_ref = timed(ours), timed(cublas)
        fl = 2.0 * M * N * K
        rows.append((n * us, n, us, fl / us / 1e6, us_ref, fl / us_ref / 1e6, us_ref / us, k))
        del As, Bs, Cs, pre, agi, res
    rows.sort(reverse=True)
    tot, tot_ref = sum(r[0] for r in rows), sum(r[1] * r[4] for r in rows)
    print(f"# workload {args.workload}, {args.batch} utt/GPU; non-batched GEMM calls of one step, isolated, operands rotated beyond L2")
    print(f"# total per step: ours {tot / 1e3:.2f} ms, cuBLAS + separate ATen epilogue passes {tot_ref / 1e3:.2f} ms "
          f"over {sum(r[1] for r in rows)} launches")
    print("# per_step_us  n  ours_us  ours_TF  cublas_us  cublas_TF  win  (ta,tb,M,N,K,bi,bo,bias,act,saved_deriv,drop,actgrad_in,actgrad,res,f32,acc)")
    for r in rows:
        print(f"{r[0]:10.1f} {r[1]:4d} {r[2]:8.1f} {r[3]:8.1f} {r[4]:9.1f} {r[5]:8.1f} {r[6]:6.2f}  {r[7]}")


if __name__ == "__main__":
    main()
