"""Quick tcgen05 GEMM throughput probe (CUDA events, L2-exceeding rotation of operands)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from espnet_slurp_b200 import ops

def probe(M, N, K, trans=False, iters=20):
    nbuf = max(2, int(300e6 // ((M * K + N * K) * 2)) + 1)
    nbuf = min(nbuf, 16)
    if not trans:
        As = [torch.randn(M, K, device="cuda").bfloat16() for _ in range(nbuf)]
        Bs = [torch.randn(N, K, device="cuda").bfloat16() for _ in range(nbuf)]
    else:
        As = [torch.randn(K, M, device="cuda").bfloat16() for _ in range(nbuf)]
        Bs = [torch.randn(K, N, device="cuda").bfloat16() for _ in range(nbuf)]
    out = torch.empty(M, N, device="cuda", dtype=torch.float32 if trans else torch.bfloat16)
    for i in range(3):
        ops.gemm(As[i % nbuf], Bs[i % nbuf], trans=trans, out=out, accumulate=trans)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        ops.gemm(As[i % nbuf], Bs[i % nbuf], trans=trans, out=out, accumulate=trans)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    a32 = As[0]; b32 = Bs[0]
    for i in range(3):
        (a32.t() @ b32) if trans else (a32 @ b32.t())
    t0.record()
    for i in range(iters):
        (As[i % nbuf].t() @ Bs[i % nbuf]) if trans else (As[i % nbuf] @ Bs[i % nbuf].t())
    t1.record(); torch.cuda.synchronize()
    ms_ref = t0.elapsed_time(t1) / iters
    tf = 2.0 * M * N * K / ms / 1e9
    print(json.dumps({"M": M, "N": N, "K": K, "trans": trans, "ms": round(ms, 4), "tflops": round(tf, 1),
                      "cublas_ms": round(ms_ref, 4), "cublas_tflops": round(2.0 * M * N * K / ms_ref / 1e9, 1)}), flush=True)

if __name__ == "__main__":
    for shp in [(8192, 8192, 8192), (7936, 2048, 512), (7936, 512, 2048), (7936, 1536, 512), (7936, 512, 512),
                (150784, 512, 4608), (7936, 512, 9728)]:
        probe(*shp)
    for shp in [(2048, 512, 7936), (512, 512, 7936), (512, 4608, 150784)]:
        probe(*shp, trans=True)
