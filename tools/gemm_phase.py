"""Time a few representative GEMM shapes of the train step in isolation (CUDA events, warm L2).
Used with B200SLU_GEMM_DEBUG / B200SLU_GEMM_BN to attribute tile time to loads / MMA / epilogue / stores."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from espnet_slurp_b200 import ops

torch.manual_seed(0)
M, d, F = 9920, 512, 2048
x = torch.randn(M, d, device="cuda").bfloat16()
h = torch.randn(M, F, device="cuda").bfloat16()
w1 = (torch.randn(F, d, device="cuda") * 0.05).bfloat16()
w2 = (torch.randn(d, F, device="cuda") * 0.05).bfloat16()
wq = (torch.randn(1536, d, device="cuda") * 0.05).bfloat16()
b1 = torch.randn(F, device="cuda"); b2 = torch.randn(d, device="cuda"); bq = torch.randn(1536, device="cuda")
res = torch.randn(M, d, device="cuda")
dact = torch.rand(M, F, device="cuda").bfloat16()
acc = torch.zeros(F, d, device="cuda")
B, T, H = 64, 155, 8
qkv = torch.randn(B * T, 3 * d, device="cuda").bfloat16()

cases = {
    "ffn1 fwd bias+swish+dact+drop": lambda: ops.gemm(x, w1, bias=b1, act="swish", act_grad_out=True, drop=(0.1, 1, 2)),
    "ffn1 fwd bias only           ": lambda: ops.gemm(x, w1, bias=b1),
    "qkv  fwd bias (N=1536)       ": lambda: ops.gemm(x, wq, bias=bq),
    "ffn2 fwd bias+drop+res f32   ": lambda: ops.gemm(h, w2, bias=b2, drop=(0.1, 1, 3), residual=res, out_dtype=torch.float32),
    "ffn1 dgrad mul (N=2048,K=512)": lambda: ops.gemm(x, w2, trans_b=True, act_grad_input=dact, act_grad="mul"),
    "ffn2 dgrad plain (N=512)     ": lambda: ops.gemm(h, w1, trans_b=True),
    "proj dgrad plain (512,512)   ": lambda: ops.gemm(x, wq[:512], trans_b=True),
    "wgrad splitK 2048x512 K=9920 ": lambda: ops.gemm(h, x, trans=True, out=acc, accumulate=True),
    "one tile 128x256x64          ": lambda: ops.gemm(x[:128, :64], w1[:256, :64]),
    "148 tiles 128x256, K=512     ": lambda: ops.gemm(x[:128 * 37], w1[:1024]),
    "148 tiles 128x256, K=2048    ": lambda: ops.gemm(h[:128 * 37], w2[:256].repeat(4, 1)),
    "attn scores (155,155,64)x512 ": lambda: ops.attn_f1(qkv[:, :d], qkv[:, d:2 * d], B, T, T, H),
}
print("env BN=%s DEBUG=%s" % (os.environ.get("B200SLU_GEMM_BN"), os.environ.get("B200SLU_GEMM_DEBUG")))
for name, fn in cases.items():
    for _ in range(3): fn()
    torch.cuda.synchronize()
    n = 20
    g = torch.cuda.CUDAGraph()   # graph replay: no host launch cost between the kernels
    with torch.cuda.graph(g):
        for _ in range(n): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record(); torch.cuda.synchronize()
    print("%s %8.1f us" % (name, e0.elapsed_time(e1) * 1e3 / n))
