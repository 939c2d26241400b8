"""Run a few representative GEMM launches (for ncu --set full captures)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from espnet_slurp_b200 import ops
torch.manual_seed(0)
B, T, H, d = 64, 155, 8, 512
M = B * T
x = torch.randn(M, 3 * d, device="cuda").bfloat16()
w1 = torch.randn(2048, d, device="cuda").bfloat16()
ln = torch.randn(M, d, device="cuda").bfloat16()
bias = torch.randn(2048, device="cuda")
for it in range(3):
    # 1: attention scores, batched (155x155x64) x 512
    s = ops.attn_f1(x[:, :d], x[:, d:2 * d], B, T, T, H)
    # 2: FFN w_1 forward with bias + swish + pre-act + dropout
    h, pre = ops.gemm(ln, w1, bias=bias, act="swish", pre_act_out=True, drop=(0.1, 1, 2))
    # 3: plain dgrad 9920x512x2048
    g = ops.gemm(h, w1, trans_b=True)
    # 4: wgrad split-K 2048x512 K=9920
    acc = torch.zeros(2048, d, device="cuda")
    ops.gemm(h, ln, trans=True, out=acc, accumulate=True)
torch.cuda.synchronize()
print("ok")
