"""Print the in-kernel phase cycle counts (B200SLU_GEMM_DEBUG bit 16) for a few shapes."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from espnet_slurp_b200 import ops
torch.manual_seed(0)
M, d, F = 9920, 512, 2048
x = torch.randn(M, d, device="cuda").bfloat16()
h = torch.randn(M, F, device="cuda").bfloat16()
w1 = (torch.randn(F, d, device="cuda") * 0.05).bfloat16()
w2 = (torch.randn(d, F, device="cuda") * 0.05).bfloat16()
b1 = torch.randn(F, device="cuda")
for rep in range(2):
    ops.gemm(x[:128 * 37], w1[:1024]);  torch.cuda.synchronize()              # 148 tiles, K=512
    ops.gemm(x[:128 * 37], w1[:1024], bias=b1[:1024]);  torch.cuda.synchronize()   # same, specialised bias kernel
    ops.gemm(h[:128 * 37], w2[:256].repeat(4, 1)); torch.cuda.synchronize()   # 148 tiles, K=2048
    ops.gemm(x, w1, bias=b1); torch.cuda.synchronize()                         # ffn1 bias only
    ops.gemm(h, w1, trans_b=True); torch.cuda.synchronize()                    # ffn2 dgrad
