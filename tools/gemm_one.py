"""Five representative GEMM launches of the train step, three rounds (ncu: -k regex:gemm_bf16 -s 10 -c 5 = third round)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from espnet_slurp_b200 import ops

torch.manual_seed(0)
M, d, F = int(os.environ.get("GEMM_M", 19840)), 512, 2048
x = torch.randn(M, d, device="cuda").bfloat16()
h = torch.randn(M, F, device="cuda").bfloat16()
w1 = (torch.randn(F, d, device="cuda") * 0.05).bfloat16()
w2 = (torch.randn(d, F, device="cuda") * 0.05).bfloat16()
b1 = torch.randn(F, device="cuda"); b2 = torch.randn(d, device="cuda")
res = torch.randn(M, d, device="cuda")
dact = torch.rand(M, F, device="cuda").bfloat16()
acc = torch.zeros(F, d, device="cuda")
for it in range(3):
    ops.gemm(x, w1, bias=b1)                                                             # 1: bias only, N=2048 K=512
    ops.gemm(x, w1, bias=b1, act="swish", act_grad_out=True, drop=(0.1, 1, 2))           # 2: FFN w_1 forward
    ops.gemm(x, w2, trans_b=True, act_grad_input=dact, act_grad="mul")                   # 3: FFN w_1 dgrad
    ops.gemm(h, w2, bias=b2, drop=(0.1, 1, 3), residual=res, alpha=0.5, out_dtype=torch.float32)  # 4: FFN w_2 forward
    ops.gemm(h, x, trans=True, out=acc, accumulate=True)                                  # 5: FFN w_1 wgrad (split-K, fp32 reduce-add)
torch.cuda.synchronize()
print("ok")
