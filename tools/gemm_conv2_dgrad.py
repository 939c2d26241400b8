"""conv2 dgrad shape of the 128-utterance step: dcol[376960, 4608] = dY2[376960, 512] . W2[512, 4608] (NT), isolated."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from espnet_slurp_b200 import ops
M, N, K = 376960 // int(os.environ.get("DIV", "1")), 4608, 512
dy = torch.randn(M, K, device="cuda").bfloat16()
w = (torch.randn(K, N, device="cuda") * 0.05).bfloat16()
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(2): ops.gemm(dy, w, trans_b=True, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): ops.gemm(dy, w, trans_b=True, out=out)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / 3
print("conv2 dgrad M=%d: %.1f us  %.0f TFLOP/s" % (M, us, 2.0 * M * N * K / us * 1e-6))
