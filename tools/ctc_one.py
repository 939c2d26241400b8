"""One shape of the CTC sweep (for ncu): python tools/ctc_one.py B T V L [reps]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ctc_sweep import make
from espnet_slurp_b200 import ops

B, T, V, L = (int(x) for x in sys.argv[1:5])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
logits, hl, lab, ll = make(B, T, V, L)
dl, dh, dlab, dll = logits.cuda(), hl.int().cuda(), lab.int().cuda(), ll.int().cuda()
for _ in range(reps):
    _, loss_sum, ws = ops.ctc_fwd(dl, V, dh, dlab, dll)
    g32, _ = ops.ctc_bwd(dl, V, dh, dlab, dll, ws, 1.0, want_f32=True, want_bf16=False)
torch.cuda.synchronize()
print(float(loss_sum))
