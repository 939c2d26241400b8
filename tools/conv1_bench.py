import sys, torch
sys.path.insert(0, ".")
from espnet_slurp_b200 import ops as K
x = torch.randn(128, 626, 80, device="cuda"); w = torch.randn(512, 9, device="cuda") * 0.3; b = torch.randn(512, device="cuda") * 0.1
fl = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for wb in (False, True):
    ts = []
    for _ in range(7):
        fl.zero_(); a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); K.conv1_fwd(x, w, b, want_bits=wb); e.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(e) * 1e3)
    print("conv1_fwd bits=%s %.1f us" % (wb, sorted(ts)[3]))
