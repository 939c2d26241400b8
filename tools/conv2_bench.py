"""conv.2 (Conv2dSubsampling's second convolution) at the bench shape: implicit GEMMs against the im2col + GEMM + col2im path.
Usage: python tools/conv2_bench.py [B T1 F1 C]"""
import sys

import torch

sys.path.insert(0, ".")
from espnet_slurp_b200 import ops as K  # noqa: E402

_a = [x for x in sys.argv[1:] if not x.startswith("--")]
B, T1, F1, C = (int(x) for x in _a[:4]) if len(_a) >= 4 else (128, 312, 39, 512)
T2, F2 = (T1 - 3) // 2 + 1, (F1 - 3) // 2 + 1
M = B * T2 * F2
g = torch.Generator(device="cuda").manual_seed(0)
h1 = torch.relu(torch.randn(B, T1, F1, C, device="cuda", generator=g)).bfloat16()
w = (torch.randn(C, 9 * C, device="cuda", generator=g) * 0.02).bfloat16()
bias = torch.randn(C, device="cuda", generator=g) * 0.1
dh2 = (torch.randn(M, C, device="cuda", generator=g) * 0.1).bfloat16()
dw = torch.zeros(C, 9 * C, device="cuda")
bits = K.relu_bits(h1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, n=5):
    fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return sorted(ts)[len(ts) // 2]


flops = 2.0 * M * C * 9 * C
if "--once" in sys.argv:  # one launch of each GEMM, for an ncu capture (-k regex:gemm_bf16)
    K.conv2_fwd(h1, w, bias)
    K.conv2_wgrad(dh2, h1, dw)
    K.conv2_dgrad_relu(dh2, w, h1, bits)
    col = K.im2col(h1)
    K.gemm(col, w, bias=bias, act="relu")
    K.gemm(dh2, col, trans=True, out=dw, accumulate=True)
    K.gemm(dh2, w, trans_b=True)
    torch.cuda.synchronize()
    sys.exit(0)
rows = []
rows.append(("fwd implicit", timed(lambda: K.conv2_fwd(h1, w, bias))))
rows.append(("wgrad implicit", timed(lambda: K.conv2_wgrad(dh2, h1, dw))))
rows.append(("dgrad implicit (4 class GEMMs)", timed(lambda: K.conv2_dgrad_relu(dh2, w, h1, bits))))
col = K.im2col(h1)
rows.append(("im2col", timed(lambda: K.im2col(h1))))
rows.append(("fwd GEMM on col", timed(lambda: K.gemm(col, w, bias=bias, act="relu"))))
rows.append(("wgrad GEMM on col", timed(lambda: K.gemm(dh2, col, trans=True, out=dw, accumulate=True))))
dcol = K.gemm(dh2, w, trans_b=True)
rows.append(("dgrad GEMM -> dcol", timed(lambda: K.gemm(dh2, w, trans_b=True))))
rows.append(("col2im + relu mask", timed(lambda: K.col2im_relu(dcol, h1))))
print(f"# conv.2 B={B} T1={T1} F1={F1} C={C}: M={M} pixels, {flops / 1e12:.3f} TFLOP per GEMM")
for name, us in rows:
    tf = flops / us / 1e6 if "GEMM" in name or "implicit" in name else float("nan")
    print(f"{name:34s} {us:9.1f} us   {tf:7.1f} TFLOP/s")
# agreement of the two paths on this data
a, b_ = K.conv2_fwd(h1, w, bias).float(), K.gemm(col, w, bias=bias, act="relu").float()
print("fwd max |diff|", float((a - b_).abs().max()), "of", float(b_.abs().max()))
a, b_ = K.conv2_dgrad_relu(dh2, w, h1, bits).float(), K.col2im_relu(dcol, h1).float()
print("dgrad max |diff|", float((a - b_).abs().max()), "of", float(b_.abs().max()))
