"""Memory-bound kernels of the train step in isolation, at the shapes of the 128-utterance step (M = 128 x 155 rows).

Each case is replayed from a CUDA graph over ROT rotating buffer sets (working set >> 126 MB L2) and reported as
algorithmic GB/s against the measured HBM copy peak (MEASURED_PEAKS.json).  python tools/membound.py [case-substring]
(MB_ONCE=1: two plain launches per case and no timing, for `ncu --set full`)
"""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from espnet_slurp_b200 import ops

torch.manual_seed(0)
dev = "cuda"
B, T, d, F, H, K = 128, int(os.environ.get("MB_T", 155)), 512, 2048, 8, 31
M = B * T
ROT = 6
peak = 6500.0
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def rnd(*s, dt=torch.float32):
    return torch.randn(*s, device=dev).to(dt)


cases = []


def case(name, nbytes, make):
    cases.append((name, nbytes, make))


def mk_ln_fwd():
    x = rnd(M, d); g = rnd(d); b = rnd(d)
    return lambda: ops.layernorm_fwd(x, g, b)
case("ln_fwd  f32 -> bf16", M * d * 6, mk_ln_fwd)

def mk_ln_bwd():
    x = rnd(M, d); g = rnd(d); b = rnd(d); dy = rnd(M, d, dt=torch.bfloat16); add = rnd(M, d)
    _, _, st = ops.layernorm_fwd(x, g, b)
    dg = torch.zeros(d, device=dev); db = torch.zeros(d, device=dev)
    return lambda: ops.layernorm_bwd(dy, x, g, st, dg, db, dx_add=add)
case("ln_bwd  dy bf16 + dx_add", M * d * (2 + 4 + 4 + 4), mk_ln_bwd)

def mk_ln_bwd_g():
    x = rnd(M, d); g = rnd(d); b = rnd(d); dy = rnd(M, d, dt=torch.bfloat16); add = rnd(M, d)
    _, _, st = ops.layernorm_fwd(x, g, b)
    dg = torch.zeros(d, device=dev); db = torch.zeros(d, device=dev)
    return lambda: ops.layernorm_bwd(dy, x, g, st, dg, db, dx_add=add, want_g=True, g_alpha=0.5, g_drop=(0.1, 1, 2))
case("ln_bwd  + bf16 g out (drop)", M * d * (2 + 4 + 4 + 4 + 2), mk_ln_bwd_g)

def mk_cast_colsum():
    x = rnd(M, d); cs = torch.zeros(d, device=dev)
    return lambda: ops.cast_bf16(x, 0.5, (0.1, 1, 2), colsum_out=cs)
case("cast_colsum f32 -> bf16 (drop)", M * d * 6, mk_cast_colsum)

def mk_colsum(n):
    def mk():
        x = rnd(M, n, dt=torch.bfloat16); cs = torch.zeros(n, device=dev)
        return lambda: ops.colsum(x, cs)
    return mk
case("colsum bf16 N=512", M * 512 * 2, mk_colsum(512))
case("colsum bf16 N=2048", M * 2048 * 2, mk_colsum(2048))

def mk_softmax_fwd():
    ld = ops.round8(T); ldb = ops.round8(2 * T - 1)
    ac = rnd(H, B, T, ld); bd = rnd(H, B, T, ldb)
    kl = torch.full((B,), T, device=dev, dtype=torch.int32)
    return lambda: ops.softmax_fwd(ac, bd, kl, T, "latest", False, 0.125, (0.1, 1, 2))
_ld, _ldb = ((T + 7) // 8) * 8, ((2 * T - 1 + 7) // 8) * 8
case("softmax_fwd rel-pos latest (drop)", H * B * T * (_ld * 4 + T * 4 + _ld * 2 * 2), mk_softmax_fwd)

def mk_softmax_bwd():
    ld = ops.round8(T); ldb = ops.round8(2 * T - 1)
    p = torch.rand(H, B, T, ld, device=dev).bfloat16(); dpd = rnd(H, B, T, ld)
    return lambda: ops.softmax_bwd(dpd, p, T, "latest", 0.125, ldb, (0.1, 1, 2))
case("softmax_bwd rel-pos latest (drop)", H * B * T * (_ld * 4 + _ld * 2 + _ld * 2 + _ldb * 2), mk_softmax_bwd)

def mk_glu_fwd():
    pw1 = rnd(M, 2 * d, dt=torch.bfloat16); w = rnd(d, K); b = rnd(d)
    return lambda: ops.glu_dwconv_fwd(pw1, w, b, B, T, K)
case("glu_dw_fwd", M * d * (4 + 2), mk_glu_fwd)

def mk_glu_bwd():
    pw1 = rnd(M, 2 * d, dt=torch.bfloat16); w = rnd(d, K); do = rnd(M, d, dt=torch.bfloat16)
    dw = torch.zeros(d, K, device=dev); dbs = torch.zeros(d, device=dev)
    return lambda: ops.glu_dwconv_bwd(do, pw1, w, B, T, K, dw, dbs)
case("glu_dw_bwd", M * d * (2 + 4 + 4), mk_glu_bwd)

def mk_bn_fwd():
    x = rnd(M, d, dt=torch.bfloat16); mr = torch.rand(2 * d, device=dev) + 0.5; g = rnd(d); b = rnd(d)
    return lambda: ops.bn_swish_fwd(x, mr, g, b)
case("bn_swish_fwd", M * d * 4, mk_bn_fwd)

def mk_bn_bwd():
    x = rnd(M, d, dt=torch.bfloat16); dy = rnd(M, d, dt=torch.bfloat16); mr = torch.rand(2 * d, device=dev) + 0.5
    g = rnd(d); b = rnd(d); dg = torch.zeros(d, device=dev); db = torch.zeros(d, device=dev)
    return lambda: ops.bn_swish_bwd(dy, x, mr, g, b, dg, db)
case("bn_swish_bwd (reduce + apply)", M * d * (4 + 4 + 2), mk_bn_bwd)

def mk_pos_bias_fwd():
    q = rnd(M, d, dt=torch.bfloat16); u = rnd(d); v = rnd(d)
    return lambda: ops.pos_bias_fwd(q, u, v)
case("pos_bias_fwd", M * d * 6, mk_pos_bias_fwd)

def mk_pos_bias_bwd():
    a = rnd(M, d, dt=torch.bfloat16); b = rnd(M, d, dt=torch.bfloat16); dq = torch.empty(M, 3 * d, device=dev, dtype=torch.bfloat16)
    du = torch.zeros(d, device=dev); dv = torch.zeros(d, device=dev)
    return lambda: ops.pos_bias_bwd(a, b, dq, du, dv)
case("pos_bias_bwd", M * d * 6, mk_pos_bias_bwd)

sel = sys.argv[1] if len(sys.argv) > 1 else ""
print("peak %.0f GB/s (MEASURED_PEAKS hbm burst)" % peak)
for name, nbytes, make in cases:
    if sel not in name:
        continue
    if os.environ.get("MB_ONCE"):  # one plain launch per case: the ncu --set full target
        fns = [make() for _ in range(2)]
        for f in fns:
            f()
        torch.cuda.synchronize()
        continue
    fns = [make() for _ in range(ROT)]
    for f in fns:
        f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(3):
            for f in fns:
                f()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (3 * ROT)
    gbs = nbytes / us * 1e-3
    print("%-36s %8.1f us  %7.1f MB  %7.0f GB/s  %.2f of peak" % (name, us, nbytes / 1e6, gbs, gbs / peak))
    del fns, g
    torch.cuda.empty_cache()
