"""Decode-loop timing of TransformerDecoder.forward_one_step: `steps` hypothesis-extension steps of a beam
over one utterance's encoder memory -- (a) with the per-layer cache (last position is the only query, memory keys / values
projected once), (b) the uncached prefix re-run, (c) the reference's own cached step in eager PyTorch on this GPU
(oracle.decoder_forward_one_step with LIBRARY_OPS: cuBLAS / ATen, fp32).

    python tools/decode_bench.py [--beam 10] [--T 155] [--steps 20] [--vocab 600]
"""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from espnet_slurp_b200 import modules as M  # noqa: E402
from espnet_slurp_b200.testing import make_sd  # noqa: E402
from oracle import model as om  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--beam", type=int, default=10)
    ap.add_argument("--T", type=int, default=155)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--vocab", type=int, default=600)
    a = ap.parse_args()
    V, d, H, F, NB = a.vocab, 512, 8, 2048, 6
    ocfg = om.DecCfg(V, d, H, F, NB)
    sd = make_sd(om.EncCfg(80, d, H, F, 0, 31, "latest"), ocfg, V, seed=0)
    dec = M.TransformerDecoder(V, d, H, F, NB).cuda().eval()
    dec.load_state_dict({k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")}, strict=False)
    g = torch.Generator().manual_seed(0)
    mem = torch.randn(1, a.T, d, generator=g).cuda().expand(a.beam, -1, -1)
    ys = torch.randint(1, V - 1, (a.beam, a.steps + 1), generator=g).cuda()

    def loop(fn):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        cache = None
        for l in range(1, a.steps + 1):
            _, cache = fn(ys[:, :l], cache)
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) * 1e3

    cached = lambda y, c: dec.forward_one_step(y, None, mem, cache=c)
    rerun = lambda y, c: dec.forward_one_step(y, None, mem, cache=None)
    om.LIBRARY_OPS = True
    osd = {k: v.cuda() for k, v in sd.items() if k.startswith("decoder.")}
    memc = mem.contiguous()
    with torch.no_grad():
        eager = lambda y, c: om.decoder_forward_one_step(y, memc, osd, ocfg, cache=c)
        res = {}
        for name, fn in (("cached", cached), ("prefix re-run", rerun), ("reference eager PyTorch (cached, fp32)", eager)):
            loop(fn)
            res[name] = min(loop(fn) for _ in range(3))
    print(f"decode loop: beam {a.beam}, memory T = {a.T}, {a.steps} steps, 6 decoder blocks, V = {V} (host wall-clock incl. launches)")
    for k, v in res.items():
        print(f"  {k:42s} {v:8.2f} ms  ({v / a.steps:.3f} ms / step)")


if __name__ == "__main__":
    main()
