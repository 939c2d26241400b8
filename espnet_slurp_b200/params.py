"""Parameter storage for the B200 step: one flat fp32 master buffer, one flat bf16 copy, one flat fp32 gradient buffer.

The drop-in contract keeps every ``nn.Parameter`` under its reference state_dict name and shape;
``ParamPack`` re-points their ``.data`` (and ``.grad``) at views of the flat buffers so that

  * the whole bf16 weight refresh is ONE cast kernel per step,
  * q/k/v (and the decoder's source-attention k/v) projections sit back to back and are used as one fused
    [3d, d] / [2d, d] GEMM operand without any copy,
  * the data-parallel gradient all-reduce runs over large contiguous chunks of one buffer.

Only two weights need a kernel-side layout that differs from the reference's and get a separate bf16 copy
(+ fp32 gradient scratch that ``finalize_grads`` permutes back):
  encoder.embed.conv.2.weight (C, C, 3, 3)  ->  [C, (kh, kw, c)]      im2col GEMM operand
  encoder.embed.out.0.weight  (d, C*F2)     ->  [d, (f, c)]           consumes the channels-last conv output directly
"""
from __future__ import annotations

from typing import Dict, List

import torch

from . import step as _step


def _order(names: List[str]) -> List[str]:
    """Reference order, except q/k/v (self_attn) and k/v (src_attn) weights then biases are made adjacent."""
    out, done = [], set()
    have = set(names)
    for n in names:
        if n in done:
            continue
        group = [n]
        if n.endswith(".self_attn.linear_q.weight"):
            pre = n[: -len("linear_q.weight")]
            group = [pre + f"linear_{x}.{y}" for y in ("weight", "bias") for x in "qkv"]
        elif n.endswith(".src_attn.linear_k.weight"):
            pre = n[: -len("linear_k.weight")]
            group = [pre + f"linear_{x}.{y}" for y in ("weight", "bias") for x in "kv"]
        for g in group:
            if g in have and g not in done:
                out.append(g)
                done.add(g)
    return out


class ParamPack:
    def __init__(self, tensors: Dict[str, torch.Tensor], device=None, adopt: bool = False, lowp=torch.bfloat16,
                 grad_alloc=None):
        """tensors: name -> fp32 parameter (reference names / shapes).  adopt=True re-points each given tensor's
        ``.data`` at its flat view (use with nn.Parameters).  lowp: storage type of the GEMM-operand copy (bf16; fp32 only
        in the CPU tests of the orchestration).  grad_alloc(n): allocator of the flat fp32 gradient buffer (data-parallel
        runs place it in memory the other ranks can map: dist.peer_grad_alloc); zeroed here."""
        names = _order([k for k, v in tensors.items() if v.is_floating_point()])
        device = device if device is not None else next(iter(tensors.values())).device
        self.device = device
        offs, off = {}, 0
        for n in names:
            offs[n] = off
            off += (tensors[n].numel() + 7) // 8 * 8
        self.total = off
        self.flat32 = torch.zeros(off, device=device, dtype=torch.float32)
        self.flat16 = torch.zeros(off, device=device, dtype=lowp)
        self.gflat = torch.zeros(off, device=device, dtype=torch.float32) if grad_alloc is None else grad_alloc(off).zero_()
        self.names, self.offs = names, offs
        self.shapes = {n: tuple(tensors[n].shape) for n in names}
        self.P: Dict[str, torch.Tensor] = {}
        self.W: Dict[str, torch.Tensor] = {}
        self.G: Dict[str, torch.Tensor] = {}       # what the kernels accumulate into
        self.Gref: Dict[str, torch.Tensor] = {}    # reference-layout views of gflat (become .grad)
        with torch.no_grad():
            for n in names:
                t = tensors[n]
                v = self.flat32[offs[n]: offs[n] + t.numel()].view(t.shape)
                v.copy_(t.detach().to(device=device, dtype=torch.float32))
                self.P[n] = v
                self.Gref[n] = self.gflat[offs[n]: offs[n] + t.numel()].view(t.shape)
                self.G[n] = self.Gref[n]
                if adopt:
                    t.data = v
                if t.dim() >= 2:
                    rows = t.shape[0]
                    self.W[n] = self.flat16[offs[n]: offs[n] + t.numel()].view(rows, -1)
                    self.G[n] = self.Gref[n].view(rows, -1)
        # fused projections (views; adjacency guaranteed by _order and by sizes being multiples of 8)
        for n in names:
            for q, fused, k in ((".self_attn.linear_q.weight", "linear_qkv", 3), (".src_attn.linear_k.weight", "linear_kv", 2)):
                if n.endswith(q):
                    pre = n[: n.rindex("linear_")]
                    rows, cols = self.shapes[n]
                    assert (rows * cols) % 8 == 0 and rows % 8 == 0, "fused projections need sizes that are multiples of 8"
                    o = offs[n]
                    self.W[pre + fused + ".weight"] = self.flat16[o: o + k * rows * cols].view(k * rows, cols)
                    self.G[pre + fused + ".weight"] = self.gflat[o: o + k * rows * cols].view(k * rows, cols)
                    ob = offs[n[: -len("weight")] + "bias"]
                    assert ob == o + k * rows * cols
                    self.P[pre + fused + ".bias"] = self.flat32[ob: ob + k * rows]
                    self.G[pre + fused + ".bias"] = self.gflat[ob: ob + k * rows]
        # the two re-laid-out weights
        self.special = {}
        for n in names:
            if n.endswith(".embed.conv.2.weight"):
                C_ = self.shapes[n][0]
                self.special[n] = ("conv2", C_)
                self.W[n] = torch.zeros(C_, 9 * C_, device=device, dtype=lowp)
                self.G[n] = torch.zeros(C_, 9 * C_, device=device, dtype=torch.float32)
            elif n.endswith(".embed.out.0.weight"):
                d_, CF = self.shapes[n]
                C_ = self.shapes[n[: -len("out.0.weight")] + "conv.2.weight"][0]
                self.special[n] = ("out", d_, C_, CF // C_)
                self.W[n] = torch.zeros(d_, CF, device=device, dtype=lowp)
                self.G[n] = torch.zeros(d_, CF, device=device, dtype=torch.float32)

    # -- per step -----------------------------------------------------------------------------------------------
    def refresh(self, cast: bool = True):
        """bf16 copies of the current fp32 masters (one cast kernel + the two re-layouts).  cast=False: ``flat16`` is
        current already (the optimizer kernel wrote it with the update); only the re-layouts are redone."""
        if cast:
            _step.K.cast_bf16(self.flat32, out=self.flat16)
        for n, sp in self.special.items():
            o, num = self.offs[n], 1
            for s in self.shapes[n]:
                num *= s
            src = self.flat16[o: o + num]
            if sp[0] == "conv2":
                C_ = sp[1]
                self.W[n].view(C_, 3, 3, C_).copy_(src.view(C_, C_, 3, 3).permute(0, 2, 3, 1))
            else:
                _, d_, C_, F2 = sp
                self.W[n].view(d_, F2, C_).copy_(src.view(d_, C_, F2).permute(0, 2, 1))

    def zero_grad(self):
        self.gflat.zero_()
        for n in self.special:
            self.G[n].zero_()

    def finalize_grads(self):
        """Bring the re-laid-out gradients back to the reference layout (into gflat)."""
        for n, sp in self.special.items():
            if sp[0] == "conv2":
                C_ = sp[1]
                self.Gref[n].copy_(self.G[n].view(C_, 3, 3, C_).permute(0, 3, 1, 2))
            else:
                _, d_, C_, F2 = sp
                self.Gref[n].view(d_, C_, F2).copy_(self.G[n].view(d_, F2, C_).permute(0, 2, 1))

    def shapes_numel(self, n: str) -> int:
        num = 1
        for s in self.shapes[n]:
            num *= s
        return num

    def grads(self) -> Dict[str, torch.Tensor]:
        return self.Gref
