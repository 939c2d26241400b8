"""CPU restatement (numpy, TEST CODE ONLY) of the library's counter-based dropout decisions
(espnet_slurp_b200/csrc/common.cuh: dropout_key / dropout_lanes8 / dropout_keep8) and the replay hook that feeds them to
the oracle's Dropout call sites.

The reference draws torch's Philox stream (torch.nn.Dropout at encoder_layer.py:97-149, attention.py:91,
positionwise_feed_forward.py:32, embedding.py:94,244, decoder_layer.py:88-131), which no other implementation can
reproduce; dropout parity is therefore checked with the keep decisions as an INPUT to both sides.  On the GPU the
decisions come from b200slu_dropout_mask (compared bit for bit with ``keep_mask`` below in tests/test_dropout_gpu.py).
"""
from __future__ import annotations

import numpy as np
import torch

_GOLD = np.uint64(0x9E3779B97F4A7C15)


def _splitmix64(z: np.ndarray) -> np.ndarray:
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def keep_mask(p: float, seed: int, site: int, n: int, dev_seed: int = 0) -> np.ndarray:
    """uint8 [n]: 1 where flat element e of dropout site (seed, site) is kept.  dev_seed: the registered device-resident
    per-step seed (b200slu_set_step_params), XORed in like the kernels do; 0 when none is registered."""
    with np.errstate(over="ignore"):
        sd = np.uint64((int(seed) ^ int(dev_seed)) & 0xFFFFFFFFFFFFFFFF)
        key = _splitmix64(np.asarray(sd + _GOLD * np.uint64(site + 1), dtype=np.uint64))
        ng = (n + 7) // 8
        g = np.arange(ng, dtype=np.uint64)
        z0 = _splitmix64(key + g * _GOLD)
        z1 = (z0 ^ (z0 >> np.uint64(29))) * np.uint64(0xD6E8FEB86659FD93)
        z1 = z1 ^ (z1 >> np.uint64(32))
    lanes = np.empty((ng, 8), dtype=np.uint32)
    for q, z in enumerate((z0, z0 >> np.uint64(32), z1, z1 >> np.uint64(32))):
        w = (z & np.uint64(0xFFFFFFFF)).astype(np.uint32)
        lanes[:, 2 * q] = w & np.uint32(0xFFFF)
        lanes[:, 2 * q + 1] = w >> np.uint32(16)
    thr = np.uint32(int(np.float32(p) * np.float32(65536.0)))
    return (lanes >= thr).astype(np.uint8).reshape(-1)[:n]


def round8(n: int) -> int:
    return (n + 7) // 8 * 8


def multiplier(drop, shape, mask_fn=keep_mask):
    """fp32 tensor of ``shape``: keep / (1 - p) of the site ``drop`` = (p, seed, site), flat row-major element index."""
    p, seed, site = drop
    n = int(np.prod(shape))
    keep = np.asarray(mask_fn(p, seed, site, n)).reshape(shape)
    inv = np.float32(1.0) / (np.float32(1.0) - np.float32(p))
    return torch.from_numpy(keep.astype(np.float32)) * float(inv)


class DropReplay:
    """The ``drop`` hook of oracle/model.py: replays the CUDA path's decisions.  sites = StepCtx.sites of the CUDA (or
    emulated) step: tag -> (p, site).  Layout per site = the tensor the reference's Dropout acts on, except attention
    probabilities, which the kernels index head-major with the key axis padded to a multiple of 8:
    ((h*B + b)*Tq + i) * round8(Tk) + j."""

    def __init__(self, sites: dict, seed: int, mask_fn=keep_mask):
        self.sites, self.seed, self.mask_fn = sites, seed, mask_fn
        self.used = set()

    def __call__(self, x: torch.Tensor, tag: str) -> torch.Tensor:
        p, site = self.sites[tag]
        self.used.add(tag)
        if p <= 0.0:
            return x
        if tag.endswith(".att"):
            B, H, Tq, Tk = x.shape
            m = multiplier((p, self.seed, site), (H, B, Tq, round8(Tk)), self.mask_fn)[..., :Tk].permute(1, 0, 2, 3)
        else:
            m = multiplier((p, self.seed, site), tuple(x.shape), self.mask_fn)
        return x * m.to(x.dtype)
