"""Synthetic model state for benchmarks, smoke runs and tests: random-init parameters under the reference's
state_dict names and shapes (there is no network for checkpoints).  No arithmetic of the path lives here."""
from __future__ import annotations

import torch


def make_sd(enc, dec, V, seed=0):
    """Random-init parameters under the reference's state_dict names/shapes."""
    g = torch.Generator().manual_seed(seed)
    d, C, F2 = enc.d, enc.d, ((enc.idim - 1) // 2 - 1) // 2
    sd = {}

    def lin(name, o, i, bias=True):
        sd[name + ".weight"] = torch.randn(o, i, generator=g) / i ** 0.5
        if bias:
            sd[name + ".bias"] = torch.randn(o, generator=g) * 0.1

    def ln(name, n):
        sd[name + ".weight"] = 1 + 0.1 * torch.randn(n, generator=g)
        sd[name + ".bias"] = 0.1 * torch.randn(n, generator=g)

    sd["encoder.embed.conv.0.weight"] = torch.randn(C, 1, 3, 3, generator=g) / 3
    sd["encoder.embed.conv.0.bias"] = torch.randn(C, generator=g) * 0.1
    sd["encoder.embed.conv.2.weight"] = torch.randn(C, C, 3, 3, generator=g) / (9 * C) ** 0.5
    sd["encoder.embed.conv.2.bias"] = torch.randn(C, generator=g) * 0.1
    lin("encoder.embed.out.0", d, C * F2)
    for n in range(enc.blocks):
        p = f"encoder.encoders.{n}"
        for a in "qkv":
            lin(f"{p}.self_attn.linear_{a}", d, d)
        lin(f"{p}.self_attn.linear_out", d, d)
        lin(f"{p}.self_attn.linear_pos", d, d, bias=False)
        sd[f"{p}.self_attn.pos_bias_u"] = torch.randn(enc.heads, d // enc.heads, generator=g) * 0.1
        sd[f"{p}.self_attn.pos_bias_v"] = torch.randn(enc.heads, d // enc.heads, generator=g) * 0.1
        for f in ("feed_forward", "feed_forward_macaron"):
            lin(f"{p}.{f}.w_1", enc.ffn, d)
            lin(f"{p}.{f}.w_2", d, enc.ffn)
        sd[f"{p}.conv_module.pointwise_conv1.weight"] = torch.randn(2 * d, d, 1, generator=g) / d ** 0.5
        sd[f"{p}.conv_module.pointwise_conv1.bias"] = torch.randn(2 * d, generator=g) * 0.1
        sd[f"{p}.conv_module.depthwise_conv.weight"] = torch.randn(d, 1, enc.kernel, generator=g) / enc.kernel ** 0.5
        sd[f"{p}.conv_module.depthwise_conv.bias"] = torch.randn(d, generator=g) * 0.1
        ln(f"{p}.conv_module.norm", d)
        sd[f"{p}.conv_module.pointwise_conv2.weight"] = torch.randn(d, d, 1, generator=g) / d ** 0.5
        sd[f"{p}.conv_module.pointwise_conv2.bias"] = torch.randn(d, generator=g) * 0.1
        for nm in ("norm_ff", "norm_mha", "norm_ff_macaron", "norm_conv", "norm_final"):
            ln(f"{p}.{nm}", d)
    ln("encoder.after_norm", d)
    sd["decoder.embed.0.weight"] = torch.randn(V, d, generator=g)
    for n in range(dec.blocks):
        p = f"decoder.decoders.{n}"
        for att in ("self_attn", "src_attn"):
            for a in ("q", "k", "v", "out"):
                lin(f"{p}.{att}.linear_{a}", d, d)
        lin(f"{p}.feed_forward.w_1", dec.ffn, d)
        lin(f"{p}.feed_forward.w_2", d, dec.ffn)
        for nm in ("norm1", "norm2", "norm3"):
            ln(f"{p}.{nm}", d)
    ln("decoder.after_norm", d)
    lin("decoder.output_layer", V, d)
    lin("ctc.ctc_lo", V, d)
    return sd
