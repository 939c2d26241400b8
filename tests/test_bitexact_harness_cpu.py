"""CPU self-check of the one-hot index-map harness of tests/test_bitexact_gpu.py: run against the plain-torch statement
of the kernels (tests/emul_ops.py, whose _shift is written independently with gather/where), it must pass -- and it must
FAIL when the statement's index map is perturbed.  (The GPU tests run the same checks against the CUDA kernels.)"""
import numpy as np
import pytest
import torch

import emul_ops as E
import test_bitexact_gpu as H


@pytest.fixture(autouse=True)
def _exact():
    old = E.exact
    E.exact = True
    yield
    E.exact = old


@pytest.mark.parametrize("mode,T", [("latest", 33), ("legacy", 33), ("latest", 64), ("legacy", 64), ("latest", 100), ("legacy", 100)])
def test_harness_accepts_the_reference_map(mode, T):
    H.check_forward_gather(E, mode, T)
    H.check_backward_scatter(E, mode, T)
    if mode == "latest":
        H.check_fused_attention(E, T)


@pytest.mark.parametrize("mode", ["latest", "legacy"])
def test_harness_rejects_an_off_by_one_map(mode, monkeypatch):
    good = E._shift

    def bad(bd, T, m):
        out = good(bd, T, m)
        out[:, :, 5, 7] = out[:, :, 5, 8]  # one wrong source element
        return out

    monkeypatch.setattr(E, "_shift", bad)
    with pytest.raises(AssertionError):
        H.check_forward_gather(E, mode, 33)
