"""Allocation-free label-smoothing loss entry point (b200slu_lsm_loss_fwd_bwd_static: scratch in a static device array,
no counterpart in the reference) against the existing entry point with a
caller-provided workspace, through the C ABI, on two calls in a row.  Collected last on purpose (pytest -x)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _call(static: bool, logits, targets, V, smoothing, scale, ws=None):
    from espnet_slurp_b200 import _lib
    from espnet_slurp_b200.ops import ptr, stream_ptr
    lib = _lib.lib()
    R, ldv = logits.shape
    ldg = (V + 7) // 8 * 8
    stats = torch.full((3,), -7.0, device="cuda")
    g32 = torch.full((R, ldg), 9.0, device="cuda")
    g16 = torch.full((R, ldg), 9.0, device="cuda", dtype=torch.bfloat16)
    if static:
        rc = lib.b200slu_lsm_loss_fwd_bwd_static(ptr(logits), ldv, R, V, ptr(targets), -1, smoothing, scale, ptr(stats), ptr(g32),
                                                 ptr(g16), ldg, stream_ptr())
    else:
        rc = lib.b200slu_lsm_loss_fwd_bwd(ptr(logits), ldv, R, V, ptr(targets), -1, smoothing, scale, ptr(ws), ptr(stats),
                                          ptr(g32), ptr(g16), ldg, stream_ptr())
    return rc, stats, g32, g16


def test_static_scratch_lsm_loss_matches_workspace_entry_point():
    from espnet_slurp_b200 import _lib
    lib = _lib.lib()
    max_rows = lib.b200slu_lsm_static_max_rows()
    assert max_rows == 32768
    g = torch.Generator().manual_seed(3)
    # two calls in a row with DIFFERENT sizes and contents (the larger first): the second must not see the first's scratch
    for R, V in ((2688, 600), (301, 50)):
        ldv = (V + 7) // 8 * 8
        logits = (torch.randn(R, ldv, generator=g) * 2).cuda()
        targets = torch.randint(0, V, (R,), generator=g).int()
        targets[::7] = -1  # ignored rows
        targets = targets.cuda()
        ws = torch.empty(lib.b200slu_lsm_workspace_size(R), device="cuda", dtype=torch.uint8)
        rc0, st0, a32, a16 = _call(False, logits, targets, V, 0.1, 0.7 / 128, ws)
        before = torch.cuda.memory_allocated()
        rc1, st1, b32, b16 = _call(True, logits, targets, V, 0.1, 0.7 / 128)
        torch.cuda.synchronize()
        assert rc0 == 0 and rc1 == 0, lib.b200slu_last_error()
        # same kernels, same row-wise arithmetic: identical results (tolerance of tests/test_losses_frontend_gpu.py: 2e-5)
        np.testing.assert_allclose(st1.cpu().numpy(), st0.cpu().numpy(), rtol=2e-5)
        assert torch.equal(b32, a32) and torch.equal(b16, a16)
        assert float(st1[2]) == float((targets >= 0).sum())
        # and against plain torch, as the existing test does
        lp = torch.log_softmax(logits[:, :V].double(), 1)
        true = torch.full((R, V), 0.1 / (V - 1), dtype=torch.float64, device="cuda")
        valid = targets >= 0
        true[valid, targets[valid].long()] = 0.9
        want = (true * (true.log() - lp)).sum(1)[valid].sum()
        np.testing.assert_allclose(float(st1[0]), float(want), rtol=2e-5)
        # the entry point itself allocates nothing: with the outputs pre-allocated, neither torch's allocator nor the
        # driver (cudaMemGetInfo) sees a single byte move across a call (the first call above already did the per-device
        # symbol lookup / lazy module load)
        stats = torch.empty(3, device="cuda")
        torch.cuda.synchronize()
        free0, alloc0, resv0 = torch.cuda.mem_get_info()[0], torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
        rc2 = lib.b200slu_lsm_loss_fwd_bwd_static(logits.data_ptr(), ldv, R, V, targets.data_ptr(), -1, 0.1, 0.7 / 128,
                                                  stats.data_ptr(), b32.data_ptr(), b16.data_ptr(), (V + 7) // 8 * 8,
                                                  torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert rc2 == 0
        assert torch.cuda.mem_get_info()[0] == free0 and torch.cuda.memory_allocated() == alloc0 and \
            torch.cuda.memory_reserved() == resv0
        assert before > 0
    # beyond the static maximum: refused with a clear message, nothing launched
    R = max_rows + 1
    logits = torch.zeros(R, 8, device="cuda")
    targets = torch.zeros(R, device="cuda", dtype=torch.int32)
    rc, st, _, _ = _call(True, logits, targets, 8, 0.1, 1.0)
    torch.cuda.synchronize()
    assert rc != 0 and b"exceed the static scratch" in lib.b200slu_last_error()
    assert float(st[0]) == -7.0
    # exactly the maximum works
    rc, st, _, _ = _call(True, logits[:max_rows], targets[:max_rows], 8, 0.1, 1.0)
    torch.cuda.synchronize()
    assert rc == 0 and float(st[2]) == max_rows
