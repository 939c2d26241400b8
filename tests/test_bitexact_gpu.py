"""The two bit-exact claims, checked ON THE DEVICE (integer / index work: no tolerance).

rel_shift (attention.py:240-263 latest, :145-165 legacy): the reference's pad/view trick is an index map, frozen from the
unmodified reference in tests/golden/rel_shift.npz for T <= 64 and restated by oracle/model.py for larger T.  Every
kernel that applies it is driven with ONE-HOT scores: exactly one element of the un-shifted matrix_bd per query row is
so large that the softmax is exactly one-hot (the other exponentials underflow to 0), so the position of the 1.0 in the
output IS the kernel's index map -- forward gather (softmax.cu warp-per-row and short-row kernels, attention.cu's
TMEM window / shift ring across query tiles) and backward scatter (softmax.cu).  All (i, j) pairs are covered by
running T batch entries, entry b testing j = (i + b) mod T for every row i.

CTC lattice (gtn_ctc.py:26-36): the blank-extended labels and skip-arc permissions as the device code computes them
(b200slu_ctc_lattice_states calls the same device functions as the lattice kernels) against the oracle's lattice.
"""
import os

import numpy as np
import pytest
import torch

from oracle import model as om

pytestmark = pytest.mark.gpu
BF = torch.bfloat16
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def K():
    from espnet_slurp_b200 import ops
    return ops


def round8(n):
    return (n + 7) // 8 * 8


def index_map(T, mode):
    """(row, col) int arrays [T, T] of the un-shifted source of shifted element (i, j); row = -1: the literal zero.
    T <= 64: straight from the reference's frozen outputs; beyond: the oracle's restatement (pinned to them for T <= 64)."""
    z = np.load(os.path.join(GOLDEN, "rel_shift.npz"))
    key = f"{mode}_{T}"
    if key in z.files:
        code = z[key].astype(np.int64)  # row * 10000 + col + 1, 0 = literal zero
        row, col = (code - 1) // 10000, (code - 1) % 10000
        row[code == 0], col[code == 0] = -1, 0
        return row, col
    if mode == "latest":
        col = om.rel_shift_index_latest(T).astype(np.int64)
        return np.broadcast_to(np.arange(T)[:, None], (T, T)).copy(), col
    row, col = om.rel_shift_index_legacy(T)
    row, col = row.astype(np.int64), col.astype(np.int64)
    row[col < 0] = -1
    col[col < 0] = 0
    return row, col


def test_golden_index_maps_available_for_both_modes():
    z = np.load(os.path.join(GOLDEN, "rel_shift.npz"))
    for T in (33, 64):
        for mode in ("latest", "legacy"):
            assert f"{mode}_{T}" in z.files
            row, col = index_map(T, mode)
            if mode == "latest":
                assert (row == np.arange(T)[:, None]).all()
                assert np.array_equal(col, om.rel_shift_index_latest(T))
            else:
                orow, ocol = om.rel_shift_index_legacy(T)
                assert np.array_equal(col[row >= 0], ocol[ocol >= 0]) and np.array_equal(row[row >= 0], orow[ocol >= 0])


def _offsets(T):
    """key offsets tested per batch entry: entry b puts the hot element of query row i at key (i + off_b) mod T; all offsets
    (= every (i, j) pair) for the exhaustive sizes, a spread of them beyond."""
    if T <= 64:
        return np.arange(T)
    return np.unique(np.concatenate([[0, 1, 2, 3, T // 2, T - 3, T - 2, T - 1], np.random.RandomState(T).randint(0, T, 8)]))


def check_forward_gather(ops, mode, T):
    row, col = index_map(T, mode)
    Pn = 2 * T - 1 if mode == "latest" else T
    ld, ld_bd = round8(T), round8(Pn)
    offs = _offsets(T)
    nb = len(offs)
    ii = np.arange(T)
    bd = np.zeros((nb, T, ld_bd), np.float32)
    for b, off in enumerate(offs):
        jj = (ii + off) % T
        r, c = row[ii, jj], col[ii, jj]
        bd[b, r[r >= 0], c[r >= 0]] = 4096.0
    dev = "cuda" if ops is not None and ops.__name__.endswith(".ops") else "cpu"
    lens = torch.full((nb,), T, dtype=torch.int32)
    p, _ = ops.softmax_fwd(torch.zeros(1, nb, T, ld).to(dev), torch.from_numpy(bd)[None].to(dev), lens.to(dev), T, mode, False, 0.125)
    p = p.float().cpu().numpy()[0][..., :T]
    for b in range(nb):
        scores = np.where(row >= 0, bd[b][np.maximum(row, 0), col], 0.0)  # the REFERENCE map applied on the host
        hot = scores == scores.max(1, keepdims=True)                      # rows without a hot source: all keys tie (uniform)
        expect = hot / hot.sum(1, keepdims=True)
        assert np.array_equal(p[b] > 0, hot), (mode, T, int(offs[b]))
        assert np.abs(p[b] - expect).max() <= 2.0 ** -9 * expect.max(), (mode, T, int(offs[b]))


def check_backward_scatter(ops, mode, T):
    """d(score)[i, j*] is the only positive entry of row i (uniform p, one-hot upstream gradient): its image in d_bd must
    sit at the reference's (row, col) and nowhere else."""
    row, col = index_map(T, mode)
    Pn = 2 * T - 1 if mode == "latest" else T
    ld, ld_bd = round8(T), round8(Pn)
    offs = _offsets(T)
    nb = len(offs)
    dev = "cuda" if ops.__name__.endswith(".ops") else "cpu"
    p = torch.zeros(1, nb, T, ld)
    p[..., :T] = 1.0 / 64  # exact in bf16; rows need not sum to 1 for the index test
    dpd = torch.zeros(1, nb, T, ld)
    ii = np.arange(T)
    for b, off in enumerate(offs):
        dpd[0, b, ii, (ii + off) % T] = 1.0
    _, dbd = ops.softmax_bwd(dpd.to(dev), p.to(BF).to(dev) if dev == "cuda" else p, T, mode, 0.125, ld_bd=ld_bd)
    dbd = dbd.float().cpu().numpy()[0]
    for b, off in enumerate(offs):
        expect = np.zeros((T, ld_bd), bool)
        jj = (ii + off) % T
        r, c = row[ii, jj], col[ii, jj]
        expect[r[r >= 0], c[r >= 0]] = True
        assert np.array_equal(dbd[b] > 0, expect), (mode, T, int(off))


FWD_CASES = [("latest", 33), ("latest", 64), ("legacy", 33), ("legacy", 64), ("latest", 155), ("latest", 230), ("latest", 600),
             ("legacy", 155), ("legacy", 300)]
BWD_CASES = [("latest", 33), ("latest", 64), ("legacy", 33), ("legacy", 64), ("latest", 155), ("latest", 600), ("legacy", 155)]


@pytest.mark.parametrize("mode,T", FWD_CASES)
def test_rel_shift_forward_gather_is_the_reference_index_map(mode, T):
    check_forward_gather(K(), mode, T)


@pytest.mark.parametrize("mode,T", BWD_CASES)
def test_rel_shift_backward_scatter_is_the_reference_index_map(mode, T):
    check_backward_scatter(K(), mode, T)


def check_fused_attention(ops, T):
    """attention.cu computes matrix_bd itself (tcgen05) and shifts it through the TMEM window + shared-memory ring, per
    128-query tile.  One-hot scores are built from binary codes: pos row n carries the code of n (two dims per bit), q_v of
    query (b, i) the code of its target source column, so (q_v . pos_n) is maximal for exactly one n, by a margin of
    1024 (x scale 0.125 x log2 e = 185 > 149: the other exponentials are exactly 0)."""
    row, col = index_map(T, "latest")
    Pn = 2 * T - 1
    nbits = int(np.ceil(np.log2(Pn)))
    assert 2 * nbits <= 64

    def code(n):
        n = np.asarray(n)
        out = np.zeros(n.shape + (64,), np.float32)
        for k in range(nbits):
            bit = (n >> k) & 1
            out[..., 2 * k] = bit
            out[..., 2 * k + 1] = 1 - bit
        return out * 32.0

    offs = np.arange(T) if T <= 64 else np.unique(np.array([0, 1, 2, 5, 63, 64, 65, T // 2, 127 % T, 128 % T, T - 2, T - 1]))
    nb = len(offs)
    ii = np.arange(T)
    qv = np.zeros((nb, T, 64), np.float32)
    want = np.zeros((nb, T), np.int64)
    for b, off in enumerate(offs):
        jj = (ii + off) % T
        qv[b] = code(col[ii, jj])
        want[b] = jj
    dev = "cuda" if ops.__name__.endswith(".ops") else "cpu"
    pos = torch.from_numpy(code(np.arange(Pn))).to(BF).to(dev)
    qv_t = torch.from_numpy(qv.reshape(nb * T, 64)).to(BF).to(dev)
    zeros = torch.zeros(nb * T, 64, dtype=BF, device=dev)
    v = torch.randn(nb * T, 64, generator=torch.Generator().manual_seed(0)).to(BF)
    lens = torch.full((nb,), T, dtype=torch.int32, device=dev)
    p, _, ctx = ops.relattn_fwd(zeros, qv_t, zeros, v.to(dev), pos, lens, nb, T, T, 1, "latest", False, 0.125)
    p = p.float().cpu().numpy()[0][..., :T]
    expect = np.zeros((nb, T, T), np.float32)
    expect[np.arange(nb)[:, None], ii[None, :], want] = 1.0
    assert np.array_equal(p, expect)
    # one-hot probabilities pick single value rows: the context is v[b, j*] exactly
    vv = v.float().numpy().reshape(nb, T, 64)
    assert np.array_equal(ctx.float().cpu().numpy().reshape(nb, T, 64), vv[np.arange(nb)[:, None], want])


@pytest.mark.parametrize("T", [33, 64, 150, 155, 192])
def test_fused_attention_rel_shift_is_the_reference_index_map(T):
    if not K().relattn_supported(T, T, 64, "latest"):
        pytest.skip("shape not on the fused kernel")
    check_fused_attention(K(), T)


def test_ctc_lattice_states_bit_exact_vs_gtn_lattice():
    rs = np.random.RandomState(0)
    B, Lmax, V = 64, 40, 12  # small vocabulary: many repeated labels (the no-skip rule)
    lens = rs.randint(0, Lmax + 1, B)
    lens[:4] = [0, 1, Lmax, 2]
    labels = rs.randint(1, V, (B, Lmax))
    labels[5, :6] = [3, 3, 5, 1, 1, 2]
    lens[5] = 6
    ext, skip = K().ctc_lattice_states(torch.from_numpy(labels).int().cuda(), torch.from_numpy(lens).int().cuda(), blank=0)
    ext, skip = ext.cpu().numpy(), skip.cpu().numpy()
    for b in range(B):
        we, ws = om.ctc_extended_labels(labels[b, : lens[b]].tolist(), 0)
        S = 2 * lens[b] + 1
        assert np.array_equal(ext[b, :S], we) and (ext[b, S:] == -1).all(), b
        assert np.array_equal(skip[b, :S], ws) and (skip[b, S:] == 0).all(), b
    assert ext[5, :13].tolist() == [0, 3, 0, 3, 0, 5, 0, 1, 0, 1, 0, 2, 0]
    assert skip[5, :13].tolist() == [0, 0, 0, 0, 0, 1, 0, 1, 0, 0, 0, 1, 0]
