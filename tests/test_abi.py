"""The C-ABI library loads and exports every symbol include/b200slu.h declares, with the argument counts the ctypes
binding assumes (no compute calls: runs without a GPU)."""
import os
import re

from espnet_slurp_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "b200slu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(b200slu_\w+)\s*\(([^;{}]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    return out


def test_every_declared_symbol_is_exported_and_bound():
    decl = header_functions()
    assert len(decl) >= 35
    lib = _lib.lib()
    for name, nargs in decl.items():
        assert hasattr(lib, name), f"{name} declared in include/b200slu.h but not exported by libb200slu.so"
        assert name in _lib._SIGNATURES, f"{name} has no ctypes signature"
        assert len(_lib._SIGNATURES[name]) == nargs, (name, len(_lib._SIGNATURES[name]), nargs)
    assert set(_lib._SIGNATURES) == set(decl)
    assert lib.b200slu_version() == 100


def test_rel_shift_index_host_function_matches_oracle():
    import numpy as np
    from oracle import model as om
    lib = _lib.lib()
    for T in (1, 2, 5, 16, 33):
        col = om.rel_shift_index_latest(T)
        row_l, col_l = om.rel_shift_index_legacy(T)
        for i in range(T):
            for j in range(T):
                v = lib.b200slu_rel_shift_index(i, j, T, 1)
                assert (v >> 16, (v & 0xFFFF) - 1) == (i, col[i, j])
                w = lib.b200slu_rel_shift_index(i, j, T, 2)
                if col_l[i, j] < 0:
                    assert w == 0
                else:
                    assert (w >> 16, (w & 0xFFFF) - 1) == (row_l[i, j], col_l[i, j])


def test_no_product_import_of_the_oracle():
    """The product package must never route through oracle/ (bench.py and smoke() may, as checker / CPU baseline)."""
    pkg = os.path.join(ROOT, "espnet_slurp_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            assert "oracle" not in open(os.path.join(pkg, fn)).read(), fn


def test_product_mel_matrix_equals_reference_golden():
    import numpy as np
    from espnet_slurp_b200.melbank import mel_matrix
    z = np.load(os.path.join(ROOT, "tests", "golden", "frontend.npz"))
    m = mel_matrix()
    assert m.is_contiguous()
    np.testing.assert_allclose(m.numpy(), z["melmat"], rtol=0, atol=1e-9)
