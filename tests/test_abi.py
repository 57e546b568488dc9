"""CPU-side checks of the product library: it loads, exports every symbol declared in
include/*.h, its host tables match the oracle, and it fails loudly without a GPU."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import harness as H

ROOT = H.ROOT


def _declared_symbols():
    """every function declared in include/*.h, macro families expanded by the C preprocessor"""
    syms = set()
    inc = os.path.join(ROOT, "include")
    for fn in sorted(os.listdir(inc)):
        if not fn.endswith(".h"):
            continue
        text = subprocess.run(["gcc", "-E", "-P", os.path.join(inc, fn)], check=True,
                              capture_output=True, text=True).stdout
        for m in re.finditer(r"\b((?:b200|av1|aom)_[A-Za-z0-9_]*)\s*\(", text):
            syms.add(m.group(1))
    return syms


def test_library_exports_declared_symbols():
    import b200av1
    L = b200av1.lib()
    missing = [s for s in sorted(_declared_symbols()) if not hasattr(L, s)]
    assert not missing, missing


def test_struct_sizes():
    import b200av1
    assert C.sizeof(b200av1.QRow) == 32
    assert b200av1.TXBLOCK_DTYPE.itemsize == 16


def test_host_tables_match_oracle():
    import b200av1
    L, O = b200av1.lib(), H.oracle()
    for ts in range(19):
        assert L.b200_tx_wide(ts) == O.av1o_tx_wide(ts)
        assert L.b200_tx_high(ts) == O.av1o_tx_high(ts)
        n = L.b200_tx_max_eob(ts)
        assert n == O.av1o_max_eob(ts)
        for tt in range(16):
            assert L.b200_tx_valid(ts, tt) == O.av1o_tx_valid(ts, tt)
            s, i = H.oracle_scan(ts, tt)
            assert np.array_equal(np.ctypeslib.as_array(L.b200_scan(ts, tt), shape=(n,)), s)
            assert np.array_equal(np.ctypeslib.as_array(L.b200_iscan(ts, tt), shape=(n,)), i)


def test_cospi_table_matches_oracle():
    """the kernels' cosine table is computed with the host libm; it must equal the reference's
    literal av1_cospi_arr_data (pinned in the oracle) -- the library also self-checks it"""
    import b200av1
    L, O = b200av1.lib(), H.oracle()
    L.b200_cospi.restype = C.POINTER(C.c_int32)
    for bit in range(10, 14):
        got = np.ctypeslib.as_array(L.b200_cospi(bit), shape=(64,))
        assert np.array_equal(got, np.ctypeslib.as_array(O.av1o_cospi(bit), shape=(64,)))


def test_no_gpu_fails_loudly():
    import b200av1
    if b200av1.lib().b200_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(b200av1.B200Error) as e:
        b200av1.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_build_quantizer_row_matches_oracle_and_reference():
    """b200_build_quantizer_row (host) against the oracle's row for every qindex / bit depth /
    sharpness, and -- where the reference build is present -- against av1_build_quantizer with
    non-zero plane deltas (u/v rows)."""
    import b200av1
    L, O = b200av1.lib(), H.oracle()
    row = b200av1.QRow()
    for bd in (8, 10, 12):
        for sharp in (0, 3, -2, 7, -7):
            for q in range(256):
                assert L.b200_build_quantizer_row(bd, q, 0, 0, sharp, C.byref(row)) == 0
                o = H.oracle_qrow(bd, q, 0, sharp)
                for name in ("zbin", "round", "quant", "quant_shift", "round_fp", "quant_fp", "dequant"):
                    assert list(getattr(row, name)) == list(getattr(o, name))[:2], (bd, sharp, q, name)
    assert L.b200_build_quantizer_row(9, 0, 0, 0, 0, C.byref(row)) != 0
    R = H.ref() if H.have_ref() else None
    if R is None:
        return

    class Quants(C.Structure):
        _fields_ = [(n, C.c_int16 * 8 * 256) for n in
                    ("y_quant", "y_quant_shift", "y_zbin", "y_round", "y_quant_fp",
                     "u_quant_fp", "v_quant_fp", "y_round_fp", "u_round_fp", "v_round_fp",
                     "u_quant", "v_quant", "u_quant_shift", "v_quant_shift", "u_zbin",
                     "v_zbin", "u_round", "v_round")]

    class Dequants(C.Structure):
        _fields_ = [(n, C.c_int16 * 8 * 256) for n in ("y", "u", "v")]

    for bd in (8, 10):
        for sharp in (0, 2, -3):
            qs, dq = Quants(), Dequants()
            ydc, udc, uac, vdc, vac = -7, 5, -12, 20, 63
            R.av1_build_quantizer(bd, ydc, udc, uac, vdc, vac, C.byref(qs), C.byref(dq), sharp)
            for q in range(0, 256, 3):
                for pl, (dcd, acd) in zip("yuv", ((ydc, 0), (udc, uac), (vdc, vac))):
                    assert L.b200_build_quantizer_row(bd, q, dcd, acd, sharp, C.byref(row)) == 0
                    for name in ("zbin", "round", "quant", "quant_shift", "round_fp", "quant_fp"):
                        theirs = getattr(qs, "%s_%s" % (pl, name))[q]
                        assert list(getattr(row, name)) == list(theirs)[:2], (bd, sharp, q, pl, name)
                    assert list(row.dequant) == list(getattr(dq, pl)[q])[:2], (bd, sharp, q, pl)
