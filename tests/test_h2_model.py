"""The float16 arithmetic model of the packed CUDA kernels (tests/h2_model.py) against the oracle, on the CPU."""
import numpy as np
import pytest

from oracle.pyoracle import CRC16, CRC24B, CRC_NONE, FLAVOUR_AVX512, FLAVOUR_GENERIC, kb_of, nshort_of
from tests import h2_model
from tools import stimulus as S


def graph_rows(bg, Z):
    import ctypes

    from oracle import pyoracle as O

    if not O.ref_available():
        pytest.skip("oracle/_ref not built")
    lib = ctypes.CDLL(O.REF_SO)
    deg = np.zeros(46, np.uint16)
    cols = np.zeros((46, 20), np.uint16)
    sh = np.zeros((46, 20), np.uint16)
    vp = ctypes.c_void_p
    M = lib.ref_graph_dump(bg, Z, vp(deg.ctypes.data), vp(cols.ctypes.data), vp(sh.ctypes.data))
    return [[(int(cols[m, k]), int(sh[m, k])) for k in range(deg[m])] for m in range(M)]


def test_scale_tables():
    x = np.arange(0, 121).astype(np.float16)
    assert (h2_model.scale_simd(x).astype(int) == (np.arange(121) * 52428) >> 16).all()
    assert (h2_model.scale_generic(x).astype(int) == (8 * np.arange(121) + 5) // 10).all()


@pytest.mark.parametrize("flavour", [FLAVOUR_AVX512, FLAVOUR_GENERIC])
@pytest.mark.parametrize("bg,Z", [(1, 16), (2, 24), (1, 36)])
def test_model_matches_oracle(oracle, flavour, bg, Z):
    rows = graph_rows(bg, Z)
    kb, ns = kb_of(bg), nshort_of(bg)
    K = kb * Z
    rng = np.random.default_rng(5 + Z + flavour)
    for trial, (sigma, amp, ncols) in enumerate([(3.0, 8.0, ns), (9.0, 8.0, ns), (20.0, 30.0, ns), (5.0, 8.0, kb + 3),
                                                  (40.0, 60.0, kb + 8)]):
        msg = rng.integers(0, 2, K, dtype=np.uint8)
        cb = S.encode(bg, Z, msg)[:ncols * Z]
        llrs = np.clip(np.round(amp * (1 - 2.0 * cb) + rng.normal(0, sigma, cb.size)), -120, 120).astype(np.int8)
        if trial == 2:
            llrs[rng.integers(0, llrs.size, 40)] = 127  # "infinite" inputs (fillers look like this)
        nz = int(np.max(np.nonzero(llrs)[0])) + 1
        L = max((nz + 2 * Z + Z - 1) // Z, kb + 4) - kb
        for max_iter in (1, 3, 6):
            it, want = oracle.decode(flavour, bg, Z, llrs, 0, max_iter, CRC_NONE)
            _, soft = h2_model.decode(rows, Z, llrs, K, L, max_iter, generic=(flavour == FLAVOUR_GENERIC))
            got = (soft[:kb].reshape(-1) <= 0).astype(np.uint8)
            assert (got == np.unpackbits(want)[:K]).all(), (bg, Z, trial, max_iter, int((got != np.unpackbits(want)[:K]).sum()))
