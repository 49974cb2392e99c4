"""LLR producers, CPU side: the plain-C restatement (oracle/pdc_oracle_demod.c) against the golden fixture
recorded from the compiled reference (tests/golden/demod_v1.npz, tools/gen_golden_demod.py) and, where oracle/_ref is present,
against the compiled reference itself on fresh inputs. Bit-exact: the LLRs are int8."""
import numpy as np
import pytest

from oracle import pyoracle
from tests.demod_cases import golden_calls, golden_scrambling
from tools import demod_stimulus as S


@pytest.fixture(scope="module")
def orc():
    return pyoracle.DemodOracle()


@pytest.fixture(scope="module")
def ref():
    if not pyoracle.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return pyoracle.DemodReference()


def test_restatement_reproduces_every_golden_call(orc):
    calls = golden_calls()
    assert len(calls) > 150
    for name, mod, sym, nv, want in calls:
        got = orc.demodulate_soft(sym, nv, mod)
        assert np.array_equal(got, want), (name, np.nonzero(got != want)[0][:4])


def test_fused_multiply_add_placement_is_observable(orc):
    """The fixture holds inputs whose LLRs change when a multiply-add is not fused: an all-unfused restatement must fail on them
    (otherwise the fixture would not pin the placement)."""
    fused = [c for c in golden_calls() if c[0].startswith("fused_")]
    assert len(fused) >= 10
    orc.set_fused_mask(0)
    try:
        wrong = sum(not np.array_equal(orc.demodulate_soft(sym, nv, mod), want) for _, mod, sym, nv, want in fused)
    finally:
        orc.set_fused_mask(0x1f)
    assert wrong == len(fused)


def test_scrambling_sequence_and_descrambling_golden(orc):
    cases, prg_far = golden_scrambling()
    for c_init, off, x, want in cases:
        assert np.array_equal(orc.descramble(x, c_init, off), want), (c_init, off)
    assert np.array_equal(orc.prg_bits(512, 0x2bcd1234, 3_000_000), prg_far)
    # TS 38.211 5.2.1 by the definition, independent of both implementations.
    c_init, n = 0x5a5a5, 200
    x1, x2 = [0] * (1600 + n + 31), [0] * (1600 + n + 31)
    x1[0] = 1
    for i in range(31):
        x2[i] = (c_init >> i) & 1
    for i in range(1600 + n):
        x1[i + 31] = x1[i + 3] ^ x1[i]
        x2[i + 31] = x2[i + 3] ^ x2[i + 2] ^ x2[i + 1] ^ x2[i]
    want = np.array([x1[i + 1600] ^ x2[i + 1600] for i in range(n)], np.uint8)
    assert np.array_equal(orc.prg_bits(n, c_init, 0), want)


def test_descrambling_is_an_involution(orc):
    x = S.valid_llrs(5000, 3)
    y = orc.descramble(x, 0x1234567, 777)
    assert np.array_equal(np.abs(y.astype(np.int16)), np.abs(x.astype(np.int16)))
    assert np.array_equal(orc.descramble(y, 0x1234567, 777), x)


def test_noiseless_constellation_points_give_the_transmitted_bits(orc):
    for mod in S.MODS:
        sym, nv, bits = S.equalised(512, mod, 5, noise_var=1e-4, special=False)
        llr = orc.demodulate_soft(sym, nv, mod)
        assert np.array_equal((llr < 0).astype(np.uint8), bits), mod


@pytest.mark.parametrize("mod", S.MODS)
def test_restatement_matches_compiled_reference_on_fresh_inputs(orc, ref, mod):
    for n in list(range(0, 70)) + [127, 128, 1000, 4099, 13104]:
        for seed in range(2):
            sym, nv, _ = S.equalised(n, mod, 7000 + 13 * n + seed, noise_var=0.02 if mod >= 6 else 0.2)
            if n == 0:
                sym, nv = np.zeros((0, 2), np.float32), np.zeros(0, np.float32)
            a, b = ref.demodulate_soft(sym, nv, mod), orc.demodulate_soft(sym, nv, mod)
            assert np.array_equal(a, b), (mod, n, seed, np.nonzero(a != b)[0][:4])


def test_near_zero_threshold_matches_compiled_reference(orc, ref):
    for mod in (4, 6, 8):
        sym, nv = S.near_zero_boundary(3000, 17 + mod)
        for L in (3, 3000):
            for i in range(0, 3000, L):
                a = ref.demodulate_soft(sym[i:i + L], nv[i:i + L], mod)
                b = orc.demodulate_soft(sym[i:i + L], nv[i:i + L], mod)
                assert np.array_equal(a, b), (mod, L, i)


def test_descrambling_matches_compiled_reference(orc, ref):
    r = np.random.default_rng(8)
    for n in (0, 1, 15, 16, 17, 31, 32, 33, 63, 64, 65, 1000, 104832):
        c_init, off = int(r.integers(0, 1 << 31)), int(r.integers(0, 2_000_000))
        x = S.valid_llrs(n, 100 + n)
        assert np.array_equal(ref.descramble(x, c_init, off), orc.descramble(x, c_init, off)), (n, c_init, off)
