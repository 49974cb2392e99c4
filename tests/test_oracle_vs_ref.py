"""Pins the C restatement (oracle/) against the UNMODIFIED reference compiled from source (oracle/_ref).

Runs wherever oracle/_ref/libsrsran_ref.so exists (this container, and the GPU box since the .so travels).
"""
import numpy as np
import pytest

from oracle.pyoracle import (CRC16, CRC24A, CRC24B, CRC_NONE, FLAVOUR_AVX512, FLAVOUR_GENERIC, LIFTING_SIZES, kb_of,
                             nshort_of)


def awgn_llrs(rng, bits, sigma, amp=8.0):
    x = amp * (1.0 - 2.0 * bits.astype(np.float64)) + rng.normal(0.0, sigma, bits.size)
    return np.clip(np.round(x), -120, 120).astype(np.int8)


@pytest.mark.parametrize("crc_id", [CRC16, CRC24A, CRC24B])
def test_crc_matches_all_reference_impls(oracle, ref_avx512, crc_id):
    rng = np.random.default_rng(10 + crc_id)
    for n in (1, 5, 8, 16, 24, 25, 100, 128, 129, 1000, 3840, 8432, 8448):
        data = rng.integers(0, 256, (n + 7) // 8, dtype=np.uint8)
        want = oracle.crc(crc_id, data, n)
        for impl in (0, 1, 2):  # generic, LUT, PCLMUL
            assert ref_avx512.crc(impl, crc_id, data, n) == want, (n, impl)


def test_crc_known_answers(oracle):
    # crc(M || crc(M)) == 0 for every polynomial (the property the early stop relies on).
    rng = np.random.default_rng(3)
    for crc_id, order in ((CRC16, 16), (CRC24A, 24), (CRC24B, 24)):
        bits = rng.integers(0, 2, 1000, dtype=np.uint8)
        c = oracle.crc(crc_id, np.packbits(bits), bits.size)
        tail = np.array([(c >> (order - 1 - i)) & 1 for i in range(order)], np.uint8)
        full = np.concatenate([bits, tail])
        assert oracle.crc(crc_id, np.packbits(full), full.size) == 0


@pytest.mark.parametrize("bg", [1, 2])
def test_encoder_matches_reference_all_lifting_sizes(oracle, ref_avx512, ref_generic, bg):
    rng = np.random.default_rng(20 + bg)
    for Z in LIFTING_SIZES:
        msg = rng.integers(0, 2, kb_of(bg) * Z, dtype=np.uint8)
        cb = oracle.encode(bg, Z, msg)
        assert (cb == ref_avx512.encode(bg, Z, msg)).all(), Z
        assert (cb == ref_generic.encode(bg, Z, msg)).all(), Z


@pytest.mark.parametrize("bg", [1, 2])
def test_rate_matcher_matches_reference(oracle, ref_generic, bg):
    rng = np.random.default_rng(30 + bg)
    for Z in (2, 7, 24, 96, 208, 384):
        K, N = kb_of(bg) * Z, nshort_of(bg) * Z
        for trial in range(6):
            Qm = (1, 2, 4, 6, 8, 2)[trial]
            rv = trial % 4
            F = int(rng.integers(0, Z)) if trial % 2 else 0
            Nref = 0 if trial < 3 else int(rng.integers((kb_of(bg) + 6) * Z, N + 50))
            E = int(rng.integers(K // 2, 3 * N)) // Qm * Qm
            msg = rng.integers(0, 2, K, dtype=np.uint8)
            msg[K - F:] = 0
            cb = oracle.encode(bg, Z, msg)
            got = oracle.rate_match(bg, Z, cb, rv, Qm, Nref, F, E)
            want = ref_generic.encode_rm(bg, Z, msg, rv, Qm, Nref, F, E)
            assert (got == want).all(), (Z, trial)


@pytest.mark.parametrize("flavour", [FLAVOUR_AVX512, FLAVOUR_GENERIC])
@pytest.mark.parametrize("bg", [1, 2])
def test_dematcher_matches_reference(oracle, ref_avx512, ref_generic, flavour, bg):
    ref = ref_avx512 if flavour == FLAVOUR_AVX512 else ref_generic
    rng = np.random.default_rng(40 + bg + flavour)
    for Z in LIFTING_SIZES:
        K, N = kb_of(bg) * Z, nshort_of(bg) * Z
        for trial in range(8):
            Qm = (1, 2, 4, 6, 8, 8, 2, 4)[trial]
            F = int(rng.integers(0, 2 * Z)) if trial % 2 else 0
            Nref = 0 if trial < 4 else int(rng.integers((kb_of(bg) + 3) * Z, N + 50))
            hi = (6 * N) if trial == 7 else (N + N // 2)
            E = max(Qm, int(rng.integers(Z, hi)) // Qm * Qm)
            # Garbage (but in-range) soft buffer pins the write set of the copy mode.
            garbage = rng.integers(-120, 121, N).astype(np.int8)
            a, b = garbage.copy(), garbage.copy()
            for tx, rv in enumerate(rng.permutation(4)[:3]):
                llrs = rng.integers(-120, 121, E).astype(np.int8)
                new = tx == 0
                assert oracle.rate_dematch(flavour, a, llrs, new, int(rv), Qm, Nref, F) == 0
                ref.rate_dematch(b, llrs, new, bg, Z, int(rv), Qm, Nref, F)
                assert (a == b).all(), (Z, trial, tx)


@pytest.mark.parametrize("flavour", [FLAVOUR_AVX512, FLAVOUR_GENERIC])
@pytest.mark.parametrize("bg", [1, 2])
def test_decoder_matches_reference_all_lifting_sizes(oracle, ref_avx512, ref_generic, flavour, bg):
    ref = ref_avx512 if flavour == FLAVOUR_AVX512 else ref_generic
    rng = np.random.default_rng(50 + bg + flavour)
    kb, ns = kb_of(bg), nshort_of(bg)
    for Z in LIFTING_SIZES:
        K = kb * Z
        for trial in range(4):
            msg = rng.integers(0, 2, K, dtype=np.uint8)
            cb = oracle.encode(bg, Z, msg)
            n = ns * Z if trial == 0 else int(rng.integers(kb + 2, ns + 1)) * Z
            llrs = awgn_llrs(rng, cb[:n], (2.0, 4.0, 6.0, 9.0)[trial])
            if trial == 3:
                llrs[rng.integers(0, n, n // 3)] = 0
            if trial == 2:
                llrs[int(rng.integers(K, n)):] = 0  # trailing zeros => fewer layers
            for crc_id in (CRC_NONE, CRC16, CRC24B):
                for max_iter in (1, 6):
                    it_o, out_o = oracle.decode(flavour, bg, Z, llrs, 0, max_iter, crc_id)
                    it_r, out_r = ref.decode(bg, Z, llrs, 0, max_iter, crc_id)
                    assert it_o == it_r, (Z, trial, crc_id, max_iter)
                    assert (np.unpackbits(out_o)[:K] == np.unpackbits(out_r)[:K]).all(), (Z, trial, crc_id)


def test_decoder_early_stop_iterations_match(oracle, ref_avx512):
    # Valid codewords carrying a CRC: the iteration count of the early stop must agree.
    rng = np.random.default_rng(60)
    for bg, Z, crc_id, ncrc in ((1, 384, CRC24B, 24), (1, 96, CRC24A, 24), (2, 56, CRC16, 16), (2, 384, CRC24B, 24)):
        K = kb_of(bg) * Z
        F = 16
        seen = set()
        for sigma in (1.0, 4.0, 6.0, 7.0, 8.0, 10.0):
            payload = rng.integers(0, 2, K - F - ncrc, dtype=np.uint8)
            c = oracle.crc(crc_id, np.packbits(payload), payload.size)
            crc_bits = np.array([(c >> (ncrc - 1 - i)) & 1 for i in range(ncrc)], np.uint8)
            msg = np.concatenate([payload, crc_bits, np.zeros(F, np.uint8)])
            cb = oracle.encode(bg, Z, msg)
            llrs = awgn_llrs(rng, cb, sigma)
            llrs[K - 2 * Z - F:K - 2 * Z] = 127  # fillers as the dematcher leaves them
            it_o, out_o = oracle.decode(FLAVOUR_AVX512, bg, Z, llrs, F, 6, crc_id)
            it_r, out_r = ref_avx512.decode(bg, Z, llrs, F, 6, crc_id)
            assert it_o == it_r and (out_o == out_r).all(), (bg, Z, sigma)
            seen.add(it_o)
            if it_o > 0:
                assert (np.unpackbits(out_o)[:K - F] == msg[:K - F]).all()
        assert len(seen) >= 3, seen  # the sweep exercised several different iteration counts


def test_decoder_reference_known_answers(oracle, ref_avx512):
    # ldpc_enc_dec_test.cpp:340-364: all-zero LLRs => nullopt and an all-ones message (force_decoding flag set).
    for bg, Z in ((1, 8), (2, 384)):
        K, n = kb_of(bg) * Z, nshort_of(bg) * Z
        zeros = np.zeros(n, np.int8)
        for force in (False, True):
            it_o, out_o = oracle.decode(FLAVOUR_AVX512, bg, Z, zeros, 0, 2, CRC_NONE, force)
            it_r, out_r = ref_avx512.decode(bg, Z, zeros, 0, 2, CRC_NONE, force)
            assert it_o == it_r == -1
            assert (np.unpackbits(out_o)[:K] == np.unpackbits(out_r)[:K]).all()
            assert np.unpackbits(out_o)[:K].all()
