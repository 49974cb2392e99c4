"""PDSCH encoder, CPU side: the TX-chain restatement (tools/stimulus/pdc_stimulus.c, exposed as
oracle.pyoracle.PdschEncOracle) against the compiled, unmodified pdsch_encoder_impl of oracle/_ref (both LDPC encoder flavours)
and against the golden codeword digests recorded from it (tests/golden/pdsch_enc_v1.npz, tools/gen_golden_enc.py)."""
import hashlib
import os

import numpy as np
import pytest

from oracle import pyoracle
from tools.enc_cases import CASES

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pdsch_enc_v1.npz")


def tb_bytes(case_index, tbs_bits):
    return np.random.default_rng(500 + case_index).integers(0, 256, tbs_bits // 8, dtype=np.uint8)


def digest(cw):
    return np.frombuffer(hashlib.sha256(np.packbits(cw).tobytes()).digest(), np.uint8)


def test_restatement_reproduces_the_golden_codewords():
    g = np.load(GOLDEN)
    orc = pyoracle.PdschEncOracle()
    assert len(g["digests"]) == len(CASES)
    for k, (bg, tbs, Qm, L, nsym, rv, Nref) in enumerate(CASES[:-1]):  # the 1.3 Mbit case runs in the GPU suite
        cw = orc.encode(tb_bytes(k, tbs), tbs, bg, rv, Qm, Nref, L, nsym)
        assert cw.size == nsym * Qm and cw.max() <= 1
        assert np.array_equal(digest(cw), g["digests"][k]), CASES[k]
        if f"cw{k}" in g.files:
            assert np.array_equal(cw, np.unpackbits(g[f"cw{k}"])[:cw.size]), CASES[k]


@pytest.mark.parametrize("generic", [False, True])
def test_restatement_matches_compiled_pdsch_encoder(generic):
    if not pyoracle.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    ref, orc = pyoracle.PdschEncReference(generic), pyoracle.PdschEncOracle()
    for k, (bg, tbs, Qm, L, nsym, rv, Nref) in enumerate(CASES[:-1]):
        for rv2 in {rv, (rv + 2) % 4}:
            tb = tb_bytes(k, tbs)
            a, b = ref.encode(tb, tbs, bg, rv2, Qm, Nref, L, nsym), orc.encode(tb, tbs, bg, rv2, Qm, Nref, L, nsym)
            assert np.array_equal(a, b), (CASES[k], rv2, np.nonzero(a != b)[0][:4])
