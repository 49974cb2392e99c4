"""Generates tests/golden/pdsch_enc_v1.npz from the COMPILED, UNMODIFIED reference (oracle/_ref: pdsch_encoder_impl with
ldpc_encoder_avx2): SHA-256 digests of every case's codeword (packed) and, for the small cases, the codeword itself.

    python tools/gen_golden_enc.py
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from tools.enc_cases import CASES  # noqa: E402


def main():
    ref = pyoracle.PdschEncReference()
    out, digests = {}, []
    for k, (bg, tbs, Qm, L, nsym, rv, Nref) in enumerate(CASES):
        tb = np.random.default_rng(500 + k).integers(0, 256, tbs // 8, dtype=np.uint8)
        cw = ref.encode(tb, tbs, bg, rv, Qm, Nref, L, nsym)
        digests.append(np.frombuffer(hashlib.sha256(np.packbits(cw).tobytes()).digest(), np.uint8))
        if cw.size <= 40000:
            out[f"cw{k}"] = np.packbits(cw)
    out["digests"] = np.stack(digests)
    path = os.path.join(ROOT, "tests", "golden", "pdsch_enc_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
