"""Generates tests/golden/ulsch_demux_v1.npz from the compiled, UNMODIFIED reference demultiplexer (oracle/_ref): for every case of
tests/demux_cases.py the SHA-256 of each of the four output streams, their lengths, and for the small cases the streams
themselves. Run in the build container (needs /root/reference compiled by oracle/build_ref.sh); the fixture travels."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import demux_oracle as D  # noqa: E402
from tests import demux_cases as C  # noqa: E402


def digest(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a, np.int8).tobytes()).digest(), np.uint8)


def main():
    ref = D.UlschDemuxReference()
    cs = C.cases()
    out = {"digests": np.zeros((len(cs), 4, 32), np.uint8), "lengths": np.zeros((len(cs), 4), np.uint32)}
    for k, cfg in enumerate(cs):
        cw = C.codeword(cfg, 100 + k)
        streams = ref.demultiplex(cfg, cw, block=0)
        again = ref.demultiplex(cfg, cw, block=997)  # block size of the on_new_block calls must not matter
        for q, (a, b) in enumerate(zip(streams, again)):
            assert np.array_equal(a, b), (k, q)
            out["digests"][k, q] = digest(a)
            out["lengths"][k, q] = a.size
        if cw.size <= 6000:
            for q, a in enumerate(streams):
                out[f"s{k}_{q}"] = a
    path = os.path.join(ROOT, "tests", "golden", "ulsch_demux_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(cs), "cases")


if __name__ == "__main__":
    main()
