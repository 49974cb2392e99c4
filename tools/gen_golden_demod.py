"""Generates tests/golden/demod_v1.npz from the COMPILED, UNMODIFIED reference (oracle/_ref, AVX512-host build of
demodulation_mapper_impl and pseudo_random_generator_impl): run it in the build container, commit the result.

    python tools/gen_golden_demod.py

Content: for every modulation a list of segments (= calls of demodulate_soft) of lengths that exercise the AVX512 / AVX2 /
one-by-one loops and their remainders, with special values mixed in; inputs on which a fused and an unfused multiply-add give
different LLRs (found with the restatement, answered by the reference); symbols on the near-zero threshold; descrambling
cases (c_init, offset, LLRs) and raw sequence bits far into the sequence."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from tools import demod_stimulus as S  # noqa: E402

LENGTHS = (1, 3, 4, 7, 9, 16, 17, 33, 67, 130, 1031)


def main():
    ref, orc = pyoracle.DemodReference(), pyoracle.DemodOracle()
    out = {}
    for mod in S.MODS:
        syms, nvs, llrs, lens = [], [], [], []
        for n in LENGTHS:
            sym, nv, _ = S.equalised(n, mod, 1000 * mod + n, noise_var=0.02 if mod >= 6 else 0.2)
            syms.append(sym), nvs.append(nv), lens.append(n)
            llrs.append(ref.demodulate_soft(sym, nv, mod))
        out[f"mod{mod}_sym"], out[f"mod{mod}_nv"] = np.concatenate(syms), np.concatenate(nvs)
        out[f"mod{mod}_llr"], out[f"mod{mod}_len"] = np.concatenate(llrs), np.array(lens, np.uint32)
    # Fused vs unfused multiply-add: segments of one vector batch (AVX2 loop of 256QAM / 64QAM) or shorter (one-by-one loop).
    r = np.random.default_rng(11)
    for mod, L, name in ((8, 4, "avx2"), (6, 16, "avx2"), (6, 3, "scalar"), (4, 3, "scalar"), (8, 3, "scalar")):
        keep_sym, keep_nv, keep_llr = [], [], []
        N = L * 150000
        sym = r.normal(0, 0.8, (N, 2)).astype(np.float32)
        nv = (np.abs(r.normal(0.05, 0.03, N)) + 1e-3).astype(np.float32)
        res = {}
        for m in (0x00, 0x1f):
            orc.set_fused_mask(m)
            res[m] = np.concatenate([orc.demodulate_soft(sym[i:i + L], nv[i:i + L], mod) for i in range(0, N, L)])
        orc.set_fused_mask(0x1f)
        qm = S.bits_per_symbol(mod)
        for s in np.unique(np.nonzero(res[0] != res[0x1f])[0] // (L * qm))[:12]:
            keep_sym.append(sym[s * L:(s + 1) * L]), keep_nv.append(nv[s * L:(s + 1) * L])
            keep_llr.append(ref.demodulate_soft(keep_sym[-1], keep_nv[-1], mod))
        if keep_sym:
            out[f"fused_{name}_mod{mod}_L{L}_sym"] = np.concatenate(keep_sym)
            out[f"fused_{name}_mod{mod}_L{L}_nv"] = np.concatenate(keep_nv)
            out[f"fused_{name}_mod{mod}_L{L}_llr"] = np.concatenate(keep_llr)
        print(mod, L, name, len(keep_sym), "discriminating segments")
    # Near-zero threshold, one-by-one loop (segments of 3 symbols) and vector loops (segments of 32).
    for mod in (4, 6, 8):
        sym, nv = S.near_zero_boundary(96, 70 + mod)
        out[f"nz_mod{mod}_sym"], out[f"nz_mod{mod}_nv"] = sym, nv
        out[f"nz_mod{mod}_llr3"] = np.concatenate([ref.demodulate_soft(sym[i:i + 3], nv[i:i + 3], mod) for i in range(0, 96, 3)])
        out[f"nz_mod{mod}_llr32"] = np.concatenate([ref.demodulate_soft(sym[i:i + 32], nv[i:i + 32], mod) for i in range(0, 96, 32)])
    # Descrambling.
    cases = [(0x12345, 0, 77), (0x7fffffff, 1, 1000), (1, 8191, 130), ((0x4601 << 15) + 935, 1362816 - 4000, 4000),
             (0, 31, 64), ((1 << 30) + 5, 100000, 257)]
    desc = []
    for k, (c_init, off, n) in enumerate(cases):
        x = S.valid_llrs(n, 90 + k)
        out[f"scr{k}_in"], out[f"scr{k}_out"] = x, ref.descramble(x, c_init, off)
        desc.append((c_init, off, n))
    out["scr_cases"] = np.array(desc, np.uint32)
    out["prg_far"] = ref.prg_bits(512, 0x2bcd1234, 3_000_000)
    path = os.path.join(ROOT, "tests", "golden", "demod_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
