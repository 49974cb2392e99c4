"""Shared helpers of the LLR-producer tests: the golden fixture as a list of (mod, symbols, noise_vars, expected LLRs) calls."""
import os

import numpy as np

from tools import demod_stimulus as S

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "demod_v1.npz")


def golden_calls():
    """Every call of demodulate_soft recorded in the fixture: (name, mod, symbols, noise_vars, reference LLRs)."""
    g = np.load(GOLDEN)
    calls = []
    for mod in S.MODS:
        qm, a = S.bits_per_symbol(mod), 0
        for n in g[f"mod{mod}_len"]:
            n = int(n)
            calls.append((f"mod{mod}_n{n}", mod, g[f"mod{mod}_sym"][a:a + n], g[f"mod{mod}_nv"][a:a + n],
                          g[f"mod{mod}_llr"][a * qm:(a + n) * qm]))
            a += n
    for key in g.files:
        if key.startswith("fused_") and key.endswith("_sym"):
            stem = key[:-4]
            mod, L = int(stem.split("_mod")[1].split("_")[0]), int(stem.split("_L")[1])
            qm = S.bits_per_symbol(mod)
            sym, nv, llr = g[stem + "_sym"], g[stem + "_nv"], g[stem + "_llr"]
            for i in range(0, nv.size, L):
                calls.append((f"{stem}_{i // L}", mod, sym[i:i + L], nv[i:i + L], llr[i * qm:(i + L) * qm]))
    for mod in (4, 6, 8):
        sym, nv = g[f"nz_mod{mod}_sym"], g[f"nz_mod{mod}_nv"]
        for L in (3, 32):
            llr = g[f"nz_mod{mod}_llr{L}"]
            for i in range(0, 96, L):
                calls.append((f"nz_mod{mod}_L{L}_{i // L}", mod, sym[i:i + L], nv[i:i + L], llr[i * mod:(i + L) * mod]))
    return calls


def golden_scrambling():
    g = np.load(GOLDEN)
    return [(int(c), int(o), g[f"scr{k}_in"], g[f"scr{k}_out"]) for k, (c, o, n) in enumerate(g["scr_cases"])], g["prg_far"]
