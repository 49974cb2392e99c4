"""Synthetic equaliser outputs for the LLR producers (soft demodulator + descrambler): constellation points of TS 38.211 5.1
plus noise, the special values the reference's own unit test cares about (demodulation_mapper_test.cpp: zero / infinite / NaN
noise variances and symbols), values that sit on interval boundaries, and values around the near-zero threshold."""
import numpy as np

MODS = (0, 1, 2, 4, 6, 8)  # 0 = pi/2-BPSK


def bits_per_symbol(mod):
    return 1 if mod == 0 else mod


def modulate(bits, mod):
    """TS 38.211 5.1.1-5.1.6; bits: uint8 array of length n * Qm. Returns complex64 (n,)."""
    b = 1.0 - 2.0 * bits.astype(np.float64)
    qm = bits_per_symbol(mod)
    b = b.reshape(-1, qm)
    if mod in (0, 1):
        z = (b[:, 0] + 1j * b[:, 0]) / np.sqrt(2)
        if mod == 0:
            z = z * np.exp(1j * (np.arange(z.size) % 2) * np.pi / 2)
        return z.astype(np.complex64)
    if mod == 2:
        return ((b[:, 0] + 1j * b[:, 1]) / np.sqrt(2)).astype(np.complex64)
    if mod == 4:
        return ((b[:, 0] * (2 - b[:, 2]) + 1j * b[:, 1] * (2 - b[:, 3])) / np.sqrt(10)).astype(np.complex64)
    if mod == 6:
        re = b[:, 0] * (4 - b[:, 2] * (2 - b[:, 4]))
        im = b[:, 1] * (4 - b[:, 3] * (2 - b[:, 5]))
        return ((re + 1j * im) / np.sqrt(42)).astype(np.complex64)
    re = b[:, 0] * (8 - b[:, 2] * (4 - b[:, 4] * (2 - b[:, 6])))
    im = b[:, 1] * (8 - b[:, 3] * (4 - b[:, 5] * (2 - b[:, 7])))
    return ((re + 1j * im) / np.sqrt(170)).astype(np.complex64)


def equalised(n, mod, seed, noise_var=0.01, special=True):
    """(symbols float32 (n, 2), noise_vars float32 (n,), transmitted bits)."""
    r = np.random.default_rng(seed)
    qm = bits_per_symbol(mod)
    bits = r.integers(0, 2, n * qm, dtype=np.uint8)
    z = modulate(bits, mod)
    z = z + (r.normal(0, np.sqrt(noise_var / 2), n) + 1j * r.normal(0, np.sqrt(noise_var / 2), n)).astype(np.complex64)
    sym = np.stack([z.real, z.imag], axis=1).astype(np.float32)
    nv = (noise_var * r.uniform(0.5, 2.0, n)).astype(np.float32)
    if special and n > 8:
        k = r.integers(0, n, max(1, n // 10))
        vals = np.array([0, 1e-10, -1e-10, 1e-9, 3e-5, -3e-5, np.inf, -np.inf, np.nan, 1e30, -1e30, 3e9, -3e9, 1e-40,
                         2 / np.sqrt(10), 2 / np.sqrt(42), 4 / np.sqrt(170), -6 / np.sqrt(42), 14 / np.sqrt(170)], np.float32)
        sym[k, r.integers(0, 2, k.size)] = vals[r.integers(0, vals.size, k.size)]
        k = r.integers(0, n, max(1, n // 20))
        nvals = np.array([0, -1, np.inf, np.nan, 1e-30, 1e30, 1e-42], np.float32)
        nv[k] = nvals[r.integers(0, nvals.size, k.size)]
    return sym, nv, bits


def near_zero_boundary(n, seed):
    """Symbols whose squared magnitude sits within a few ulps of the 1e-9 near-zero threshold (math_utils.h:39,93-96)."""
    r = np.random.default_rng(seed)
    re = r.uniform(1e-6, 3.1e-5, n).astype(np.float32)
    im = np.sqrt(np.maximum(np.float64(1e-9) - re.astype(np.float64) ** 2, 0)).astype(np.float32)
    im = (im.view(np.int32) + r.integers(-3, 4, n).astype(np.int32)).view(np.float32)
    sym = np.stack([re * r.choice([-1, 1], n).astype(np.float32), im * r.choice([-1, 1], n).astype(np.float32)], axis=1)
    nv = r.uniform(1e-9, 1e-7, n).astype(np.float32)
    return sym.astype(np.float32), nv


def valid_llrs(n, seed):
    """int8 LLRs as the demodulator emits them: [-120, 120], with a few +-127 (infinity) mixed in (log_likelihood_ratio.h:65)."""
    r = np.random.default_rng(seed)
    x = r.integers(-120, 121, n).astype(np.int8)
    if n > 4:
        k = r.integers(0, n, max(1, n // 16))
        x[k] = r.choice(np.array([-127, 127], np.int8), k.size)
    return x
