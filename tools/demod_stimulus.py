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


def gold_sequence(n, c_init):
    """TS 38.211 5.2.1: c(i) = x1(i + 1600) ^ x2(i + 1600), i < n. The first 31 * 32 bits bit by bit, the rest in vector steps:
    over GF(2) the recurrences also hold at every power-of-two stride s (x(m) = x(m - 28 s) ^ x(m - 31 s) for x1)."""
    total = 1600 + n
    out = []
    for taps, init in (((0, 3), 1), ((0, 1, 2, 3), c_init & 0x7fffffff)):
        x = np.zeros(max(total, 1024) + 31, np.uint8)
        for i in range(31):
            x[i] = (init >> i) & 1
        L = 31
        while L < 1024:
            x[L] = np.bitwise_xor.reduce([x[L - 31 + t] for t in taps])
            L += 1
        while L < total:
            s = 1
            while 31 * s * 2 <= L:
                s *= 2
            m = min(28 * s, total - L)
            acc = np.zeros(m, np.uint8)
            for t in taps:
                acc ^= x[L - 31 * s + t * s:L - 31 * s + t * s + m]
            x[L:L + m] = acc
            L += m
        out.append(x[1600:total])
    return out[0] ^ out[1]


def valid_llrs(n, seed):
    """int8 LLRs as the demodulator emits them: [-120, 120], with a few +-127 (infinity) mixed in (log_likelihood_ratio.h:65)."""
    r = np.random.default_rng(seed)
    x = r.integers(-120, 121, n).astype(np.int8)
    if n > 4:
        k = r.integers(0, n, max(1, n // 16))
        x[k] = r.choice(np.array([-127, 127], np.int8), k.size)
    return x


def config2_transmission(n_tbs, sigma2, seed):
    """n_tbs config-2 transport blocks (273 PRB, 4 layers, 256QAM): scrambled, modulated, noisy equaliser outputs in the
    order the demodulator sees them (13 data OFDM symbols of 13104 symbols each), plus the descriptors of every stage."""
    from srsran_project_b200 import binding as B
    from tools import stimulus as T

    c = T.CONFIG2
    seg = T.segmentation(c["bg"], c["tbs"])
    Es = T.rm_lengths(seg["C"], c["G"], c["nof_layers"], c["Qm"])
    nsym_tb, per_call = c["G"] // 8, 273 * 12 * 4
    assert nsym_tb % per_call == 0
    symbols, noise = np.zeros((n_tbs * nsym_tb, 2), np.float32), np.zeros(n_tbs * nsym_tb, np.float32)
    segs = np.zeros(n_tbs * (nsym_tb // per_call), B.DEMOD_SEG_DTYPE)
    tbs_packed = []
    r = np.random.default_rng(seed)
    for t in range(n_tbs):
        tb = r.integers(0, 2, c["tbs"], dtype=np.uint8)
        tbs_packed.append(np.packbits(tb))
        _, bits, _ = T.tb_encode(c["bg"], c["tbs"], tb, c["rv"], c["Qm"], 0, c["nof_layers"], c["G"])
        c_init = ((0x4601 + t) << 15) + 41 + t
        tx = bits ^ gold_sequence(c["G"], c_init)
        z = modulate(tx, 8)
        z = z + (r.normal(0, np.sqrt(sigma2 / 2), z.size) + 1j * r.normal(0, np.sqrt(sigma2 / 2), z.size)).astype(np.complex64)
        symbols[t * nsym_tb:(t + 1) * nsym_tb, 0], symbols[t * nsym_tb:(t + 1) * nsym_tb, 1] = z.real, z.imag
        noise[t * nsym_tb:(t + 1) * nsym_tb] = np.float32(sigma2)
        for k in range(nsym_tb // per_call):
            s = segs[t * (nsym_tb // per_call) + k]
            s["sym_offset"], s["llr_offset"] = t * nsym_tb + k * per_call, (t * nsym_tb + k * per_call) * 8
            s["nof_symbols"], s["mod"], s["flags"] = per_call, 8, B.PDC_DEMOD_F_DESCRAMBLE
            s["c_init"], s["scramble_offset"] = c_init, k * per_call * 8
    descs = B.make_descs(n_tbs * seg["C"])
    tbd = np.zeros(n_tbs, B.TB_DESC_DTYPE)
    stride = (seg["K"] + 31) // 32 * 4
    tb_stride = (c["tbs"] // 8 + 3) // 4 * 4  # pdc_tb_desc.tb_offset must be a multiple of 4
    for t in range(n_tbs):
        o = t * c["G"]
        for q in range(seg["C"]):
            d = descs[t * seg["C"] + q]
            d["bg"], d["Z"], d["Qm"], d["rv"], d["crc_type"], d["max_iter"] = 1, seg["Z"], 8, 0, B.PDC_CRC24B, 6
            d["flags"], d["E"], d["nof_filler"] = B.PDC_F_NEW_DATA | B.PDC_F_EARLY_STOP, Es[q], seg["F"]
            d["llr_offset"], d["soft_offset"], d["out_offset"] = o, (t * seg["C"] + q) * 25344, (t * seg["C"] + q) * stride
            o += Es[q]
        tbd[t]["first_cb"], tbd[t]["nof_cbs"], tbd[t]["tbs_bits"], tbd[t]["tb_offset"] = t * seg["C"], seg["C"], c["tbs"], t * tb_stride
    return dict(symbols=symbols, noise=noise, segs=segs, descs=descs, tbd=tbd, tbs=tbs_packed, G=c["G"], C=seg["C"],
                stride=stride, tbs_bytes=c["tbs"] // 8, tb_stride=tb_stride)
