"""numpy float16 model of the arithmetic the packed fp16x2 CUDA kernels use (srsran_project_b200/csrc/pdc_kernel_h2*.cuh).

The kernels keep soft bits and messages as IEEE half floats: integers of magnitude <= 2048 are exact, the reference's
"+-127 = infinity" becomes a real +-inf, and every operation below is one half2 instruction on the GPU. This model
exists so that the formulation (equality trick instead of argmin, promotion by overflow, sticky infinities, the
fused-multiply-add scalings) can be checked against the oracle on the CPU, without a GPU.
"""
import numpy as np

F16 = np.float16


def fma16(a, b, c):
    """Fused multiply-add with ONE rounding to half (products/sums of our operands are exact in float64)."""
    with np.errstate(over="ignore", invalid="ignore"):
        return (a.astype(np.float64) * np.float64(b) + np.float64(c)).astype(F16)


def scale_simd(m):
    """(x * 52428) >> 16 for integers 0..120: HFMA2(x, 0.8h, 1023.5) -> max(.,1024) -> -1024."""
    r = fma16(m, F16(0.8), 1023.5)
    return (np.maximum(r, F16(1024.0)).astype(np.float64) - 1024.0).astype(F16)


def scale_generic(m):
    """round(0.8f * x): HFMA2(x, 0.8h, 1024) - 1024."""
    return (fma16(m, F16(0.8), 1024.0).astype(np.float64) - 1024.0).astype(F16)


def decode(graph_rows, Z, llrs, K, n_layers, max_iter, generic=False, check=None):
    """graph_rows: list of rows, each a list of (col, shift). llrs: int8 input (already de-matched), zero beyond.
    Returns soft values (float16, natural order) after each iteration through `check(it, soft)` -> stop flag."""
    n_cols = max(c for row in graph_rows[:n_layers] for c, _ in row) + 1
    soft = np.zeros((n_cols, Z), F16)
    x = np.clip(llrs.astype(np.int32), -64, 64)
    for n in range(min(x.size, (n_cols - 2) * Z)):
        soft[2 + n // Z, n % Z] = x[n]
    c2v = [np.zeros((len(row), Z), F16) for row in graph_rows[:n_layers]]
    sign = lambda a: np.signbit(a)
    for it in range(max_iter):
        for m in range(n_layers):
            row = graph_rows[m]
            u = np.zeros((len(row), Z), F16)
            for k, (c, s) in enumerate(row):
                S = np.roll(soft[c], -s)  # S[i] = soft[c][(i + s) mod Z]
                with np.errstate(invalid="ignore"):
                    u[k] = S - c2v[m][k]
            a = np.abs(u)
            # two smallest magnitudes capped at 120 (the kernel finds them with a min / max tree; same multiset result)
            srt = np.sort(np.vstack([a.astype(np.float32), np.full((2, Z), 120.0, np.float32)]), axis=0)
            m1, m2 = srt[0].astype(F16), srt[1].astype(F16)
            par = np.zeros(Z, bool)
            for k in range(len(row)):
                par ^= sign(u[k])
            s1 = scale_generic(m1) if generic else scale_simd(m1)
            s2 = scale_generic(m2) if generic else scale_simd(m2)
            for k, (c, s) in enumerate(row):
                # sign-normalised form of pdc_kernel_h2.cuh::h2_row pass 2
                mag = np.where(a[k] == m1, s2, s1)
                ms = np.where(par, -mag, mag).astype(F16)                       # magnitude of the new message, row-parity sign
                sg = np.where(sign(u[k]), F16(-1.0), F16(1.0)).astype(F16)       # copysign(1, u)
                with np.errstate(invalid="ignore"):
                    qa = np.clip(a[k].astype(np.float64) - 336.0, 0.0, 1.0).astype(F16)  # 1 iff u is infinite
                w0 = fma16(qa, F16(65504.0), np.minimum(a[k], F16(120.0)))                # min(|u|, 120), +inf iff u is infinite
                with np.errstate(over="ignore"):
                    w = (w0.astype(np.float64) + ms.astype(np.float64)).astype(F16)
                qb = np.clip(w.astype(np.float64) - 120.0, 0.0, 1.0).astype(F16)          # 1 iff w > 120 (also when w is +inf)
                sm = fma16(qb, F16(65504.0), w)                                            # w, or +inf: 121 + 65504 rounds to inf
                cn = (ms.astype(np.float64) * sg.astype(np.float64)).astype(F16)          # may be -0: harmless
                # the kernel adds the storage bias in the same fused operation, which turns a -0 product into +0
                Snew = (fma16(sm, sg, 1152.0).astype(np.float64) - 1152.0).astype(F16)
                c2v[m][k] = cn
                soft[c] = np.roll(Snew, s)
        if check is not None and check(it + 1, soft):
            return it + 1, soft
    return None, soft
