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
            m1 = np.full(Z, 120.0, F16)
            m2 = np.full(Z, 120.0, F16)
            par = np.zeros(Z, bool)
            for k in range(len(row)):
                m2 = np.minimum(m2, np.maximum(m1, a[k]))
                m1 = np.minimum(m1, a[k])
                par ^= sign(u[k])
            s1 = scale_generic(m1) if generic else scale_simd(m1)
            s2 = scale_generic(m2) if generic else scale_simd(m2)
            for k, (c, s) in enumerate(row):
                mag = np.where(a[k] == m1, s2, s1)
                neg = par ^ sign(u[k])
                cn = np.where(neg, -mag, mag).astype(F16)  # may be -0: harmless
                v = np.maximum(np.minimum(u[k], F16(120.0)), F16(-120.0))
                t = (v + cn).astype(F16)
                with np.errstate(invalid="ignore"):
                    xx = np.maximum(np.abs(t), (a[k].astype(np.float64) - 216.0).astype(F16))
                q = np.clip(xx.astype(np.float64) - 120.0, 0.0, 1.0).astype(F16)
                G = fma16(q, F16(65504.0), 1.0)
                with np.errstate(over="ignore"):
                    Snew = (t.astype(np.float64) * G.astype(np.float64)).astype(F16)
                c2v[m][k] = cn
                soft[c] = np.roll(Snew, s)
        if check is not None and check(it + 1, soft):
            return it + 1, soft
    return None, soft
