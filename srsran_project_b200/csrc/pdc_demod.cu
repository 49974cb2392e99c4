// LLR producers in front of the code-block decode chain: soft demodulation of equalised symbols into
// int8 log-likelihood ratios and descrambling, one kernel, LLRs born in HBM where pdc_decode_device reads them.
//
// What it replaces in the reference (bit-exact against the AVX512 build of those files, oracle/pdc_oracle_demod.c):
//   demodulation_mapper_impl::demodulate_soft   lib/phy/upper/channel_modulation/demodulation_mapper_impl.cpp:79-106
//     per modulation                            demodulation_mapper_{qpsk,qam16,qam64,qam256}.cpp, avx512_helpers.h, avx2_helpers.h
//   revert_scrambling + pseudo_random_generator lib/phy/upper/channel_processors/pusch/pusch_demodulator_impl.cpp:60-222,261-262,401-405
//
// A segment is ONE call of demodulate_soft (in the reference: the resource elements of one OFDM symbol, or the part of it
// that fits the codeword buffer's block, pusch_demodulator_impl.cpp:370-395). That matters for parity: the reference runs
// AVX512 batches, then AVX2 batches, then single symbols, and the three loops round differently (fused or separate
// multiply-add, reciprocal or division, per-component or per-symbol near-zero test, half-even or half-away rounding), so
// the arithmetic a symbol gets depends on its index within the call and on the call's length. Every loop is restated here.
//
// This translation unit is compiled WITHOUT --use_fast_math and with -fmad=false: every operation is the IEEE
// single-precision one the host performs (subnormals included), and a fused multiply-add appears exactly where the
// compiled reference has one (written as fmaf).
//
// Mapping: persistent CTAs walk the list of 1024-symbol chunks (a chunk lies inside one segment): eight warps with 4 consecutive symbols per thread (4 Qm output
// bytes, one aligned vector store). Bound: 20 B of HBM traffic per 256QAM symbol against ~150 issue slots — the kernel sits
// between the HBM and the issue roofline.
//
// Scrambling sequence (TS 38.211 5.2.1) without a serial walk: a ninth warp jumps both 31-bit LFSRs to the chunk's first bit with
// precomputed powers of the transition matrix (binary expansion of the position: one 31-lane parity + ballot per set bit),
// every lane then jumps a further 32 l bits to produce word l, and the remaining words follow from the word-level
// recurrence W[j + 31] = W[j + 3] ^ W[j] (the characteristic polynomial over GF(2) commutes with squaring, so what holds
// for bits at distance 1 holds for words at distance 32). The eight arithmetic warps meet it at one barrier before the stores.
#include "../../include/pusch_dec_cuda.h"
#include "pdc_device.cuh"
#include <cmath>
#include <cstdint>
#include <cuda_runtime.h>

#define PDC_DEMOD_CHUNK 1024 /* symbols per CTA */
#define PDC_DEMOD_THREADS 288 /* eight arithmetic warps (4 symbols per thread) + one warp that generates the scrambling words */
#define PDC_DEMOD_JUMP_BITS 28

struct pdc_demod_tables {
  float2   interval[7][16];                      // (slope, intercept): 64QAM tables 0..2, 256QAM tables 3..6
  float    inv_width[7];                         // 1 / interval width
  float    width[7];
  float    half_n[7];                            // nof_intervals / 2
  int      n_minus_1[7];
  float    n_minus_1_f[7];
  float    qam16_gain, qam16_thr;                // 4 / sqrt(10), 2 / sqrt(10)
};
__constant__ pdc_demod_tables c_demod;
// Rows of M^(2^j) for both LFSRs (jump) and of M^(32 l), [lfsr][row][l] (word bootstrap).
__device__ uint32_t d_prg_jump[2][PDC_DEMOD_JUMP_BITS][32];
__device__ uint32_t d_prg_word[2][31][32];

static constexpr float NEAR_ZERO = 1e-9f;

// ---- quantisers (NaN -> 0 in all three) ------------------------------------------------------------------------------------
// mm512::quantize_ps: scale (one rounding), clip to +-120, convert with round-half-even. Clipping after the conversion is the
// same function, and cvt.rni maps NaN to 0 as the masked multiply does.
__device__ __forceinline__ int q_avx512(float v, float scale)
{
  const int i = __float2int_rn(__fmul_rn(scale, v));
  return max(min(i, 120), -120);
}
// mm256::quantize_ps: the same function (scale, clip, round-half-even, NaN -> out of bounds -> 0).
__device__ __forceinline__ int q_avx2(float v, float scale)
{
  return q_avx512(v, scale);
}
// log_likelihood_ratio::quantize: clip, divide by the range, times 120, round half away from zero.
__device__ __forceinline__ int q_scalar(float v, float range)
{
  float c = v;
  if (fabsf(v) > range) {
    c = copysignf(range, v);
  }
  const float r = roundf(__fmul_rn(__fdiv_rn(c, range), 120.0f));
  return (r != r) ? 0 : (int)r;
}

// ---- interval functions -----------------------------------------------------------------------------------------------------
// x86 cvtps2dq gives INT_MIN for NaN and for anything outside int32; cvt.rmi saturates instead. Only the positive side
// differs after the clamp to [0, n - 1] (INT_MAX is never a legitimate result: floats that large are multiples of 128).
__device__ __forceinline__ int interval_index(float s, int n_minus_1)
{
  const int i = __float2int_rd(s);
  return (i == 0x7fffffff) ? 0 : max(min(i, n_minus_1), 0);
}
__device__ __forceinline__ float interval_avx512(float v, float rcp, int t, const float2 (*tab)[16])
{
  const int    i  = interval_index(fmaf(v, c_demod.inv_width[t], c_demod.half_n[t]), c_demod.n_minus_1[t]);
  const float2 si = tab[t][i];
  const float  r  = __fmul_rn(fmaf(si.x, v, si.y), rcp);
  return (fabsf(v) > NEAR_ZERO) ? r : 0.0f;
}
__device__ __forceinline__ float interval_avx2(float v, float rcp, int t, const float2 (*tab)[16])
{
  // floor(v * (1 / width)) converted, THEN + n / 2 as an integer.
  const int    f  = __float2int_rd(__fmul_rn(v, c_demod.inv_width[t]));
  const int    h  = (int)c_demod.half_n[t];
  const int    i  = (f == 0x7fffffff || f == (int)0x80000000) ? 0 : max(min(f + h, c_demod.n_minus_1[t]), 0);
  const float2 si = tab[t][i];
  const float  r  = __fmul_rn(fmaf(si.x, v, si.y), rcp); // the compiler fuses the reference's mul + add
  return (fabsf(v) <= NEAR_ZERO) ? 0.0f : r;
}
__device__ __forceinline__ float interval_scalar(float v, float rcp, int t, const float2 (*tab)[16])
{
  const int    f  = __float2int_rd(__fdiv_rn(v, c_demod.width[t]));
  const int    h  = (int)c_demod.half_n[t];
  const int    i  = (f == 0x7fffffff || f == (int)0x80000000) ? 0 : max(min(f + h, c_demod.n_minus_1[t]), 0);
  const float2 si = tab[t][i];
  return __fmul_rn(fmaf(si.x, v, si.y), rcp);
}

enum { PATH_SCALAR = 0, PATH_AVX2 = 1, PATH_AVX512 = 2 };

// ---- the common case of the AVX512 loop, without conversions or integer clamps -------------------------------------------------
// When |re|, |im| < 1e8 and the noise reciprocal is finite, nothing in mm512::interval_function / quantize_ps can saturate a
// conversion or produce a NaN, and both roundings can be done by the adder: floor(x) for 0 <= x < 2^22 is the low mantissa of
// x + 2^23 rounded DOWN, round-half-even(y) for |y| <= 120 is the low byte of y + 1.5 * 2^23 rounded to nearest. The clamps
// move in front of the roundings (they commute with them at integer bounds). Same results as interval_avx512 + q_avx512, with
// the work on the FMA pipe instead of the quarter-rate conversion unit and the integer min / max.
template <int T0, int NT>
__device__ __forceinline__ void demod_interval_fast(uint32_t (&w)[2 * NT], const float (&re)[4], const float (&im)[4],
                                                    const float (&rcp)[4], uint32_t tab_addr)
{
  uint32_t z[4][2 * NT]; // magic sums: the low byte is the LLR
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const float v[2]  = {re[s], im[s]};
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const bool live = fabsf(v[c]) > NEAR_ZERO;
#pragma unroll
      for (int k = 0; k < NT; ++k) {
        const int      t    = T0 + k;
        const float    sidx = fminf(fmaxf(fmaf(v[c], c_demod.inv_width[t], c_demod.half_n[t]), 0.0f), c_demod.n_minus_1_f[t]);
        const uint32_t bits = __float_as_uint(__fadd_rd(sidx, 8388608.0f));
        float2         si;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(si.x), "=f"(si.y) : "r"(tab_addr + 128u * t + ((bits & 15u) << 3)));
        float r = __fmul_rn(fmaf(si.x, v[c], si.y), rcp[s]);
        r       = live ? r : 0.0f;
        const float y = fminf(fmaxf(__fmul_rn(6.0f, r), -120.0f), 120.0f);
        z[s][2 * k + c] = __float_as_uint(__fadd_rn(y, 12582912.0f));
      }
    }
  }
#pragma unroll
  for (int word = 0; word < 2 * NT; ++word) {
    uint32_t b[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int byte = 4 * word + q;
      b[q]           = z[byte / (2 * NT)][byte % (2 * NT)];
    }
    w[word] = __byte_perm(__byte_perm(b[0], b[1], 0x0040), __byte_perm(b[2], b[3], 0x0040), 0x5410);
  }
}

// One symbol of 64QAM (T0 = 0, NT = 3) or 256QAM (T0 = 3, NT = 4): q[2 k], q[2 k + 1] = bits 2k, 2k + 1.
template <int T0, int NT>
__device__ __forceinline__ void demod_interval(int (&q)[8], float re, float im, float nv, int path,
                                               const float2 (*tab)[16])
{
  const float scale = 6.0f; // 120 / 20
  if (path == PATH_AVX512) {
    const float rcp = (nv > 0.0f) ? __fdiv_rn(1.0f, nv) : 0.0f;
#pragma unroll
    for (int k = 0; k < NT; ++k) {
      q[2 * k]     = q_avx512(interval_avx512(re, rcp, T0 + k, tab), scale);
      q[2 * k + 1] = q_avx512(interval_avx512(im, rcp, T0 + k, tab), scale);
    }
  } else if (path == PATH_AVX2) {
    const float rcp = (nv > 0.0f) ? __fdiv_rn(1.0f, nv) : 0.0f;
#pragma unroll
    for (int k = 0; k < NT; ++k) {
      q[2 * k]     = q_avx2(interval_avx2(re, rcp, T0 + k, tab), scale);
      q[2 * k + 1] = q_avx2(interval_avx2(im, rcp, T0 + k, tab), scale);
    }
  } else {
    const bool  nz  = fmaf(re, re, __fmul_rn(im, im)) < NEAR_ZERO; // is_near_zero(cf_t), fused by the compiler
    const float rcp = (nv > 0.0f) ? __fdiv_rn(1.0f, nv) : 0.0f;
#pragma unroll
    for (int k = 0; k < NT; ++k) {
      q[2 * k]     = nz ? 0 : q_scalar(interval_scalar(re, rcp, T0 + k, tab), 20.0f);
      q[2 * k + 1] = nz ? 0 : q_scalar(interval_scalar(im, rcp, T0 + k, tab), 20.0f);
    }
  }
}

__device__ __forceinline__ int demod_bpsk(float re, float im, float nv)
{
  if (!(nv > 0.0f)) {
    return 0;
  }
  const float gain = 2.0f * 1.41421356237309504880f;
  return q_scalar(__fdiv_rn(__fmul_rn(gain, __fadd_rn(re, im)), nv), 24.0f);
}
__device__ __forceinline__ void demod_qpsk(int (&q)[8], float re, float im, float nv, int path)
{
  const float gain = 2.0f * 1.41421356237309504880f;
  if (path == PATH_AVX2) {
    const float rcp = (nv > 0.0f) ? __fdiv_rn(1.0f, nv) : 0.0f;
    q[0]            = q_avx2(__fmul_rn(__fmul_rn(gain, re), rcp), 5.0f);
    q[1]            = q_avx2(__fmul_rn(__fmul_rn(gain, im), rcp), 5.0f);
  } else if (!(nv > 0.0f)) {
    q[0] = q[1] = 0;
  } else {
    q[0] = q_scalar(__fdiv_rn(__fmul_rn(gain, re), nv), 24.0f);
    q[1] = q_scalar(__fdiv_rn(__fmul_rn(gain, im), nv), 24.0f);
  }
}
__device__ __forceinline__ void demod_qam16(int (&q)[8], float re, float im, float nv, int path)
{
  const float gain = c_demod.qam16_gain, thr = c_demod.qam16_thr;
  const float x[2] = {re, im};
  if (path == PATH_AVX2) {
    const float rcp = (nv > 0.0f) ? __fdiv_rn(1.0f, nv) : 0.0f;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const float first  = __fmul_rn(gain, x[c]);
      const float second = __fsub_rn(__fmul_rn(2.0f, first), copysignf(0.8f, x[c]));
      float       l01    = (fabsf(x[c]) > thr) ? second : first;
      float       l23    = __fsub_rn(0.8f, fabsf(first));
      l01 = __fmul_rn(l01, rcp), l23 = __fmul_rn(l23, rcp);
      if (fabsf(x[c]) <= NEAR_ZERO) {
        l01 = 0.0f, l23 = 0.0f;
      }
      q[c]     = q_avx2(l01, 6.0f);
      q[2 + c] = q_avx2(l23, 6.0f);
    }
    return;
  }
  const bool nz = fmaf(re, re, __fmul_rn(im, im)) < NEAR_ZERO;
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    if (nz || !(nv > 0.0f)) {
      q[c] = q[2 + c] = 0;
      continue;
    }
    float l = __fmul_rn(gain, x[c]);
    if (fabsf(x[c]) > thr) {
      l = __fsub_rn(__fmul_rn(2.0f, l), copysignf(0.8f, x[c]));
    }
    q[c]     = q_scalar(__fdiv_rn(l, nv), 20.0f);
    q[2 + c] = q_scalar(__fdiv_rn(fmaf(-gain, fabsf(x[c]), 0.8f), nv), 20.0f); // 0.8 - gain |x|, fused by the compiler
  }
}

// ---- scrambling sequence ----------------------------------------------------------------------------------------------------
// State s of an LFSR at position p: bit i = x(p + i), 31 bits. Returns M^pos s, computed by the 31 low lanes of a warp.
__device__ __forceinline__ uint32_t prg_jump(const uint32_t (*jump)[PDC_DEMOD_JUMP_BITS][32], uint32_t s, uint32_t pos,
                                             int lfsr, int lane)
{
  for (int j = 0; j < PDC_DEMOD_JUMP_BITS && (pos >> j) != 0; ++j) {
    if ((pos >> j) & 1u) {
      const uint32_t row = jump[lfsr][j][lane];
      s                  = __ballot_sync(0xffffffffu, (__popc(row & s) & 1) != 0) & 0x7fffffffu;
    }
  }
  return s;
}
// Word l of the sequence that starts at the position of state s: bits 32 l .. 32 l + 31.
__device__ __forceinline__ uint32_t prg_word(const uint32_t (*word)[31][32], uint32_t s, int lfsr, int lane)
{
  uint32_t t = 0;
#pragma unroll
  for (int r = 0; r < 31; ++r) {
    t |= (uint32_t)(__popc(word[lfsr][r][lane] & s) & 1) << r;
  }
  const uint32_t taps = lfsr ? 0xfu : 0x9u;
  return t | ((uint32_t)(__popc(t & taps) & 1) << 31);
}
// What the sequence warp remembers between chunks: w1 / w2 hold `n` words of the two LFSR sequences from bit `start` on.
struct prg_history {
  uint32_t c_init, start;
  int      n;
};
// Fills seq[0 .. nwords) with the scrambling sequence from bit `pos` on (LSB of a word first); one warp. If the words kept
// from the previous chunk reach 31 words past `pos` (the next chunk of the same codeword: consecutive chunks, and consecutive
// OFDM symbols, continue the sequence), the recurrence simply carries on from them; otherwise both LFSRs jump.
__device__ __forceinline__ void prg_generate(const uint32_t (*jump)[PDC_DEMOD_JUMP_BITS][32], const uint32_t (*word)[31][32],
                                             uint32_t* seq, uint32_t* w1, uint32_t* w2, int nwords, uint32_t c_init,
                                             uint32_t pos, int lane, prg_history& h)
{
  const uint32_t d = pos - h.start;
  if (h.n != 0 && h.c_init == c_init && pos >= h.start && (d & 31u) == 0 && (int)(d >> 5) + 32 <= h.n) {
    const int      k = (int)(d >> 5);
    const uint32_t a = w1[k + lane], b = w2[k + lane];
    __syncwarp();
    w1[lane] = a, w2[lane] = b;
  } else {
    const uint32_t s1 = prg_jump(jump, 1u, 1600u + pos, 0, lane);
    const uint32_t s2 = prg_jump(jump, c_init & 0x7fffffffu, 1600u + pos, 1, lane);
    w1[lane] = prg_word(word, s1, 0, lane);
    w2[lane] = prg_word(word, s2, 1, lane);
  }
  __syncwarp();
  // 32 words more than the chunk needs: the seed of the next one.
  const int total = nwords + 32;
  for (int base = 32; base < total; base += 28) {
    if (lane < 28) {
      const int j   = base + lane - 31;
      w1[base + lane] = w1[j + 3] ^ w1[j];
      w2[base + lane] = w2[j + 3] ^ w2[j + 2] ^ w2[j + 1] ^ w2[j];
    }
    __syncwarp();
  }
  for (int j = lane; j < nwords; j += 32) {
    seq[j] = w1[j] ^ w2[j];
  }
  h.c_init = c_init, h.start = pos;
  h.n      = 32 + ((total - 32 + 27) / 28) * 28;
}

// Negates the bytes of x whose bit in the low nibble of `bits` is set (byte k <-> bit k): -x = (x ^ 0xff) + 1 per byte.
__device__ __forceinline__ uint32_t negate_bytes(uint32_t x, uint32_t bits)
{
  const uint32_t one = ((bits & 0xfu) * 0x00204081u) & 0x01010101u;
  const uint32_t a   = x ^ (one * 0xffu);
  return ((a & 0x7f7f7f7fu) + one) ^ (a & 0x80808080u);
}

// Named barriers of the producer / consumer pipeline: ids 1, 2 = "sequence buffer b is full", 3, 4 = "buffer b is free again".
__device__ __forceinline__ void bar_sync(int id)
{
  asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(PDC_DEMOD_THREADS) : "memory");
}
__device__ __forceinline__ void bar_arrive(int id)
{
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "n"(PDC_DEMOD_THREADS) : "memory");
}

struct pdc_demod_chunk {
  uint32_t seg;
  uint32_t first; // first symbol of the chunk within its segment
};

// Arithmetic first, one CTA barrier (the scrambling words of warp 0 are complete), then sign flips and stores: warp 0's
// sequence generation overlaps the other warps' arithmetic instead of preceding it.
template <int MOD>
__device__ __forceinline__ void demod_chunk(const pdc_demod_seg& sg, uint32_t first, const float2* __restrict__ symbols,
                                            const float* __restrict__ noise, int8_t* __restrict__ llrs,
                                            const float2 (*tab)[16], const uint32_t* seq, bool descramble, int tid, int buf)
{
  constexpr int QM   = (MOD == 0) ? 1 : MOD;
  constexpr int B512 = (MOD == 6) ? 32 : ((MOD == 8) ? 8 : 0);
  constexpr int B256 = (MOD == 2) ? 16 : ((MOD == 4) ? 8 : ((MOD == 6) ? 16 : ((MOD == 8) ? 4 : 0)));
  const uint32_t nsym = sg.nof_symbols;
  const uint32_t end512 = B512 ? (nsym / B512) * B512 : 0u;
  const uint32_t end256 = B256 ? (nsym / B256) * B256 : 0u;
  const uint32_t i0   = first + 4u * (uint32_t)tid;
  const uint32_t nvalid = (i0 < nsym) ? min(4u, nsym - i0) : 0u;
  const size_t   g      = (size_t)sg.sym_offset + i0;

  float re[4], im[4], nv[4];
  if (nvalid == 4 && (g & 3) == 0 && ((reinterpret_cast<uintptr_t>(symbols) | reinterpret_cast<uintptr_t>(noise)) & 15) == 0) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(symbols + g));
    const float4 b = __ldg(reinterpret_cast<const float4*>(symbols + g) + 1);
    const float4 n = __ldg(reinterpret_cast<const float4*>(noise + g));
    re[0] = a.x, im[0] = a.y, re[1] = a.z, im[1] = a.w, re[2] = b.x, im[2] = b.y, re[3] = b.z, im[3] = b.w;
    nv[0] = n.x, nv[1] = n.y, nv[2] = n.z, nv[3] = n.w;
  } else {
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const bool   ok = (uint32_t)s < nvalid;
      const float2 z  = ok ? __ldg(symbols + g + s) : make_float2(0.f, 0.f);
      re[s] = z.x, im[s] = z.y, nv[s] = ok ? __ldg(noise + g + s) : 0.f;
    }
  }

  // 4 Qm bytes, symbol s at bytes s Qm .. s Qm + Qm - 1.
  uint32_t w[QM];
#pragma unroll
  for (int k = 0; k < QM; ++k) {
    w[k] = 0;
  }
  bool done = false;
  if constexpr (MOD == 6 || MOD == 8) {
    if (nvalid == 4 && i0 + 3 < end512) {
      float rcp[4];
      bool  ok = true;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        rcp[s] = (nv[s] > 0.0f) ? __fdiv_rn(1.0f, nv[s]) : 0.0f;
        ok     = ok && (fabsf(re[s]) < 1e8f) && (fabsf(im[s]) < 1e8f) && (rcp[s] < 3.0e38f);
      }
      if (ok) {
        demod_interval_fast<(MOD == 6) ? 0 : 3, QM / 2>(w, re, im, rcp, (uint32_t)__cvta_generic_to_shared(tab));
        done = true;
      }
    }
  }
#pragma unroll
  for (int s = 0; s < 4 && !done; ++s) {
    const uint32_t i    = i0 + s;
    const int      path = (B512 != 0 && i < end512) ? PATH_AVX512 : ((B256 != 0 && i < end256) ? PATH_AVX2 : PATH_SCALAR);
    int            q[8];
    if constexpr (MOD == 0) {
      q[0] = (i & 1u) ? demod_bpsk(im[s], -re[s], nv[s]) : demod_bpsk(re[s], im[s], nv[s]);
    } else if constexpr (MOD == 1) {
      q[0] = demod_bpsk(re[s], im[s], nv[s]);
    } else if constexpr (MOD == 2) {
      demod_qpsk(q, re[s], im[s], nv[s], path);
    } else if constexpr (MOD == 4) {
      demod_qam16(q, re[s], im[s], nv[s], path);
    } else if constexpr (MOD == 6) {
      demod_interval<0, 3>(q, re[s], im[s], nv[s], path, tab);
    } else {
      demod_interval<3, 4>(q, re[s], im[s], nv[s], path, tab);
    }
#pragma unroll
    for (int b = 0; b < QM; ++b) {
      const int byte = s * QM + b;
      w[byte >> 2] |= ((uint32_t)q[b] & 0xffu) << (8 * (byte & 3));
    }
  }

  bar_sync(1 + buf); // the producer warp has filled seq
  const uint32_t b0 = 4u * QM * (uint32_t)tid;
  const uint32_t lo = seq[b0 >> 5], hi = seq[(b0 >> 5) + 1];
  bar_arrive(3 + buf); // ... and this thread has read its part of it
  if (nvalid == 0) {
    return;
  }
  if (descramble) {
    const uint32_t c  = __funnelshift_r(lo, hi, b0 & 31u);
#pragma unroll
    for (int k = 0; k < QM; ++k) {
      w[k] = negate_bytes(w[k], c >> (4 * k));
    }
  }

  int8_t* dst = llrs + (size_t)sg.llr_offset + (size_t)i0 * QM;
  const uintptr_t a = reinterpret_cast<uintptr_t>(dst);
  if (nvalid == 4 && (a & 3) == 0) {
    if constexpr (QM % 4 == 0) {
      if ((a & 15) == 0) {
#pragma unroll
        for (int k = 0; k < QM; k += 4) {
          *reinterpret_cast<uint4*>(dst + 4 * k) = make_uint4(w[k], w[k + 1], w[k + 2], w[k + 3]);
        }
        return;
      }
    }
    if constexpr (QM % 2 == 0) {
      if ((a & 7) == 0) {
#pragma unroll
        for (int k = 0; k < QM; k += 2) {
          *reinterpret_cast<uint2*>(dst + 4 * k) = make_uint2(w[k], w[k + 1]);
        }
        return;
      }
    }
#pragma unroll
    for (int k = 0; k < QM; ++k) {
      *reinterpret_cast<uint32_t*>(dst + 4 * k) = w[k];
    }
    return;
  }
#pragma unroll
  for (int k = 0; k < QM; ++k) {
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      if ((uint32_t)(4 * k + b) < nvalid * QM) {
        dst[4 * k + b] = (int8_t)(w[k] >> (8 * b));
      }
    }
  }
}

// Persistent CTAs (as many as fit the device at once), each walking a contiguous range of the chunk list. The ninth warp runs
// one chunk AHEAD of the eight arithmetic warps through two sequence buffers, so the latency of its dependent chain (jump,
// bootstrap, recurrence) overlaps the arithmetic of the previous chunk instead of adding to every chunk's lifetime.
__global__ void __launch_bounds__(PDC_DEMOD_THREADS, 4)
    pdc_demod_kernel(const pdc_demod_seg* __restrict__ segs, const pdc_demod_chunk* __restrict__ chunks, uint32_t nchunks,
                     const float2* __restrict__ symbols, const float* __restrict__ noise, int8_t* __restrict__ llrs)
{
  __shared__ float2   s_tab[7][16];
  __shared__ uint32_t s_seq[2][260], s_w1[320], s_w2[320];
  // The jump matrices live in shared memory for the lifetime of the persistent CTA: the dependent chain of the sequence warp
  // is one load per step, and from HBM-streaming L1 / L2 that load was the whole cost of a chunk (8.5 us per chunk measured).
  __shared__ uint32_t s_jump[2][PDC_DEMOD_JUMP_BITS][32], s_word[2][31][32];
  const int           tid = threadIdx.x;
  if (tid < 7 * 16) {
    s_tab[tid >> 4][tid & 15] = c_demod.interval[tid >> 4][tid & 15];
  }
  for (int q = tid; q < 2 * PDC_DEMOD_JUMP_BITS * 32; q += PDC_DEMOD_THREADS) {
    (&s_jump[0][0][0])[q] = (&d_prg_jump[0][0][0])[q];
  }
  for (int q = tid; q < 2 * 31 * 32; q += PDC_DEMOD_THREADS) {
    (&s_word[0][0][0])[q] = (&d_prg_word[0][0][0])[q];
  }
  __syncthreads();
  // A contiguous range of the chunk list per CTA: consecutive chunks of a codeword continue the scrambling sequence.
  const uint32_t per = nchunks / gridDim.x, rem = nchunks % gridDim.x;
  const uint32_t c0  = blockIdx.x * per + min(blockIdx.x, rem), c1 = c0 + per + (blockIdx.x < rem ? 1u : 0u);
  prg_history    hist;
  hist.c_init = 0, hist.start = 0, hist.n = 0;
  int it = 0;
  for (uint32_t c = c0; c < c1; ++c, ++it) {
    const int             buf = it & 1;
    const pdc_demod_chunk ch  = chunks[c];
    const pdc_demod_seg   sg  = segs[ch.seg];
    const int             qm  = (sg.mod == 0) ? 1 : sg.mod;
    const bool            descramble = (sg.flags & PDC_DEMOD_F_DESCRAMBLE) != 0;
    if (tid >= 256) {
      if (it >= 2) {
        bar_sync(3 + buf); // every arithmetic thread has read the words of two chunks ago
      }
      if (descramble) {
        // 1024 Qm sequence bits, one word more for the unaligned window of the last thread.
        prg_generate(s_jump, s_word, s_seq[buf], s_w1, s_w2, 32 * qm + 1, sg.c_init, sg.scramble_offset + ch.first * (uint32_t)qm, tid - 256, hist);
      }
      bar_arrive(1 + buf);
      continue;
    }
    switch (sg.mod) {
      case 0:
        demod_chunk<0>(sg, ch.first, symbols, noise, llrs, s_tab, s_seq[buf], descramble, tid, buf);
        break;
      case 1:
        demod_chunk<1>(sg, ch.first, symbols, noise, llrs, s_tab, s_seq[buf], descramble, tid, buf);
        break;
      case 2:
        demod_chunk<2>(sg, ch.first, symbols, noise, llrs, s_tab, s_seq[buf], descramble, tid, buf);
        break;
      case 4:
        demod_chunk<4>(sg, ch.first, symbols, noise, llrs, s_tab, s_seq[buf], descramble, tid, buf);
        break;
      case 6:
        demod_chunk<6>(sg, ch.first, symbols, noise, llrs, s_tab, s_seq[buf], descramble, tid, buf);
        break;
      default:
        demod_chunk<8>(sg, ch.first, symbols, noise, llrs, s_tab, s_seq[buf], descramble, tid, buf);
        break;
    }
  }
}

#include "pdc_kernel_demux.cuh"

// ---- host side ---------------------------------------------------------------------------------------------------------------
namespace {
struct gf2_mat {
  uint32_t row[31];
};
gf2_mat gf2_mul(const gf2_mat& a, const gf2_mat& b) // (a b) s = a (b s)
{
  gf2_mat c;
  for (int r = 0; r != 31; ++r) {
    uint32_t acc = 0;
    for (int k = 0; k != 31; ++k) {
      if ((a.row[r] >> k) & 1u) {
        acc ^= b.row[k];
      }
    }
    c.row[r] = acc;
  }
  return c;
}
gf2_mat gf2_step(int lfsr) // one step: s'[r] = s[r + 1]; s'[30] = taps . s
{
  gf2_mat m;
  for (int r = 0; r != 30; ++r) {
    m.row[r] = 1u << (r + 1);
  }
  m.row[30] = lfsr ? 0xfu : 0x9u;
  return m;
}
} // namespace

// Host copy of the jump matrices: the demultiplexer's planner jumps the two registers on the host (a few hundred word operations
// per OFDM symbol that holds placeholder bits), so no thread on the device walks that dependent chain.
static uint32_t h_prg_jump[2][PDC_DEMOD_JUMP_BITS][32];
static bool     h_prg_jump_ready = false;

// States of the two LFSRs of TS 38.211 5.2.1 at sequence position pos (bit i = x(1600 + pos + i)); false before the tables exist.
extern "C" bool pdc_prg_states(uint32_t c_init, uint32_t pos, uint32_t* s1, uint32_t* s2)
{
  if (!h_prg_jump_ready) {
    return false;
  }
  uint32_t s[2] = {1u, c_init & 0x7fffffffu};
  for (int lfsr = 0; lfsr != 2; ++lfsr) {
    const uint32_t p = 1600u + pos;
    for (int j = 0; j < PDC_DEMOD_JUMP_BITS && (p >> j) != 0; ++j) {
      if ((p >> j) & 1u) {
        uint32_t t = 0;
        for (int r = 0; r != 31; ++r) {
          t |= (uint32_t)(__builtin_popcount(h_prg_jump[lfsr][j][r] & s[lfsr]) & 1) << r;
        }
        s[lfsr] = t;
      }
    }
  }
  *s1 = s[0], *s2 = s[1];
  return true;
}

extern "C" cudaError_t pdc_demod_init_device()
{
  static pdc_demod_tables t;
  const float             u10 = 1.0f / std::sqrt(10.0f), u42 = 1.0f / std::sqrt(42.0f), u170 = 1.0f / std::sqrt(170.0f);
  // Generating integers of the tables of demodulation_mapper_qam64.cpp:44-84 and demodulation_mapper_qam256.cpp:44-164:
  // slope = k / sqrt(M) (one single-precision product), intercept = num / den (one single-precision quotient).
  static const int slope_k[7][16] = {{16, 12, 8, 4, 4, 8, 12, 16},
                                     {8, 4, 4, 8, -8, -4, -4, -8},
                                     {4, -4, 4, -4},
                                     {32, 28, 24, 20, 16, 12, 8, 4, 4, 8, 12, 16, 20, 24, 28, 32},
                                     {16, 12, 8, 4, 4, 8, 12, 16, -16, -12, -8, -4, -4, -8, -12, -16},
                                     {8, 4, 4, 8, -8, -4, -4, -8, 8, 4, 4, 8, -8, -4, -4, -8},
                                     {4, -4, 4, -4, 4, -4, 4, -4}};
  static const int icpt_n[7][16]  = {{24, 12, 4, 0, 0, -4, -12, -24},
                                     {20, 8, 8, 12, 12, 8, 8, 20},
                                     {12, -4, -4, 12},
                                     {112, 84, 60, 40, 24, 12, 4, 0, 0, -4, -12, -24, -40, -60, -84, -112},
                                     {88, 60, 36, 16, 16, 28, 36, 40, 40, 36, 28, 16, 16, 36, 60, 88},
                                     {52, 24, 24, 44, -20, -8, -8, -12, -12, -8, -8, -20, 44, 24, 24, 52},
                                     {28, -20, 12, -4, -4, 12, -20, 28}};
  static const int nint[7] = {8, 8, 4, 16, 16, 16, 8}, wmul[7] = {2, 2, 4, 2, 2, 2, 4};
  for (int k = 0; k != 7; ++k) {
    const float unit = (k < 3) ? u42 : u170, den = (k < 3) ? 21.0f : 85.0f;
    t.width[k]       = (float)wmul[k] * unit;
    t.inv_width[k]   = 1.0f / t.width[k];
    t.half_n[k]      = (float)(nint[k] / 2);
    t.n_minus_1[k]   = nint[k] - 1;
    t.n_minus_1_f[k] = (float)(nint[k] - 1);
    for (int i = 0; i != 16; ++i) {
      t.interval[k][i] = (i < nint[k]) ? make_float2((float)slope_k[k][i] * unit, (float)icpt_n[k][i] / den)
                                       : make_float2(0.f, 0.f);
    }
  }
  t.qam16_gain = 4.0f * u10, t.qam16_thr = 2.0f * u10;
  cudaError_t e = cudaMemcpyToSymbol(c_demod, &t, sizeof(t));
  if (e != cudaSuccess) {
    return e;
  }
  static uint32_t word[2][31][32];
  uint32_t(&jump)[2][PDC_DEMOD_JUMP_BITS][32] = h_prg_jump;
  for (int lfsr = 0; lfsr != 2; ++lfsr) {
    gf2_mat p = gf2_step(lfsr), m32;
    for (int j = 0; j != PDC_DEMOD_JUMP_BITS; ++j) {
      for (int r = 0; r != 32; ++r) {
        jump[lfsr][j][r] = (r < 31) ? p.row[r] : 0u;
      }
      if (j == 5) {
        m32 = p;
      }
      p = gf2_mul(p, p);
    }
    gf2_mat acc; // identity
    for (int r = 0; r != 31; ++r) {
      acc.row[r] = 1u << r;
    }
    for (int l = 0; l != 32; ++l) {
      for (int r = 0; r != 31; ++r) {
        word[lfsr][r][l] = acc.row[r];
      }
      acc = gf2_mul(m32, acc);
    }
  }
  h_prg_jump_ready = true;
  e                = cudaMemcpyToSymbol(d_prg_jump, jump, sizeof(jump));
  if (e != cudaSuccess) {
    return e;
  }
  return cudaMemcpyToSymbol(d_prg_word, word, sizeof(word));
}

extern "C" cudaError_t pdc_demod_launch(const pdc_demod_seg* d_segs, const void* d_chunks, uint32_t nchunks,
                                        const float* d_symbols, const float* d_noise, int8_t* d_llrs, cudaStream_t stream)
{
  if (nchunks == 0) {
    return cudaSuccess;
  }
  static int resident = 0; // persistent grid: CTAs that fit the device at once
  if (resident == 0) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pdc_demod_kernel, PDC_DEMOD_THREADS, 0);
    resident = sms * (per_sm > 0 ? per_sm : 1);
  }
  const uint32_t grid = nchunks < (uint32_t)resident ? nchunks : (uint32_t)resident;
  pdc_demod_kernel<<<grid, PDC_DEMOD_THREADS, 0, stream>>>(d_segs, reinterpret_cast<const pdc_demod_chunk*>(d_chunks), nchunks,
                                                           reinterpret_cast<const float2*>(d_symbols), d_noise, d_llrs);
  return cudaGetLastError();
}
