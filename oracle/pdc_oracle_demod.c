/* TEST INFRASTRUCTURE ONLY — not part of the product path (see pdc_oracle.h).
 *
 * Plain-C restatement of the LLR producers in front of the PUSCH code-block decode chain:
 *   - soft demodulation, demodulation_mapper_impl::demodulate_soft
 *       lib/phy/upper/channel_modulation/demodulation_mapper_impl.cpp:33-106 (BPSK, pi/2-BPSK, dispatch)
 *       lib/phy/upper/channel_modulation/demodulation_mapper_qpsk.cpp:38-171
 *       lib/phy/upper/channel_modulation/demodulation_mapper_qam16.cpp:43-275
 *       lib/phy/upper/channel_modulation/demodulation_mapper_qam64.cpp:44-463
 *       lib/phy/upper/channel_modulation/demodulation_mapper_qam256.cpp:44-425
 *       helpers avx512_helpers.h:45-264, avx2_helpers.h:38-272, demodulation_mapper_intervals.h:35-64,
 *       log_likelihood_ratio.cpp:90-99 (scalar quantiser)
 *   - descrambling, pusch_demodulator_impl.cpp:60-222 (revert_scrambling) over the TS 38.211 5.2.1 Gold sequence that
 *       lib/phy/upper/sequence_generators/pseudo_random_generator_impl.cpp generates (c_init = rnti 2^15 + n_id,
 *       pusch_demodulator_impl.cpp:261-262).
 *
 * The reference's result depends on WHICH vector loop a symbol falls into: one call processes its symbols in AVX512
 * batches, then AVX2 batches, then one by one, and the three loops round differently (fused multiply-add or not,
 * multiplication by a reciprocal or division, per-component or per-symbol near-zero test, round-half-even or
 * round-half-away). The flavour restated here is the reference built for an AVX512 host (-march=native there; the
 * explicit flags are in oracle/build_ref.sh), which is what tests/test_demod_oracle.py pins it against, bit for bit.
 * GCC contracts a * b + c into one fused operation wherever the reference writes it without intrinsics barriers
 * (C++ default -ffp-contract=fast): the contracted places are marked FUSED below and were established against the
 * compiled reference.
 *
 * Build: -O2 -ffp-contract=off (the Makefile passes it): every fused operation below is an explicit fmaf().
 */
#include "pdc_oracle.h"
#include <math.h>
#include <stdint.h>
#include <string.h>

static const float NEAR_ZERO = 1e-9f; /* math_utils.h:39 */
static const float SQRT2F    = 1.41421356237309504880f; /* M_SQRT2f32 */

/* Places where the compiler may have fused a * b + c in the compiled reference; bit i set = fused. Settable so that the
 * pin test can show that exactly one setting reproduces the compiled reference. */
unsigned orc_demod_fused_mask = 0x1f;
#define FUSED_SCALAR_INTERVAL 0x01u
#define FUSED_AVX2_INTERVAL 0x02u
#define FUSED_QAM16_SCALAR_23 0x04u
#define FUSED_ABS_SQ 0x08u
#define FUSED_QAM16_SECOND 0x10u

static float mul_add(float a, float b, float c, unsigned which)
{
  if (orc_demod_fused_mask & which) {
    return fmaf(a, b, c);
  }
  volatile float p = a * b;
  return p + c;
}

/* _mm512_cvtps_epi32 / _mm256_cvtps_epi32 under the default rounding mode: nearest even; NaN and out-of-range give
 * the "integer indefinite" value 0x80000000. */
static int32_t cvtps_epi32(float v)
{
  if (!(v >= -2147483648.0f && v < 2147483648.0f)) {
    return INT32_MIN;
  }
  return (int32_t)rintf(v);
}

static int clampi(int64_t v, int lo, int hi)
{
  return (int)(v < lo ? lo : (v > hi ? hi : v));
}

/* ---- quantisers ------------------------------------------------------------------------------------------------------ */
/* log_likelihood_ratio::quantize, log_likelihood_ratio.cpp:90-99. */
static int8_t quantize_scalar(float value, float range_limit)
{
  float clipped = value;
  if (fabsf(value) > range_limit) {
    clipped = copysignf(range_limit, value);
  }
  const float r = roundf(clipped / range_limit * 120.0f);
  if (isnan(r)) {
    return 0; /* static_cast<int8_t>(NaN): cvttss2si gives 0x80000000, whose low byte is 0 */
  }
  return (int8_t)(int32_t)r;
}
/* mm256::quantize_ps, avx2_helpers.h:122-168. */
static int8_t quantize_avx2(float value, float range_limit)
{
  const float scale = 120.0f / range_limit;
  value             = value * scale;
  if (value > 120.0f) {
    value = 120.0f;
  }
  if (value < -120.0f) {
    value = -120.0f;
  }
  value         = rintf(value);      /* _MM_FROUND_NINT; NaN stays NaN */
  int32_t i = cvtps_epi32(value);
  if (i > 120 || i < -120) {
    i = 0; /* check_bounds_epi32 */
  }
  return (int8_t)i;
}
/* mm512::quantize_ps / scale_clip_and_round_ps, avx512_helpers.h:45-102. */
static int8_t quantize_avx512(float value, float range_limit)
{
  const float scale = 120.0f / range_limit;
  value             = isnan(value) ? 0.0f : scale * value; /* maskz_mul_round: rounding mode of the product, not rint */
  value             = (value < 120.0f) ? value : 120.0f;   /* _mm512_min_ps */
  value             = (value > -120.0f) ? value : -120.0f; /* _mm512_max_ps */
  return (int8_t)cvtps_epi32(value);
}

/* ---- interval functions ---------------------------------------------------------------------------------------------- */
typedef struct {
  float    width;
  unsigned n;
  float    slope[16];
  float    intercept[16];
} interval_tab;

/* demodulation_mapper_intervals.h:35-64. */
static float interval_scalar(float value, float rcp_noise, const interval_tab* t)
{
  const float f   = floorf(value / t->width);
  int64_t     idx = (f >= -2147483648.0f && f < 2147483648.0f) ? (int64_t)(int32_t)f : (int64_t)INT32_MIN;
  idx += (int)t->n / 2;
  const int i = clampi(idx, 0, (int)t->n - 1);
  float     l = mul_add(t->slope[i], value, t->intercept[i], FUSED_SCALAR_INTERVAL); /* FUSED */
  return l * rcp_noise;
}
/* mm256::interval_function, avx2_helpers.h:175-258. */
static float interval_avx2(float value, float rcp_noise, const interval_tab* t)
{
  const float   s   = floorf(value * (1.0f / t->width));
  const int64_t idx = (int64_t)cvtps_epi32(s) + (int)t->n / 2; /* _mm256_add_epi32 wraps, but INT_MIN + n/2 stays negative */
  const int     i   = clampi(idx, 0, (int)t->n - 1);
  float         r   = mul_add(t->slope[i], value, t->intercept[i], FUSED_AVX2_INTERVAL) * rcp_noise; /* FUSED */
  if (fabsf(value) <= NEAR_ZERO) {
    r = 0.0f;
  }
  return r;
}
/* mm512::interval_function, avx512_helpers.h:134-246. */
static float interval_avx512(float value, float rcp_noise, const interval_tab* t)
{
  const float s = floorf(fmaf(value, 1.0f / t->width, (float)((int)t->n / 2)));
  const int   i = clampi((int64_t)cvtps_epi32(s), 0, (int)t->n - 1);
  float       r = fmaf(t->slope[i], value, t->intercept[i]) * rcp_noise;
  if (!(fabsf(value) > NEAR_ZERO)) {
    r = 0.0f;
  }
  return r;
}

static float safe_rcp(float noise)
{
  return (noise > 0.0f) ? 1.0f / noise : 0.0f;
}

/* Tables (demodulation_mapper_qam64.cpp:44-84, demodulation_mapper_qam256.cpp:44-164): slopes are k / sqrt(M) products in
 * single precision, intercepts are quotients of small integers; both are written here as the generating integers. */
static interval_tab tab64[3], tab256[4];
static int          tabs_ready;
static void         make_tab(interval_tab* t, float unit, int width_mult, unsigned n, const int* slope_k, const int* icpt_num,
                             float icpt_den)
{
  t->width = (float)width_mult * unit;
  t->n     = n;
  for (unsigned i = 0; i != 16; ++i) {
    t->slope[i] = t->intercept[i] = 0.0f;
  }
  for (unsigned i = 0; i != n; ++i) {
    t->slope[i]     = (float)slope_k[i] * unit;
    t->intercept[i] = (float)icpt_num[i] / icpt_den;
  }
}
static void init_tabs(void)
{
  if (tabs_ready) {
    return;
  }
  const float u42 = 1.0f / sqrtf(42.0f), u170 = 1.0f / sqrtf(170.0f);
  static const int s64_01[8] = {16, 12, 8, 4, 4, 8, 12, 16}, i64_01[8] = {24, 12, 4, 0, 0, -4, -12, -24};
  static const int s64_23[8] = {8, 4, 4, 8, -8, -4, -4, -8}, i64_23[8] = {20, 8, 8, 12, 12, 8, 8, 20};
  static const int s64_45[4] = {4, -4, 4, -4}, i64_45[4] = {12, -4, -4, 12};
  make_tab(&tab64[0], u42, 2, 8, s64_01, i64_01, 21.0f);
  make_tab(&tab64[1], u42, 2, 8, s64_23, i64_23, 21.0f);
  make_tab(&tab64[2], u42, 4, 4, s64_45, i64_45, 21.0f);
  static const int s256_01[16] = {32, 28, 24, 20, 16, 12, 8, 4, 4, 8, 12, 16, 20, 24, 28, 32};
  static const int i256_01[16] = {112, 84, 60, 40, 24, 12, 4, 0, 0, -4, -12, -24, -40, -60, -84, -112};
  static const int s256_23[16] = {16, 12, 8, 4, 4, 8, 12, 16, -16, -12, -8, -4, -4, -8, -12, -16};
  static const int i256_23[16] = {88, 60, 36, 16, 16, 28, 36, 40, 40, 36, 28, 16, 16, 36, 60, 88};
  static const int s256_45[16] = {8, 4, 4, 8, -8, -4, -4, -8, 8, 4, 4, 8, -8, -4, -4, -8};
  static const int i256_45[16] = {52, 24, 24, 44, -20, -8, -8, -12, -12, -8, -8, -20, 44, 24, 24, 52};
  static const int s256_67[8]  = {4, -4, 4, -4, 4, -4, 4, -4}, i256_67[8] = {28, -20, 12, -4, -4, 12, -20, 28};
  make_tab(&tab256[0], u170, 2, 16, s256_01, i256_01, 85.0f);
  make_tab(&tab256[1], u170, 2, 16, s256_23, i256_23, 85.0f);
  make_tab(&tab256[2], u170, 2, 16, s256_45, i256_45, 85.0f);
  make_tab(&tab256[3], u170, 4, 8, s256_67, i256_67, 85.0f);
  tabs_ready = 1;
}

enum { PATH_SCALAR = 0, PATH_AVX2 = 1, PATH_AVX512 = 2 };

static int near_zero_symbol(float re, float im)
{
  return mul_add(re, re, im * im, FUSED_ABS_SQ) < NEAR_ZERO; /* abs_sq, math_utils.h:80-84. FUSED */
}

static void demod_interval_symbol(int8_t* out, float re, float im, float noise, int path, const interval_tab* tabs,
                                  int ntab)
{
  const float range = 20.0f;
  if (path == PATH_SCALAR) {
    if (near_zero_symbol(re, im)) {
      memset(out, 0, 2 * (size_t)ntab);
      return;
    }
    const float rcp = (noise > 0.0f) ? 1.0f / noise : 0.0f;
    for (int k = 0; k != ntab; ++k) {
      out[2 * k]     = quantize_scalar(interval_scalar(re, rcp, &tabs[k]), range);
      out[2 * k + 1] = quantize_scalar(interval_scalar(im, rcp, &tabs[k]), range);
    }
    return;
  }
  const float rcp = safe_rcp(noise);
  for (int k = 0; k != ntab; ++k) {
    if (path == PATH_AVX2) {
      out[2 * k]     = quantize_avx2(interval_avx2(re, rcp, &tabs[k]), range);
      out[2 * k + 1] = quantize_avx2(interval_avx2(im, rcp, &tabs[k]), range);
    } else {
      out[2 * k]     = quantize_avx512(interval_avx512(re, rcp, &tabs[k]), range);
      out[2 * k + 1] = quantize_avx512(interval_avx512(im, rcp, &tabs[k]), range);
    }
  }
}

/* demodulation_mapper_impl.cpp:33-41. */
static int8_t demod_bpsk(float re, float im, float noise)
{
  if (!(noise > 0.0f)) {
    return 0;
  }
  const float l = 2.0f * SQRT2F * (re + im) / noise;
  return quantize_scalar(l, 24.0f);
}

static void demod_qpsk_symbol(int8_t* out, float re, float im, float noise, int path)
{
  const float gain = 2.0f * SQRT2F;
  if (path == PATH_AVX2) {
    const float rcp = safe_rcp(noise);
    out[0]          = quantize_avx2((gain * re) * rcp, 24.0f);
    out[1]          = quantize_avx2((gain * im) * rcp, 24.0f);
    return;
  }
  if (!(noise > 0.0f)) {
    out[0] = out[1] = 0;
    return;
  }
  out[0] = quantize_scalar(gain * re / noise, 24.0f);
  out[1] = quantize_scalar(gain * im / noise, 24.0f);
}

static void demod_qam16_symbol(int8_t* out, float re, float im, float noise, int path)
{
  const float u10 = 1.0f / sqrtf(10.0f);
  const float gain_first = 4.0f * u10, thr = 2.0f * u10;
  const float x[2] = {re, im};
  if (path == PATH_AVX2) {
    const float rcp = safe_rcp(noise);
    for (int c = 0; c != 2; ++c) {
      const float first  = gain_first * x[c];
      const float second = mul_add(2.0f, first, -copysignf(0.8f, x[c]), FUSED_QAM16_SECOND); /* exact either way */
      float       l01    = (fabsf(x[c]) > thr) ? second : first;
      float       l23    = 0.8f - fabsf(first);
      l01 *= rcp, l23 *= rcp;
      if (fabsf(x[c]) <= NEAR_ZERO) {
        l01 = 0.0f, l23 = 0.0f;
      }
      out[c]     = quantize_avx2(l01, 20.0f);
      out[2 + c] = quantize_avx2(l23, 20.0f);
    }
    return;
  }
  if (near_zero_symbol(re, im)) {
    memset(out, 0, 4);
    return;
  }
  for (int c = 0; c != 2; ++c) {
    if (!(noise > 0.0f)) {
      out[c] = out[2 + c] = 0;
      continue;
    }
    float l = gain_first * x[c];
    if (fabsf(x[c]) > thr) {
      l = mul_add(2.0f, l, -copysignf(0.8f, x[c]), FUSED_QAM16_SECOND);
    }
    out[c]      = quantize_scalar(l / noise, 20.0f);
    float l23   = mul_add(-gain_first, fabsf(x[c]), 0.8f, FUSED_QAM16_SCALAR_23); /* FUSED */
    out[2 + c]  = quantize_scalar(l23 / noise, 20.0f);
  }
}

/* One call of demodulate_soft. mod: 0 pi/2-BPSK, 1 BPSK, 2 QPSK, 4 16QAM, 6 64QAM, 8 256QAM. symbols: re, im pairs. */
void orc_demodulate_soft(int8_t* llrs, const float* symbols, const float* noise_vars, unsigned nof_symbols, int mod)
{
  init_tabs();
  /* Sizes of the vector batches of each loop, in symbols: {AVX512, AVX2}. */
  unsigned b512 = 0, b256 = 0;
  switch (mod) {
    case 2:
      b256 = 16; /* demodulation_mapper_qpsk.cpp:141-151 */
      break;
    case 4:
      b256 = 8; /* demodulation_mapper_qam16.cpp:235-243 */
      break;
    case 6:
      b512 = 32, b256 = 16; /* demodulation_mapper_qam64.cpp:403-425 */
      break;
    case 8:
      b512 = 8, b256 = 4; /* demodulation_mapper_qam256.cpp:365-385 */
      break;
    default:
      break;
  }
  const unsigned end512 = b512 ? (nof_symbols / b512) * b512 : 0;
  const unsigned end256 = b256 ? (nof_symbols / b256) * b256 : 0;
  for (unsigned i = 0; i != nof_symbols; ++i) {
    const int   path = (i < end512) ? PATH_AVX512 : ((i < end256) ? PATH_AVX2 : PATH_SCALAR);
    const float re = symbols[2 * i], im = symbols[2 * i + 1], nv = noise_vars[i];
    switch (mod) {
      case 0: /* demodulation_mapper_impl.cpp:54-77: odd symbols are rotated by -90 degrees first */
        llrs[i] = (i & 1) ? demod_bpsk(im, -re, nv) : demod_bpsk(re, im, nv);
        break;
      case 1:
        llrs[i] = demod_bpsk(re, im, nv);
        break;
      case 2:
        demod_qpsk_symbol(llrs + 2 * (size_t)i, re, im, nv, path);
        break;
      case 4:
        demod_qam16_symbol(llrs + 4 * (size_t)i, re, im, nv, path);
        break;
      case 6:
        demod_interval_symbol(llrs + 6 * (size_t)i, re, im, nv, path, tab64, 3);
        break;
      default:
        demod_interval_symbol(llrs + 8 * (size_t)i, re, im, nv, path, tab256, 4);
        break;
    }
  }
}

/* ---- scrambling sequence, TS 38.211 5.2.1 ---------------------------------------------------------------------------- */
/* c(n) = x1(n + 1600) ^ x2(n + 1600); x1(n + 31) = x1(n + 3) ^ x1(n), x1(0) = 1; x2(n + 31) = x2(n + 3) ^ x2(n + 2) ^
 * x2(n + 1) ^ x2(n), x2(0..30) = bits of c_init. Bit by bit, no fast-forward tables: the oracle is the slow, obvious one. */
void orc_prg_bits(uint8_t* out, unsigned n, uint32_t c_init, unsigned offset)
{
  uint32_t x1 = 1u, x2 = c_init & 0x7fffffffu; /* bit i = x(n + i) */
  const unsigned long long total = 1600ull + offset + n;
  for (unsigned long long k = 0; k != total; ++k) {
    if (k >= 1600ull + offset) {
      out[k - 1600ull - offset] = (uint8_t)((x1 ^ x2) & 1u);
    }
    const uint32_t n1 = ((x1 >> 3) ^ x1) & 1u;
    const uint32_t n2 = ((x2 >> 3) ^ (x2 >> 2) ^ (x2 >> 1) ^ x2) & 1u;
    x1 = (x1 >> 1) | (n1 << 30);
    x2 = (x2 >> 1) | (n2 << 30);
  }
}

/* revert_scrambling, pusch_demodulator_impl.cpp:60-222: out = -in where the sequence bit is 1 (two's complement negation:
 * -128 stays -128). */
void orc_descramble(int8_t* out, const int8_t* in, unsigned n, uint32_t c_init, unsigned offset)
{
  uint32_t x1 = 1u, x2 = c_init & 0x7fffffffu;
  const unsigned long long total = 1600ull + offset + n;
  for (unsigned long long k = 0; k != total; ++k) {
    if (k >= 1600ull + offset) {
      const unsigned i = (unsigned)(k - 1600ull - offset);
      out[i]           = ((x1 ^ x2) & 1u) ? (int8_t)(uint8_t)(0u - (uint8_t)in[i]) : in[i];
    }
    const uint32_t n1 = ((x1 >> 3) ^ x1) & 1u;
    const uint32_t n2 = ((x2 >> 3) ^ (x2 >> 2) ^ (x2 >> 1) ^ x2) & 1u;
    x1 = (x1 >> 1) | (n1 << 30);
    x2 = (x2 >> 1) | (n2 << 30);
  }
}
