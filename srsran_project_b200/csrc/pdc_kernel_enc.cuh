// PDSCH transport-block encoder on the device (the downlink mirror of the decode chain): TB CRC attachment,
// segmentation, code-block CRC, LDPC encoding, rate matching with bit interleaving and code-block concatenation.
//
// What it replaces in the reference:
//   pdsch_encoder_impl::encode          lib/phy/upper/channel_processors/pdsch/pdsch_encoder_impl.cpp:31-80
//   ldpc_segmenter_tx_impl              lib/phy/upper/channel_coding/ldpc/ldpc_segmenter_tx_impl.cpp (TB CRC, CB CRC, fillers,
//                                       zero padding; segment sizes ldpc.h:128-214, rate-matched lengths
//                                       ldpc_segmenter_helpers.h:82-100)
//   ldpc_encoder_impl / _generic / _avx2 lib/phy/upper/channel_coding/ldpc/ldpc_encoder_impl.cpp, ldpc_encoder_generic.cpp
//                                       (systematic bits, the four core parity blocks through the double-diagonal structure,
//                                       extension parity blocks), TS 38.212 5.3.2
//   ldpc_rate_matcher_impl              lib/phy/upper/channel_coding/ldpc/ldpc_rate_matcher_impl.cpp (select_bits from the
//                                       circular buffer skipping the filler window, interleave_bits), TS 38.212 5.4.2
//
// Two launches per batch of transport blocks:
//   pdc_enc_crc_kernel   one CTA per code block: the code block's share of the TB CRC. The CRC is linear: the remainder of the
//                        whole block is the xor over code blocks of (remainder of the code block's payload) x^(bits that
//                        follow it, CRC included) mod P; the first factor is a xor of per-bit powers, the second a
//                        square-and-multiply chain of 24-bit carry-less products. 152 CTAs do in parallel what the
//                        reference does in one serial pass over 160 KB.
//   pdc_enc_cb_kernel    one CTA per code block: payload bits (+ TB CRC and zero padding in the last block) -> CB CRC24B ->
//                        core and extension parity (only the rows the rate matcher will read) -> bit selection and interleaving
//                        straight into the codeword, one byte per bit as pdsch_encoder::encode delivers it.
// Bits live one per byte in shared memory: the circulant P^s of a Z-bit block is an index rotation, valid for every one of
// the 51 lifting sizes. All arithmetic is GF(2): bit-exact by construction where the index algebra is right.
#pragma once
#include "pdc_device.cuh"

// x^(2^k) mod P, k < 24, per CRC type (index crc_type - 1), and x^(2^k + j) mod P for j < 24: multiplication of a 24-bit
// remainder by x^(2^k) is then the xor of the rows its set bits select (one warp: a lane per bit).
__constant__ uint32_t c_enc_xpow2[3][24];
__device__ uint32_t   d_enc_xpow2_rows[3][24][32];

__device__ __forceinline__ uint32_t pdc_crc_poly(int crc_type)
{
  return (crc_type == PDC_CRC16) ? 0x11021u : ((crc_type == PDC_CRC24A) ? 0x1864CFBu : 0x1800063u);
}
// a(x) b(x) mod P(x) over GF(2).
__device__ __forceinline__ uint32_t pdc_crc_mulmod(uint32_t a, uint32_t b, int crc_type)
{
  const int      order = (crc_type == PDC_CRC16) ? 16 : 24;
  const uint32_t poly = pdc_crc_poly(crc_type), top = 1u << order;
  uint32_t       r = 0;
  for (int i = order - 1; i >= 0; --i) {
    r <<= 1;
    if (r & top) {
      r ^= poly;
    }
    if ((b >> i) & 1u) {
      r ^= a;
    }
  }
  return r;
}
// a(x) x^n mod P(x).
__device__ __forceinline__ uint32_t pdc_crc_shift(uint32_t a, uint32_t n, int crc_type)
{
  for (int k = 0; k < 24 && (n >> k) != 0; ++k) {
    if ((n >> k) & 1u) {
      a = pdc_crc_mulmod(a, c_enc_xpow2[crc_type - 1][k], crc_type);
    }
  }
  return a;
}

// a(x) x^n mod P(x), computed by one full warp (every lane passes the same a and n and gets the result).
__device__ __forceinline__ uint32_t pdc_crc_shift_warp(uint32_t a, uint32_t n, int crc_type, int lane)
{
  for (int k = 0; k < 24 && (n >> k) != 0; ++k) {
    if ((n >> k) & 1u) {
      const uint32_t term = ((a >> lane) & 1u) ? d_enc_xpow2_rows[crc_type - 1][k][lane] : 0u;
      a                   = __reduce_xor_sync(0xffffffffu, term);
    }
  }
  return a;
}

// Bit n of (payload ++ TB CRC ++ zero padding) of a transport block.
__device__ __forceinline__ uint32_t pdc_enc_tb_bit(const pdc_dev_enc_cb& cb, uint32_t n, uint32_t tb_crc, int tb_crc_bits)
{
  if (n < cb.tbs_bits) {
    return (cb.tb[n >> 3] >> (7 - (n & 7))) & 1u;
  }
  const uint32_t k = n - cb.tbs_bits;
  return (k < (uint32_t)tb_crc_bits) ? ((tb_crc >> (tb_crc_bits - 1 - k)) & 1u) : 0u;
}

// XOR of `v` over the CTA; every thread gets the result. scratch: 32 words.
__device__ __forceinline__ uint32_t pdc_enc_block_xor(uint32_t v, uint32_t* scratch, int tid, int nth)
{
  v = __reduce_xor_sync(0xffffffffu, v);
  __syncthreads();
  if ((tid & 31) == 0) {
    scratch[tid >> 5] = v;
  }
  __syncthreads();
  uint32_t r = 0;
  for (int w = 0; w < (nth >> 5); ++w) {
    r ^= scratch[w];
  }
  return r;
}

// 32 bits of the packed transport block starting at bit g (MSB first in the bytes), as a word whose bit j is stream bit g + j.
// The caller guarantees g + 32 <= number of bits in the block.
__device__ __forceinline__ uint32_t pdc_enc_tb_word(const uint8_t* __restrict__ tb, uint32_t g)
{
  const uint32_t B = g >> 3, sh = g & 7u;
  const uint32_t hi = ((uint32_t)tb[B] << 24) | ((uint32_t)tb[B + 1] << 16) | ((uint32_t)tb[B + 2] << 8) | (uint32_t)tb[B + 3];
  uint32_t       be = hi << sh; // big-endian window: stream bit g at bit 31
  if (sh != 0) {
    be |= (uint32_t)tb[B + 4] >> (8 - sh);
  }
  return __brev(be);
}

// b(x) x^(8 k) mod P for every byte b, k < 7, per CRC type: reduces a 32-bit word (k = 0..3) and multiplies a remainder by
// x^32 (k = 4..6) with one lookup per byte.
__device__ uint32_t d_enc_crc_bytes[3][7][256];

// CRC of a D-bit message, i.e. (sum of bit_i x^(D - 1 - i)) x^order mod P, by one warp. fetch(w) returns the word whose bit j
// is message bit 32 w + j (only bits below D are looked at). Each lane runs Horner's rule over a contiguous run of whole words
// (seven byte-table lookups per word), its remainder is moved to its place with one x^j table entry per set bit, and the warp
// xors the lanes; the bits behind the last whole word are folded bit by bit. About 200 warp instructions for an 8424-bit code
// block, where one table entry per set message bit costs about 4000.
template <class F>
__device__ __forceinline__ uint32_t pdc_enc_crc_warp(F fetch, uint32_t D, int crc_type, const uint32_t* __restrict__ crc_pow, int lane)
{
  const uint32_t  order = (crc_type == PDC_CRC16) ? 16u : 24u;
  const uint32_t* pw    = crc_pow + (size_t)(crc_type - 1) * PDC_CRC_TABLE_LEN;
  const uint32_t(*T)[256] = d_enc_crc_bytes[crc_type - 1];
  const uint32_t nfull = D >> 5, tail = D & 31u;
  const uint32_t q  = (nfull + 31u) >> 5; // whole words per lane
  const uint32_t w0 = min(nfull, (uint32_t)lane * q), w1 = min(nfull, w0 + q);
  uint32_t       acc = 0;
  for (uint32_t w = w0; w < w1; ++w) {
    const uint32_t W = __brev(fetch(w)); // bit i <-> x^i
    acc = __ldg(&T[4][acc & 255u]) ^ __ldg(&T[5][(acc >> 8) & 255u]) ^ __ldg(&T[6][acc >> 16]) ^ __ldg(&T[0][W & 255u]) ^
          __ldg(&T[1][(W >> 8) & 255u]) ^ __ldg(&T[2][(W >> 16) & 255u]) ^ __ldg(&T[3][W >> 24]);
  }
  uint32_t       out = 0;
  const uint32_t n   = D - 32u * w1 + order; // x^n: the bits that follow this lane's run, and the CRC's own x^order
#pragma unroll
  for (int b = 0; b < 24; ++b) {
    if ((acc >> b) & 1u) {
      out ^= __ldg(pw + n + b);
    }
  }
  if (lane == 0 && tail != 0) {
    const uint32_t v = fetch(nfull);
    for (uint32_t j = 0; j < tail; ++j) {
      if ((v >> j) & 1u) {
        out ^= pw[tail - 1u - j + order];
      }
    }
  }
  return __reduce_xor_sync(0xffffffffu, out);
}

// The code block's share of the TB CRC (ldpc_segmenter_tx_impl: the CRC the segmenter computes before cutting the block).
__global__ void __launch_bounds__(128) pdc_enc_crc_kernel(const pdc_dev_enc_tb* __restrict__ tbs, const uint32_t* __restrict__ crc_pow)
{
  // A warp per code block, four code blocks per CTA.
  const int      lane = threadIdx.x & 31;
  const uint32_t r    = blockIdx.x * 4u + (threadIdx.x >> 5);
  if (r >= tbs[blockIdx.y].nof_cbs) {
    return;
  }
  const pdc_dev_enc_cb cb = pdc_enc_cb_of(tbs[blockIdx.y], r);
  if (cb.first_bit >= cb.tbs_bits) {
    return; // nothing of the payload in this code block
  }
  const uint32_t nbits = min(cb.nof_bits, cb.tbs_bits - cb.first_bit); // payload bits carried here
  auto           fetch = [&](uint32_t w) {
    const uint32_t n0 = 32u * w;
    if (n0 + 32u <= nbits) {
      return pdc_enc_tb_word(cb.tb, cb.first_bit + n0);
    }
    uint32_t v = 0;
    for (uint32_t j = 0; n0 + j < nbits; ++j) {
      const uint32_t g = cb.first_bit + n0 + j;
      v |= (uint32_t)((cb.tb[g >> 3] >> (7 - (g & 7))) & 1u) << j;
    }
    return v;
  };
  // Remainder of this code block's payload as if it ended the transport block (the CRC's x^order included) ...
  const uint32_t rem = pdc_enc_crc_warp(fetch, nbits, cb.tb_crc_type, crc_pow, lane);
  // ... moved to its place: the payload bits that follow it.
  const uint32_t after = cb.tbs_bits - (cb.first_bit + nbits);
  const uint32_t share = pdc_crc_shift_warp(rem, after, cb.tb_crc_type, lane);
  if (lane == 0) {
    atomicXor(cb.tb_crc, share);
  }
}


// ---- word path: lifting sizes that are multiples of 32 (12 of the 51, among them every Z >= 256) -------------------------------
// A Z-bit block is nw = Z / 32 words, bit i of the block = bit (i & 31) of word i >> 5. The circulant P^s is a funnel shift
// across neighbouring words of the block.
__device__ __forceinline__ uint32_t pdc_enc_rot_word(const uint32_t* blk, int w, int s, int nw)
{
  int a = w + (s >> 5);
  a -= (a >= nw) ? nw : 0;
  int b = a + 1;
  b -= (b >= nw) ? nw : 0;
  return __funnelshift_r(blk[a], blk[b], (unsigned)(s & 31));
}

// Four consecutive bits t .. t + 3 of the compacted circular buffer (t + 3 < Lc, not straddling the filler window), bit k of the
// result = position t + k.
__device__ __forceinline__ uint32_t pdc_enc_nibble(const uint32_t* cbuf, int t, int n_info, int fw)
{
  const int idx = (t < n_info) ? t : t + fw;
  const int w   = idx >> 5;
  return __funnelshift_r(cbuf[w], cbuf[w + 1], (unsigned)(idx & 31)) & 0xfu; // one word of slack behind the buffer
}

// Interleaver for Qm = 8 and 4: a thread takes four consecutive columns i of the Qm x (E / Qm) interleaver, i.e. 4 Qm
// consecutive output bytes. Row q contributes four consecutive buffer bits (one funnel shift); with the nibbles of four rows
// side by side in one register, (X >> k) & 0x01010101 is a finished output word. Returns false if the geometry does not fit
// (the caller then runs the general loop).
template <int QM>
__device__ __forceinline__ bool pdc_enc_rate_match_fast(const pdc_dev_enc_cb& cb, const uint32_t* cbuf, const int* base,
                                                        int Lc, int n_info, int fw, int tid, int nth)
{
  const int Kq = (int)cb.E / QM;
  if ((Kq & 3) != 0 || (reinterpret_cast<uintptr_t>(cb.out) & 15) != 0) {
    return false;
  }
  int bq[QM];
#pragma unroll
  for (int q = 0; q < QM; ++q) {
    bq[q] = base[q];
  }
  for (int i0 = 4 * tid; i0 < Kq; i0 += 4 * nth) {
    uint32_t nib[QM];
#pragma unroll
    for (int q = 0; q < QM; ++q) {
      int t = bq[q] + i0;
      while (t >= Lc) {
        t -= Lc;
      }
      if (t + 3 < Lc && (t + 3 < n_info || t >= n_info)) {
        nib[q] = pdc_enc_nibble(cbuf, t, n_info, fw);
      } else { // wraps around the buffer or jumps over the filler window inside these four bits
        nib[q] = 0;
        for (int k = 0; k < 4; ++k) {
          int tk = t + k;
          tk -= (tk >= Lc) ? Lc : 0;
          const int idx = (tk < n_info) ? tk : tk + fw;
          nib[q] |= ((cbuf[idx >> 5] >> (idx & 31)) & 1u) << k;
        }
      }
    }
    uint32_t out[QM];
#pragma unroll
    for (int h = 0; h < QM / 4; ++h) {
      const uint32_t X = nib[4 * h] | (nib[4 * h + 1] << 8) | (nib[4 * h + 2] << 16) | (nib[4 * h + 3] << 24);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        out[k * (QM / 4) + h] = (X >> k) & 0x01010101u; // bytes (i0 + k) QM + 4 h .. + 3
      }
    }
    uint4* dst = reinterpret_cast<uint4*>(cb.out + (size_t)i0 * QM);
#pragma unroll
    for (int v = 0; v < QM / 4; ++v) {
      dst[v] = make_uint4(out[4 * v], out[4 * v + 1], out[4 * v + 2], out[4 * v + 3]);
    }
  }
  return true;
}

template <int QM>
__device__ __forceinline__ void pdc_enc_rate_match_words(const pdc_dev_enc_cb& cb, const uint32_t* cbuf, const int* base,
                                                         int Lc, int n_info, int fw, int tid, int nth)
{
  const int E = (int)cb.E;
  const bool aligned = (reinterpret_cast<uintptr_t>(cb.out) & 3) == 0;
  for (int o4 = tid * 4; o4 < E; o4 += nth * 4) {
    uint32_t word = 0;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int o = o4 + b;
      if (o < E) {
        const int i = o / QM, q = o - i * QM;
        int       t = base[q] + i;
        while (t >= Lc) {
          t -= Lc;
        }
        const int idx = (t < n_info) ? t : t + fw;
        word |= ((cbuf[idx >> 5] >> (idx & 31)) & 1u) << (8 * b);
      }
    }
    if (o4 + 4 <= E && aligned) {
      *reinterpret_cast<uint32_t*>(cb.out + o4) = word;
    } else {
      for (int b = 0; b < 4 && o4 + b < E; ++b) {
        cb.out[o4 + b] = (uint8_t)(word >> (8 * b));
      }
    }
  }
}

__device__ __forceinline__ void pdc_enc_cb_words(const pdc_dev_enc_cb& cb, const uint32_t* __restrict__ gr,
                                                 const uint32_t* __restrict__ crc_pow, uint32_t* cww, uint32_t* scratch,
                                                 int* s_base, int tid, int nth)
{
  const int bgi = cb.bg - 1, Z = cb.Z, nw = Z >> 5;
  const int Kb = (cb.bg == 1) ? 22 : 10, Nfull = (cb.bg == 1) ? 68 : 52, M = Nfull - Kb;
  const int rows = min((int)cb.rows_needed, M);
  uint32_t* lam  = cww + (Kb + rows) * nw;
  // ---- message words -----------------------------------------------------------------------------------------------------------
  const int      tb_crc_bits = (cb.tb_crc_type == PDC_CRC16) ? 16 : 24;
  const uint32_t tb_crc      = *cb.tb_crc;
  const uint32_t D           = cb.nof_bits;
  for (int w = tid; w < (Kb + rows) * nw; w += nth) {
    const uint32_t n0 = 32u * (uint32_t)w;
    uint32_t       v  = 0;
    if (n0 + 32u <= D && cb.first_bit + n0 + 32u <= cb.tbs_bits) {
      v = pdc_enc_tb_word(cb.tb, cb.first_bit + n0);
    } else if (n0 < D) {
      for (uint32_t j = 0; j < 32u && n0 + j < D; ++j) {
        v |= pdc_enc_tb_bit(cb, cb.first_bit + n0 + j, tb_crc, tb_crc_bits) << j;
      }
    }
    cww[w] = v;
  }
  __syncthreads();
  if (cb.cb_crc) {
    if (tid < 32) {
      const uint32_t acc = pdc_enc_crc_warp([&](uint32_t w) { return cww[w]; }, D, PDC_CRC24B, crc_pow, tid);
      if (tid < 24) {
        const uint32_t n = D + (uint32_t)tid;
        atomicOr(&cww[n >> 5], ((acc >> (23 - tid)) & 1u) << (n & 31u));
      }
    }
    __syncthreads();
  }
  // ---- core parity: lambda_m word by word, the edges of a row split over four threads (4 rows x nw words x 4 slices) ---------------
  for (int q = tid; q < 5 * nw; q += nth) {
    lam[q] = 0; // lambda and the sum block behind it
  }
  __syncthreads();
  for (int q = tid; q < 16 * nw; q += nth) {
    const int slice = q & 3, mw = q >> 2, m = mw / nw, w = mw - m * nw;
    const int e0 = c_bg_rows.start[bgi][m], e1 = e0 + c_bg_rows.deg[bgi][m];
    uint32_t  b  = 0;
    for (int e = e0 + slice; e < e1; e += 4) {
      const uint32_t ent = gr[e];
      const int      col = (int)(ent & 0xffu);
      if (col < Kb) {
        b ^= pdc_enc_rot_word(cww + col * nw, w, (int)(ent >> 8), nw);
      }
    }
    atomicXor(&lam[mw], b);
  }
  const int c[4] = {cb.core_shift[0], cb.core_shift[1], cb.core_shift[2], cb.core_shift[3]}, bshift = cb.bshift;
  __syncthreads();
  uint32_t* p1  = cww + Kb * nw;
  uint32_t* sum = lam + 4 * nw; // nw words more of scratch
  for (int w = tid; w < nw; w += nth) {
    sum[w] = lam[w] ^ lam[nw + w] ^ lam[2 * nw + w] ^ lam[3 * nw + w];
  }
  __syncthreads();
  for (int w = tid; w < nw; w += nth) {
    p1[w] = pdc_enc_rot_word(sum, w, (Z - bshift) % Z, nw); // sum = P^b p1
  }
  __syncthreads();
  for (int m = 0; m < 3; ++m) {
    uint32_t* pn = cww + (Kb + 1 + m) * nw;
    for (int w = tid; w < nw; w += nth) {
      uint32_t b = lam[m * nw + w];
      if (c[m] >= 0) {
        b ^= pdc_enc_rot_word(p1, w, c[m], nw);
      }
      if (m > 0) {
        b ^= cww[(Kb + m) * nw + w];
      }
      pn[w] = b;
    }
    __syncthreads();
  }
  // ---- extension rows -------------------------------------------------------------------------------------------------------------
  for (int q = tid; q < (rows - 4) * nw; q += nth) {
    const int m = 4 + q / nw, w = q - (m - 4) * nw;
    const int e0 = c_bg_rows.start[bgi][m], e1 = e0 + c_bg_rows.deg[bgi][m];
    uint32_t  b  = 0;
    for (int e = e0; e < e1; ++e) {
      const uint32_t ent = gr[e];
      const int      col = (int)(ent & 0xffu);
      if (col != Kb + m) {
        b ^= pdc_enc_rot_word(cww + col * nw, w, (int)(ent >> 8), nw);
      }
    }
    cww[(Kb + m) * nw + w] = b;
  }
  // ---- rate matching ----------------------------------------------------------------------------------------------------------------
  const int N = (Nfull - 2) * Z, Ncb = (cb.Nref > 0) ? min((int)cb.Nref, N) : N;
  const int n_sys = (Kb - 2) * Z, n_info = n_sys - (int)cb.nof_filler;
  const int fw = max(0, min(n_sys, Ncb) - min(n_info, Ncb)), Lc = Ncb - fw;
  const int f1[4] = {0, 17, 33, 56}, f2[4] = {0, 13, 25, 43};
  const int fr = (cb.bg == 1) ? f1[cb.rv & 3] : f2[cb.rv & 3];
  const int k0 = (int)(((long long)fr * Ncb) / N) * Z;
  const int t0 = (k0 < n_info) ? k0 : ((k0 >= n_sys) ? k0 - fw : n_info);
  const int Kq = (int)cb.E / cb.Qm;
  if (tid < cb.Qm) {
    s_base[tid] = (int)(((long long)t0 + (long long)tid * Kq) % Lc); // start of interleaver row q in the compacted buffer
  }
  __syncthreads();
  const uint32_t* cbuf = cww + 2 * nw;
  switch (cb.Qm) {
    case 1:
      pdc_enc_rate_match_words<1>(cb, cbuf, s_base, Lc, n_info, fw, tid, nth);
      break;
    case 2:
      pdc_enc_rate_match_words<2>(cb, cbuf, s_base, Lc, n_info, fw, tid, nth);
      break;
    case 4:
      if (!pdc_enc_rate_match_fast<4>(cb, cbuf, s_base, Lc, n_info, fw, tid, nth)) {
        pdc_enc_rate_match_words<4>(cb, cbuf, s_base, Lc, n_info, fw, tid, nth);
      }
      break;
    case 6:
      pdc_enc_rate_match_words<6>(cb, cbuf, s_base, Lc, n_info, fw, tid, nth);
      break;
    default:
      if (!pdc_enc_rate_match_fast<8>(cb, cbuf, s_base, Lc, n_info, fw, tid, nth)) {
        pdc_enc_rate_match_words<8>(cb, cbuf, s_base, Lc, n_info, fw, tid, nth);
      }
      break;
  }
}

// cw: (Kb + rows) Z bytes, one bit each: systematic blocks, then parity blocks (lifting sizes that are not multiples of 32;
// the others take pdc_enc_cb_words above).
__global__ void __launch_bounds__(256) pdc_enc_cb_kernel(const pdc_dev_enc_tb* __restrict__ tbs, const uint32_t* __restrict__ graphs,
                                                         const uint32_t* __restrict__ crc_pow)
{
  extern __shared__ __align__(16) unsigned char cw[];
  __shared__ uint32_t  scratch[32];
  if (blockIdx.x >= tbs[blockIdx.y].nof_cbs) {
    return;
  }
  const pdc_dev_enc_cb cb  = pdc_enc_cb_of(tbs[blockIdx.y], blockIdx.x);
  const int            tid = threadIdx.x, nth = blockDim.x;
  const int            bgi = cb.bg - 1, Z = cb.Z;
  const int            Kb = (cb.bg == 1) ? 22 : 10, Nfull = (cb.bg == 1) ? 68 : 52, M = Nfull - Kb;
  const int            rows = min((int)cb.rows_needed, M);
  const uint32_t*      gr   = graphs + (size_t)cb.graph_idx * PDC_GRAPH_STRIDE;
  const int            K    = Kb * Z;
  if ((Z & 31) == 0) {
    __shared__ int s_base[8];
    pdc_enc_cb_words(cb, gr, crc_pow, reinterpret_cast<uint32_t*>(cw), scratch, s_base, tid, nth);
    return;
  }

  // ---- message: payload (++ TB CRC ++ padding) ++ CB CRC ++ fillers (zero while encoding) -------------------------------------
  const int      tb_crc_bits = (cb.tb_crc_type == PDC_CRC16) ? 16 : 24;
  const uint32_t tb_crc      = *cb.tb_crc;
  for (int n = tid; n < (Kb + rows) * Z; n += nth) {
    cw[n] = (n < (int)cb.nof_bits) ? (unsigned char)pdc_enc_tb_bit(cb, cb.first_bit + (uint32_t)n, tb_crc, tb_crc_bits) : 0;
  }
  __syncthreads();
  if (cb.cb_crc) {
    const uint32_t* pw  = crc_pow + (size_t)(PDC_CRC24B - 1) * PDC_CRC_TABLE_LEN;
    const int       D   = (int)cb.nof_bits;
    uint32_t        acc = 0;
    for (int n = tid; n < D; n += nth) {
      if (cw[n]) {
        acc ^= pw[D - 1 - n + 24];
      }
    }
    acc = pdc_enc_block_xor(acc, scratch, tid, nth);
    if (tid < 24) {
      cw[D + tid] = (unsigned char)((acc >> (23 - tid)) & 1u);
    }
    __syncthreads();
  }

  // ---- core parity (ldpc_encoder_impl.cpp high-rate region; TS 38.212 5.3.2): lambda_m = sum over the systematic columns
  //      of P^shift s_col for the four core rows, kept in the parity blocks 1..3 and in a spare block ----------------------------
  unsigned char* p1  = cw + K;                 // parity block of column Kb
  unsigned char* lam = cw + (Kb + rows) * Z;   // 4 Z bytes of scratch behind the code word
  for (int q = tid; q < 4 * Z; q += nth) {
    const int m = q / Z, i = q - m * Z;
    const int e0 = c_bg_rows.start[bgi][m], e1 = e0 + c_bg_rows.deg[bgi][m];
    unsigned  b  = 0;
    for (int e = e0; e < e1; ++e) {
      const uint32_t ent = gr[e];
      const int      col = (int)(ent & 0xffu);
      if (col < Kb) {
        int j = i + (int)(ent >> 8);
        j -= (j >= Z) ? Z : 0;
        b ^= cw[col * Z + j];
      }
    }
    lam[q] = (unsigned char)b;
  }
  __syncthreads();
  // Column Kb holds three circulants in the core rows, two of them equal: the sum of the four rows leaves P^b p1.
  const int c[4] = {cb.core_shift[0], cb.core_shift[1], cb.core_shift[2], cb.core_shift[3]}, bshift = cb.bshift;
  for (int i = tid; i < Z; i += nth) {
    int j = i + bshift;
    j -= (j >= Z) ? Z : 0;
    p1[j] = lam[i] ^ lam[Z + i] ^ lam[2 * Z + i] ^ lam[3 * Z + i];
  }
  __syncthreads();
  // Rows 0..2 give the next three parity blocks in turn: lambda_m + P^c_m p1 + p_m (m > 0) + p_(m+1) = 0.
  for (int m = 0; m < 3; ++m) {
    unsigned char* pn = cw + (Kb + 1 + m) * Z;
    for (int i = tid; i < Z; i += nth) {
      unsigned b = lam[m * Z + i];
      if (c[m] >= 0) {
        int j = i + c[m];
        j -= (j >= Z) ? Z : 0;
        b ^= p1[j];
      }
      if (m > 0) {
        b ^= cw[(Kb + m) * Z + i];
      }
      pn[i] = (unsigned char)b;
    }
    __syncthreads();
  }
  // ---- extension rows: the parity block of row m >= 4 is the sum of everything else on the row (all of it known by now) --------
  for (int q = tid; q < (rows - 4) * Z; q += nth) {
    const int m = 4 + q / Z, i = q - (m - 4) * Z;
    const int e0 = c_bg_rows.start[bgi][m], e1 = e0 + c_bg_rows.deg[bgi][m];
    unsigned  b  = 0;
    for (int e = e0; e < e1; ++e) {
      const uint32_t ent = gr[e];
      const int      col = (int)(ent & 0xffu);
      if (col != Kb + m) {
        int j = i + (int)(ent >> 8);
        j -= (j >= Z) ? Z : 0;
        b ^= cw[col * Z + j];
      }
    }
    cw[(Kb + m) * Z + i] = (unsigned char)b;
  }
  __syncthreads();

  // ---- rate matching (ldpc_rate_matcher_impl.cpp: select_bits + interleave_bits) ------------------------------------------------
  // Circular buffer = code word without its first 2 Z bits, cut at Ncb; the filler window [n_info, n_sys) is never sent.
  // t-th transmittable position: t < n_info ? t : t + fw. Output bit i Qm + q = selected bit q (E / Qm) + i.
  const int    N = (Nfull - 2) * Z, Ncb = (cb.Nref > 0) ? min((int)cb.Nref, N) : N;
  const int    n_sys = (Kb - 2) * Z, n_info = n_sys - (int)cb.nof_filler;
  const int    fw = max(0, min(n_sys, Ncb) - min(n_info, Ncb)), Lc = Ncb - fw;
  const int    f1[4] = {0, 17, 33, 56}, f2[4] = {0, 13, 25, 43};
  const int    fr = (cb.bg == 1) ? f1[cb.rv & 3] : f2[cb.rv & 3];
  const int    k0 = (int)(((long long)fr * Ncb) / N) * Z; // floor(f Ncb / N) Z: exact in integers
  const int    t0 = (k0 < n_info) ? k0 : ((k0 >= n_sys) ? k0 - fw : n_info); // a start inside the window moves to its end
  const int    E = (int)cb.E, Qm = cb.Qm, Kq = E / Qm;
  const unsigned char* cbuf = cw + 2 * Z;
  for (int o4 = tid * 4; o4 < E; o4 += nth * 4) {
    uint32_t word = 0;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int o = o4 + b;
      if (o < E) {
        const int i = o / Qm, q = o - i * Qm;
        int       t = (t0 + q * Kq + i) % Lc;
        const int idx = (t < n_info) ? t : t + fw;
        word |= (uint32_t)cbuf[idx] << (8 * b);
      }
    }
    if (o4 + 4 <= E && ((reinterpret_cast<uintptr_t>(cb.out) & 3) == 0)) {
      *reinterpret_cast<uint32_t*>(cb.out + o4) = word;
    } else {
      for (int b = 0; b < 4 && o4 + b < E; ++b) {
        cb.out[o4 + b] = (uint8_t)(word >> (8 * b));
      }
    }
  }
}

extern "C" cudaError_t pdc_enc_init_device()
{
  static const uint32_t poly[3]  = {0x11021u, 0x1864CFBu, 0x1800063u};
  static const int      order[3] = {16, 24, 24};
  uint32_t              tab[3][24];
  for (int c = 0; c != 3; ++c) {
    const uint32_t top = 1u << order[c];
    // x^(2^k): square repeatedly (a(x)^2 mod P by shift-and-add)
    uint32_t v = 2u; // x
    for (int k = 0; k != 24; ++k) {
      tab[c][k]  = v;
      uint32_t r = 0;
      for (int i = order[c] - 1; i >= 0; --i) {
        r <<= 1;
        if (r & top) {
          r ^= poly[c];
        }
        if ((v >> i) & 1u) {
          r ^= v;
        }
      }
      v = r;
    }
  }
  cudaError_t e = cudaMemcpyToSymbol(c_enc_xpow2, tab, sizeof(tab));
  if (e != cudaSuccess) {
    return e;
  }
  static uint32_t bytes[3][7][256];
  for (int c = 0; c != 3; ++c) {
    const uint32_t top = 1u << order[c];
    for (int b = 0; b != 256; ++b) {
      // b(x) mod P is b itself (8 bits < order), then eight multiplications by x per table.
      uint32_t v = (uint32_t)b;
      for (int k = 0; k != 7; ++k) {
        bytes[c][k][b] = v;
        for (int i = 0; i != 8; ++i) {
          v <<= 1;
          if (v & top) {
            v ^= poly[c];
          }
        }
      }
    }
  }
  e = cudaMemcpyToSymbol(d_enc_crc_bytes, bytes, sizeof(bytes));
  if (e != cudaSuccess) {
    return e;
  }
  static uint32_t rows[3][24][32];
  for (int c = 0; c != 3; ++c) {
    const uint32_t top = 1u << order[c];
    for (int k = 0; k != 24; ++k) {
      uint32_t v = tab[c][k]; // x^(2^k + j) = x^(2^k) x^j
      for (int j = 0; j != 32; ++j) {
        rows[c][k][j] = (j < order[c]) ? v : 0u;
        v <<= 1;
        if (v & top) {
          v ^= poly[c];
        }
      }
    }
  }
  e = cudaMemcpyToSymbol(d_enc_xpow2_rows, rows, sizeof(rows));
  if (e != cudaSuccess) {
    return e;
  }
  return cudaFuncSetAttribute(pdc_enc_cb_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
}

// Threads per code block of pdc_enc_cb_kernel. A code block is a chain of ten short dependent phases, most of them narrower than
// a CTA: the device overlaps the code blocks that are resident, so the CTA is kept small (registers allow twelve per SM).
#define PDC_ENC_CB_THREADS 128
extern "C" cudaError_t pdc_enc_launch(const pdc_dev_enc_tb* d_tbs, uint32_t ntb, uint32_t max_cbs, const uint32_t* graphs,
                                      const uint32_t* crc_pow, size_t smem, cudaStream_t stream)
{
  // grid.y is limited to 65535 transport blocks per launch
  for (uint32_t t0 = 0; t0 < ntb; t0 += 65535u) {
    const uint32_t nt = (ntb - t0 < 65535u) ? ntb - t0 : 65535u;
    pdc_enc_crc_kernel<<<dim3((max_cbs + 3) / 4, nt), 128, 0, stream>>>(d_tbs + t0, crc_pow);
    pdc_enc_cb_kernel<<<dim3(max_cbs, nt), PDC_ENC_CB_THREADS, smem, stream>>>(d_tbs + t0, graphs, crc_pow);
  }
  return cudaGetLastError();
}
