// Device-side building blocks shared by the code-block kernels: LLR arithmetic, the position-parallel rate dematcher
// and the bit-parallel CRC.
//
// Arithmetic contract (reference file:line):
//   HARQ combine             ldpc_rate_dematcher_avx512_impl.cpp:29-62 (vector body) / log_likelihood_ratio.cpp:40-73 (tail)
//   variable-to-check        ldpc_decoder_avx512.cpp:86-126
//   min / second min / sign  ldpc_decoder_avx512.cpp:128-170, scale avx512_support.h:65-107, generic ldpc_decoder_generic.cpp:70-79
//   soft-bit update          ldpc_decoder_avx512.cpp:229-270 (promotion to +-127)
//   CRC                      crc_calculator_generic_impl.cpp:31-56,111-133
#pragma once
#include "pdc_common.h"
#include <cuda_runtime.h>

#define PDC_CRC_TABLE_LEN (PDC_MAX_K + 32) /* x^j mod P for j < K + order */
#define PDC_CRC_BIT_PAD 800                 /* zeros in front of the bit table: filler positions index below the order */
#define PDC_CRC_BIT_TABLE_LEN (PDC_CRC_BIT_PAD + PDC_MAX_K + 32)
/* Two copies per CRC type, the second one shifted by one entry, so that the pair of entries a thread needs for two
   neighbouring bits is an aligned 64-bit load whatever the parity of the number of filler bits. */
#define PDC_CRC_BIT_STRIDE (PDC_CRC_BIT_TABLE_LEN + 2)

struct pdc_launch_params {
  const pdc_dev_cb* cbs;
  const uint32_t*   graphs;  // [2 * 51][PDC_GRAPH_STRIDE]: col | (shift mod Z) << 8, rows concatenated
  const uint32_t*   crc_pow; // [3][PDC_CRC_TABLE_LEN]: x^j mod P for CRC16, CRC24A, CRC24B
  const uint32_t*   crc_bit; // [3][2][PDC_CRC_BIT_STRIDE]: copy v, entry v + PAD + j = x^j mod P for j >= order, 0 below
  const uint32_t*   order;   // optional permutation of CB indices (nullptr = identity)
  int               n;
};

// One code block of the PDSCH encoder (pdc_kernel_enc.cuh), planned on the host (pdc_runtime.cu: enc_plan).
struct pdc_dev_enc_cb {
  const uint8_t* tb;        // packed transport block (MSB first)
  uint8_t*       out;       // this code block's E bytes of the codeword (one bit per byte)
  uint32_t*      tb_crc;    // the transport block's CRC accumulator (xor of the code blocks' shares)
  uint32_t       tbs_bits;  // transport block size
  uint32_t       first_bit; // first TB bit (of payload ++ TB CRC ++ padding) this code block carries
  uint32_t       nof_bits;  // how many it carries: K' - CB CRC length
  uint32_t       E, Nref, nof_filler;
  uint16_t       Z, graph_idx;
  uint8_t        bg, Qm, rv, tb_crc_type; // tb_crc_type: PDC_CRC16 | PDC_CRC24A
  uint8_t        cb_crc;                  // 1: append CRC24B
  uint8_t        rows_needed;             // parity rows the rate matcher reads (>= 4)
  uint16_t       bshift;                  // shift b of the core parity solve: the four core rows sum to P^b p1
  int16_t        core_shift[4];           // shift of column Kb in core row m, -1 if the row does not hold it
};

// One transport block of the PDSCH encoder as the host plans it; the kernels derive their code block (blockIdx.x) from it, so
// a submission uploads one small structure per transport block, not one per code block.
struct pdc_dev_enc_tb {
  const uint8_t* tb;
  uint8_t*       out;      // the transport block's codeword (one bit per byte)
  uint32_t*      tb_crc;
  uint32_t       tbs_bits, nof_bits; // nof_bits: TB bits per code block (K' - CB CRC length)
  uint32_t       nof_cbs, nof_short; // code blocks; how many of them (the first ones) have the short rate-matched length
  uint32_t       E_short, E_long, Nref, nof_filler;
  uint16_t       Z, graph_idx;
  uint8_t        bg, Qm, rv, tb_crc_type, cb_crc, rows_short, rows_long, pad;
  uint16_t       bshift;
  int16_t        core_shift[4];
};
__device__ __forceinline__ pdc_dev_enc_cb pdc_enc_cb_of(const pdc_dev_enc_tb& t, uint32_t r)
{
  pdc_dev_enc_cb c;
  const bool     is_short = r < t.nof_short;
  c.tb = t.tb, c.tb_crc = t.tb_crc;
  c.out = t.out + (is_short ? (size_t)r * t.E_short : (size_t)t.nof_short * t.E_short + (size_t)(r - t.nof_short) * t.E_long);
  c.tbs_bits = t.tbs_bits, c.nof_bits = t.nof_bits, c.first_bit = r * t.nof_bits;
  c.E = is_short ? t.E_short : t.E_long, c.Nref = t.Nref, c.nof_filler = t.nof_filler;
  c.Z = t.Z, c.graph_idx = t.graph_idx, c.bg = t.bg, c.Qm = t.Qm, c.rv = t.rv, c.tb_crc_type = t.tb_crc_type;
  c.cb_crc = t.cb_crc, c.rows_needed = is_short ? t.rows_short : t.rows_long, c.bshift = t.bshift;
  c.core_shift[0] = t.core_shift[0], c.core_shift[1] = t.core_shift[1], c.core_shift[2] = t.core_shift[2];
  c.core_shift[3] = t.core_shift[3];
  return c;
}

// One OFDM symbol of one codeword for the UL-SCH demultiplexer, as the host plans it (ulsch_demultiplex_impl.cpp:330-463 in
// closed form: every RE set of the reference is "every d-th candidate, the first m of them"); the kernel rebuilds the sets.
struct pdc_dev_demux_sym {
  const int8_t* in;                           // first soft bit of the symbol (descrambled)
  int8_t *      sch, *harq, *csi1, *csi2;     // where this symbol's share of each stream starts
  uint32_t      nof_re, nbpr;                 // REs of the symbol, soft bits per RE (Qm * layers)
  uint32_t      x1_state, x2_state;           // the two scrambling LFSRs at the symbol's first soft bit (placeholder symbols only)
  uint32_t      rvd_d, rvd_m;                 // reserved REs: of all REs
  uint32_t      harq_d, harq_m;               // HARQ-ACK: of all REs (more than 2 bits) or of the reserved ones (harq_of_rvd)
  uint32_t      csi1_d, csi1_m;               // CSI Part 1: of the REs that are neither HARQ-ACK nor reserved
  uint32_t      csi2_d, csi2_m;               // CSI Part 2: of the REs that are neither HARQ-ACK (> 2 bits) nor CSI Part 1
  uint8_t       Qm, harq_of_rvd;
  uint8_t       harq_ph, csi1_ph, csi2_ph;    // 1 | 2: placeholder handling for that many bits, 0: plain copy
  uint8_t       pad[3];
};

// Row structure of the two base graphs (degrees and prefix sums), uploaded once per process.
struct pdc_bg_rows {
  uint16_t start[2][48];
  uint8_t  deg[2][48];
};
#if defined(PDC_KERNEL_TU)
__constant__ pdc_bg_rows c_bg_rows; // defined once, in pdc_kernels.cu
#endif

#define PDC_MAX_DYN_SMEM (200 * 1024) /* opt-in dynamic shared memory per CTA */

__device__ __forceinline__ int pdc_clamp(int x, int lo, int hi)
{
  return max(lo, min(x, hi));
}

// log_likelihood_ratio::operator+ (saturating sum with the +-infinity special cases).
__device__ __forceinline__ int pdc_llr_add_generic(int a, int b)
{
  if (a == -b) {
    return 0;
  }
  if (a == PDC_LLR_INF || a == -PDC_LLR_INF) {
    return a;
  }
  if (b == PDC_LLR_INF || b == -PDC_LLR_INF) {
    return b;
  }
  return pdc_clamp(a + b, -PDC_LLR_MAX, PDC_LLR_MAX);
}

// _mm512_adds_epi8 then clamp to +-120.
__device__ __forceinline__ int pdc_llr_add_simd(int a, int b)
{
  return pdc_clamp(pdc_clamp(a + b, -128, 127), -PDC_LLR_MAX, PDC_LLR_MAX);
}

__device__ __forceinline__ int pdc_v2c(int soft, int c2v)
{
  int v = pdc_clamp(soft - c2v, -PDC_LLR_MAX, PDC_LLR_MAX);
  return (soft >= PDC_LLR_INF || soft <= -PDC_LLR_INF) ? soft : v;
}

__device__ __forceinline__ int pdc_soft_update(int v2c, int c2v)
{
  int t = v2c + c2v;
  t     = (t > PDC_LLR_MAX) ? PDC_LLR_INF : ((t < -PDC_LLR_MAX) ? -PDC_LLR_INF : t);
  return (v2c >= PDC_LLR_INF || v2c <= -PDC_LLR_INF) ? v2c : t;
}

// Normalisation of a (non-negative, <= 120) minimum. AVX2/AVX512: (x * 52428) >> 16; generic: round(0.8f * x), which
// for integers 0..120 equals (8x + 5) / 10 (0.8x is never half-way between two integers).
__device__ __forceinline__ int pdc_scale(int x, bool generic)
{
  return generic ? ((8 * x + 5) / 10) : ((x * 52428) >> 16);
}

// Rate dematching + HARQ combining of one code block, one buffer position per thread and step. Every position's final
// value depends only on its old value and on the LLRs that land on it, in round order, so positions are independent.
__device__ __forceinline__ void pdc_dematch(const pdc_dematch_plan& pl, const int8_t* __restrict__ in,
                                            int8_t* __restrict__ buf, bool generic, int tid, int nth)
{
  const int F      = pl.n_sys - pl.n_info;
  const int Kq     = pl.E / pl.Qm;
  const int zero_to = pl.new_data ? ((pl.k0 < pl.n_info) ? pl.k0 : pl.n_info) : 0;
  const int fz_from = pl.final_zero ? (pl.N - (pl.Ncb - pl.end_idx)) : pl.N;
  const int simd_w  = generic ? 0 : 64;
  for (int p = tid; p < pl.N; p += nth) {
    int  v       = buf[p];
    bool in_fill = (p >= pl.n_info) && (p < pl.n_sys);
    if (p < zero_to) {
      v = 0;
    }
    if (pl.new_data && in_fill) {
      v = PDC_LLR_INF;
    }
    if (!in_fill && p < pl.Ncb) {
      int gp = (p < pl.n_info) ? p : (p - F);
      for (int r = 0; r < pl.rounds; ++r) {
        int start = (r == 0) ? pl.p0 : 0;
        if (p < start) {
          continue;
        }
        int base = (r == 0) ? 0 : (pl.n0 + (r - 1) * pl.W);
        int gs   = (start < pl.n_info) ? start : (start - F);
        int t    = base + gp - gs;
        if (t >= pl.E) {
          break;
        }
        int x = (pl.Qm == 1) ? in[t] : in[(t % Kq) * pl.Qm + t / Kq];
        if (pl.new_data && r == 0) {
          v = x;
        } else {
          // Which combine_softbits() call covers this position, and where in it (vector body or scalar tail).
          int left    = pl.E - base;
          int sys_len = (start < pl.n_info) ? min(pl.n_info - start, left) : 0;
          int cs, cl;
          if (p < pl.n_info) {
            cs = start, cl = sys_len;
          } else {
            cs = max(start, pl.n_sys);
            cl = min(pl.Ncb - cs, left - sys_len);
          }
          bool vec = (simd_w != 0) && ((p - cs) < (cl / simd_w) * simd_w);
          v        = vec ? pdc_llr_add_simd(v, x) : pdc_llr_add_generic(v, x);
        }
      }
    }
    if (p >= fz_from) {
      v = 0;
    }
    buf[p] = (int8_t)v;
  }
}

// First transmission that fits the circular buffer in one lap (new_data, rounds == 1): every written position receives
// exactly one LLR, so the walk of allot_llrs (:134-195) is a pure scatter. Input-major: a thread reads whole Qm-byte
// symbols (one vector load) and scatters them to the Qm de-interleaved streams; for a fixed stream neighbouring threads
// write neighbouring bytes. Same write set as pdc_dematch(): zero [0, min(k0, n_info)), fillers = +127, the LLRs, and the
// reference's final zero of the last (Ncb - end) entries of the N-long buffer.
template <int QM>
__device__ __forceinline__ void pdc_dematch_scatter(const pdc_dematch_plan& pl, const int8_t* __restrict__ in,
                                                    int8_t* __restrict__ buf, int tid, int nth)
{
  const int  F   = pl.n_sys - pl.n_info;
  const int  Kq  = pl.E / QM;
  const int  gs0 = (pl.p0 < pl.n_info) ? pl.p0 : (pl.p0 - F);
  const bool vec = (QM == 1) || ((reinterpret_cast<uintptr_t>(in) & (uintptr_t)((QM == 6) ? 1 : (QM - 1))) == 0);
  for (int i = tid; i < Kq; i += nth) {
    int8_t x[QM];
    if (vec && QM == 8) {
      const int2 w = *reinterpret_cast<const int2*>(in + (size_t)i * 8);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        x[j]     = (int8_t)(w.x >> (8 * j));
        x[4 + j] = (int8_t)(w.y >> (8 * j));
      }
    } else if (vec && QM == 4) {
      const int w = *reinterpret_cast<const int*>(in + (size_t)i * 4);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        x[j] = (int8_t)(w >> (8 * j));
      }
    } else if (vec && (QM == 2 || QM == 6)) {
#pragma unroll
      for (int j = 0; j < QM; j += 2) {
        const short w = *reinterpret_cast<const short*>(in + (size_t)i * QM + j);
        x[j]          = (int8_t)w;
        x[j + 1]      = (int8_t)(w >> 8);
      }
    } else {
#pragma unroll
      for (int j = 0; j < QM; ++j) {
        x[j] = in[(size_t)i * QM + j];
      }
    }
#pragma unroll
    for (int j = 0; j < QM; ++j) {
      const int gidx = gs0 + j * Kq + i;
      const int p    = (gidx < pl.n_info) ? gidx : (gidx + F);
      buf[p]         = x[j];
    }
  }
}

__device__ __forceinline__ void pdc_dematch_first_tx(const pdc_dematch_plan& pl, const int8_t* __restrict__ in,
                                                     int8_t* __restrict__ buf, int tid, int nth)
{
  const int zero_to = (pl.k0 < pl.n_info) ? pl.k0 : pl.n_info;
  const int fz_from = pl.final_zero ? (pl.N - (pl.Ncb - pl.end_idx)) : pl.N;
  for (int p = tid; p < zero_to; p += nth) {
    buf[p] = 0;
  }
  for (int p = pl.n_info + tid; p < pl.n_sys && p < fz_from; p += nth) {
    buf[p] = PDC_LLR_INF;
  }
  // Final zero region [fz_from, N): bytes up to the first 16-byte boundary, then whole vectors, then the tail.
  {
    const uintptr_t a0   = reinterpret_cast<uintptr_t>(buf + fz_from);
    int             head = (int)((16 - (a0 & 15)) & 15);
    head                 = min(head, pl.N - fz_from);
    const int nvec       = (pl.N - fz_from - head) >> 4;
    for (int p = tid; p < head; p += nth) {
      buf[fz_from + p] = 0;
    }
    int4* v = reinterpret_cast<int4*>(buf + fz_from + head);
    for (int k = tid; k < nvec; k += nth) {
      v[k] = make_int4(0, 0, 0, 0);
    }
    for (int p = fz_from + head + 16 * nvec + tid; p < pl.N; p += nth) {
      buf[p] = 0;
    }
  }
  switch (pl.Qm) {
    case 1:
      pdc_dematch_scatter<1>(pl, in, buf, tid, nth);
      break;
    case 2:
      pdc_dematch_scatter<2>(pl, in, buf, tid, nth);
      break;
    case 4:
      pdc_dematch_scatter<4>(pl, in, buf, tid, nth);
      break;
    case 6:
      pdc_dematch_scatter<6>(pl, in, buf, tid, nth);
      break;
    default:
      pdc_dematch_scatter<8>(pl, in, buf, tid, nth);
      break;
  }
}

// Block-wide XOR of one 32-bit value per thread; result valid in every thread. scratch: >= 32 words of shared memory.
__device__ __forceinline__ uint32_t pdc_block_xor(uint32_t x, uint32_t* scratch, int tid, int nth)
{
  x        = __reduce_xor_sync(0xffffffffu, x);
  int warp = tid >> 5, nw = (nth + 31) >> 5;
  __syncthreads();
  if ((tid & 31) == 0) {
    scratch[warp] = x;
  }
  __syncthreads();
  uint32_t r = 0;
  for (int w = 0; w < nw; ++w) {
    r ^= scratch[w];
  }
  return r;
}
