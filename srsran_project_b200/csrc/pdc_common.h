// Shared host/device definitions of the PUSCH code-block decode chain: geometry of a code block, the write plan of
// the rate dematcher and the device-side CB descriptor.
//
// Reference behaviour restated here (arithmetic only, structure is our own):
//   ldpc_rate_dematcher_impl::rate_dematch / allot_llrs   lib/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher_impl.cpp:46-201
//   ldpc_decoder_impl::decode (layer count)               lib/phy/upper/channel_coding/ldpc/ldpc_decoder_impl.cpp:83-115
#pragma once
#include "../../include/pusch_dec_cuda.h"
#include <stdint.h>

#if defined(__CUDACC__)
#define PDC_HD __host__ __device__ __forceinline__
#else
#define PDC_HD inline
#endif

#define PDC_MAX_Z 384
#define PDC_MAX_N (66 * PDC_MAX_Z) /* 25344: largest soft buffer, also the HARQ arena slot size */
#define PDC_MAX_K (22 * PDC_MAX_Z) /* 8448 */
#define PDC_LLR_MAX 120
#define PDC_LLR_INF 127
#define PDC_NOF_LIFTING_SIZES 51
#define PDC_GRAPH_STRIDE 320 /* uint32 entries per (bg, Z) graph: col | shift << 8 */

struct pdc_geom {
  int Kb, n_short, M; // info nodes, nodes after puncturing, check nodes
  int Z, K, N;        // lifting size, message bits, soft-buffer length
};

PDC_HD bool pdc_geom_init(pdc_geom& g, int bg, int Z)
{
  if (bg == 1) {
    g.Kb = 22, g.n_short = 66, g.M = 46;
  } else if (bg == 2) {
    g.Kb = 10, g.n_short = 50, g.M = 42;
  } else {
    return false;
  }
  g.Z = Z, g.K = g.Kb * Z, g.N = g.n_short * Z;
  return true;
}

// Position of Z in the ascending list of the 51 lifting sizes, -1 if invalid. Z = a * 2^j, a in {2,3,5,7,9,11,13,15}.
PDC_HD int pdc_lifting_pos(int Z)
{
  const int sizes[PDC_NOF_LIFTING_SIZES] = {2,   3,   4,   5,   6,   7,   8,   9,   10,  11,  12,  13,  14,
                                            15,  16,  18,  20,  22,  24,  26,  28,  30,  32,  36,  40,  44,
                                            48,  52,  56,  60,  64,  72,  80,  88,  96,  104, 112, 120, 128,
                                            144, 160, 176, 192, 208, 224, 240, 256, 288, 320, 352, 384};
  for (int i = 0; i != PDC_NOF_LIFTING_SIZES; ++i) {
    if (sizes[i] == Z) {
      return i;
    }
  }
  return -1;
}

// The rate dematcher's walk over the circular buffer, in closed form. The reference loop (allot_llrs :134-195) runs
// "rounds": round 0 starts at k0 (or at n_sys when k0 falls on/after the filler window) and ends at Ncb, every later
// round covers [0, n_info) u [n_sys, Ncb). Fillers [n_info, n_sys) are never visited.
struct pdc_dematch_plan {
  int N, Ncb, Z;
  int n_info, n_sys; // filler window [n_info, n_sys)
  int k0;            // start of the circular walk
  int p0;            // first visited position
  int n0;            // LLRs consumed by round 0 when the input is long enough
  int W;             // LLRs consumed by every later round (= Ncb - F)
  int rounds;        // number of rounds that consume at least one LLR
  int end_idx;       // tmp_idx when the loop ends (position after the last write, modulo Ncb)
  int E, Qm;
  bool new_data;
  bool final_zero;   // copy mode survived (single round): zero [N - (Ncb - end_idx), N) when end_idx != 0 (:198-200)
};

PDC_HD bool pdc_dematch_plan_init(pdc_dematch_plan& p, const pdc_geom& g, int E, int Qm, int rv, int Nref,
                                  int nof_filler, bool new_data)
{
  if (rv < 0 || rv > 3 || !(Qm == 1 || Qm == 2 || Qm == 4 || Qm == 6 || Qm == 8) || E <= 0 || (E % Qm) != 0) {
    return false;
  }
  p.N      = g.N;
  p.Z      = g.Z;
  p.Ncb    = (Nref > 0) ? ((Nref < g.N) ? Nref : g.N) : g.N;
  p.n_sys  = (g.Kb - 2) * g.Z;
  p.n_info = p.n_sys - nof_filler;
  if (nof_filler < 0 || nof_filler >= p.n_sys || p.Ncb <= p.n_sys) {
    return false;
  }
  // k0 = floor(f * Ncb / N) * Z (:104-105). The reference computes the quotient in double; f*Ncb < 2^21 and the
  // division is correctly rounded, so the floor equals the integer quotient.
  const int f1[4] = {0, 17, 33, 56};
  const int f2[4] = {0, 13, 25, 43};
  int       f     = (g.Kb == 22) ? f1[rv] : f2[rv];
  p.k0            = (int)(((long long)f * p.Ncb) / g.N) * g.Z;
  p.p0            = (p.k0 >= p.n_info && p.k0 < p.n_sys) ? p.n_sys : p.k0;
  int F           = nof_filler;
  p.W             = p.Ncb - F;
  int g0          = (p.p0 < p.n_info) ? p.p0 : (p.p0 - F);
  p.n0            = p.W - g0;
  p.E             = E;
  p.Qm            = Qm;
  p.new_data      = new_data;
  if (E <= p.n0) {
    p.rounds = 1;
  } else {
    p.rounds = 1 + (E - p.n0 + p.W - 1) / p.W;
  }
  // tmp_idx when the loop ends (allot_llrs :136-190), following the last round literally: a systematic chunk (if the
  // round starts before the filler window), an unconditional jump to n_sys, a parity chunk, then modulo Ncb.
  int consumed_before_last = (p.rounds == 1) ? 0 : (p.n0 + (p.rounds - 2) * p.W);
  int left                 = E - consumed_before_last;
  int start                = (p.rounds == 1) ? p.p0 : 0;
  int base                 = start;
  if (start < p.n_info) {
    int a = p.n_info - start;
    if (a > left) {
      a = left;
    }
    left -= a;
    base = p.n_sys;
  }
  int b = p.Ncb - base;
  if (b > left) {
    b = left;
  }
  int end = base + b;
  p.end_idx    = end % p.Ncb;
  p.final_zero = new_data && (p.rounds == 1) && (p.end_idx != 0);
  return true;
}

// Upper bound on nz = 1 + index of the last non-zero soft bit that the decoder can see after this dematch, used to
// bound the number of layers (and with it the shared memory) before the data is known.
PDC_HD int pdc_dematch_nz_bound(const pdc_dematch_plan& p)
{
  if (!p.new_data) {
    return p.N; // older transmissions may have written anywhere
  }
  if (p.rounds > 1) {
    return (p.Ncb < p.N) ? p.N : p.Ncb; // wrapped: everything up to Ncb may be non-zero; beyond Ncb the buffer is stale
  }
  if (p.final_zero) {
    // Written positions end at end_idx; [N-(Ncb-end_idx), N) is zeroed; with Ncb < N the stretch
    // [Ncb, N-(Ncb-end_idx)) keeps whatever the buffer held.
    return (p.Ncb < p.N) ? (p.N - (p.Ncb - p.end_idx)) : p.end_idx;
  }
  return p.Ncb < p.N ? p.N : p.Ncb;
}

// nof_layers of ldpc_decoder_impl::decode (:102-115) for a given nz.
PDC_HD int pdc_layers_for_nz(const pdc_geom& g, int nz)
{
  int len = nz + 2 * g.Z;
  int mn  = g.K + 4 * g.Z;
  if (len < mn) {
    len = mn;
  }
  len = ((len + g.Z - 1) / g.Z) * g.Z;
  return len / g.Z - g.Kb;
}

// pdc_dev_cb.flags only (never in pdc_cb_desc.flags): the host planner found that this code block is a first transmission from
// the start of the buffer that fits the packed kernel's staging tile (new data, rv 0, no LBRM, one lap, n_info <= E,
// E + F <= (Kb + 2) Z): the kernel then needs no dematcher plan (no integer divisions), only n_info = n_sys - F and
// end = E + F.
#define PDC_DEVF_FIRST_TX_TILE 0x80u

// Device-side view of one transport block (built by the host runtime from pdc_tb_desc).
struct pdc_dev_tb {
  uint8_t*       out;      // tbs_bits / 8 payload bytes, 4-byte aligned
  pdc_tb_result* res;
  uint32_t*      terms;    // [nof_cbs] scratch: every code block's share of the TB CRC remainder
  uint32_t       first_cb, nof_cbs, tbs_bits;
  uint32_t       cb_bits;  // C > 1: data bits per code block D = K - F - 24
  uint32_t       last_bits; // C > 1: bits of payload || TB CRC in the last code block, tbs + 24 - (C - 1) D (then zero padding)
};

// Device-side view of one code block (built by the host runtime from pdc_cb_desc).
struct pdc_dev_cb {
  const int8_t*  llrs; // E rate-matched LLRs, or the n_in soft bits in decode-only mode
  int8_t*        soft; // N-entry soft buffer (in/out); nullptr in decode-only mode
  uint8_t*       out;  // packed hard bits, 4-byte aligned
  pdc_cb_result* res;
  uint32_t       E, Nref, nof_filler;
  uint16_t       Z;
  uint8_t        bg, Qm, rv, crc_type, max_iter, flags;
  uint16_t       graph_idx;  // (bg - 1) * 51 + lifting position
  uint16_t       max_layers; // host bound used to size shared memory
};
