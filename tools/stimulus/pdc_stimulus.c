/* STIMULUS GENERATOR (test / benchmark input only; not on the product path, not the oracle).
 *
 * The decode chain needs valid codewords as input. This file restates the TRANSMIT side of TS 38.212: Section 5.1 CRC
 * attachment, 5.2.2 code-block segmentation, 5.3.2 LDPC encoding and 5.4.2 rate matching + bit interleaving. The
 * encoder and rate matcher are checked bit for bit against the reference's ldpc_encoder_{generic,avx2}
 * (lib/phy/upper/channel_coding/ldpc/ldpc_encoder_impl.cpp) and ldpc_rate_matcher_impl (ldpc_rate_matcher_impl.cpp) in
 * tests/test_oracle_vs_ref.py, the segmentation against ldpc_segmenter_rx_impl (ldpc_segmenter_rx_impl.cpp:52-106) in
 * tests/test_stimulus.py. Bits are one per byte (0/1).
 */
#include "../../srsran_project_b200/csrc/ldpc_bg_tables.h"
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  int             Kb, N_full, N_short, M, nedges;
  const uint8_t*  row_deg;
  const uint8_t*  edge_col;
  const uint16_t* edge_shift;
  int             row_start[47];
} orc_graph;

static int lifting_set_index(int Z)
{
  static const int base[8] = {2, 3, 5, 7, 9, 11, 13, 15};
  if (Z < 2 || Z > 384) {
    return -1;
  }
  for (int i = 0; i != 8; ++i) {
    int z = base[i];
    while (z < Z) {
      z *= 2;
    }
    if (z == Z) {
      return i;
    }
  }
  return -1;
}

static int orc_graph_init(orc_graph* g, int bg, int Z)
{
  int ils = lifting_set_index(Z);
  if (ils < 0 || (bg != 1 && bg != 2)) {
    return -1;
  }
  if (bg == 1) {
    g->Kb = 22, g->N_full = 68, g->N_short = 66, g->M = 46, g->nedges = PDC_BG1_EDGES;
    g->row_deg = pdc_bg1_row_deg, g->edge_col = pdc_bg1_edge_col, g->edge_shift = pdc_bg1_edge_shift[ils];
  } else {
    g->Kb = 10, g->N_full = 52, g->N_short = 50, g->M = 42, g->nedges = PDC_BG2_EDGES;
    g->row_deg = pdc_bg2_row_deg, g->edge_col = pdc_bg2_edge_col, g->edge_shift = pdc_bg2_edge_shift[ils];
  }
  int acc = 0;
  for (int m = 0; m != g->M; ++m) {
    g->row_start[m] = acc;
    acc += g->row_deg[m];
  }
  g->row_start[g->M] = acc;
  return 0;
}

/* acc[i] ^= x[(i + shift) mod Z] — multiplication of a length-Z block by the circulant P^shift. */
static void xor_rot(uint8_t* acc, const uint8_t* x, int shift, int Z)
{
  for (int i = 0; i != Z; ++i) {
    acc[i] ^= x[(i + shift) % Z];
  }
}

/* Shift of edge (m, col) or -1 when the base graph has no such edge. */
static int edge_shift(const orc_graph* g, int m, int col, int Z)
{
  for (int e = g->row_start[m]; e != g->row_start[m + 1]; ++e) {
    if (g->edge_col[e] == col) {
      return g->edge_shift[e] % Z;
    }
  }
  return -1;
}

/* msg_bits: K = Kb*Z bits (fillers as 0). cb_bits: N = (N_full-2)*Z bits, i.e. the codeword with its first 2Z
 * systematic bits punctured, as ldpc_encoder_buffer::write_codeblock(offset 0) delivers it. */
int stim_ldpc_encode(int bg, int Z, const uint8_t* msg_bits, uint8_t* cb_bits)
{
  orc_graph g;
  if (orc_graph_init(&g, bg, Z) != 0) {
    return -1;
  }
  int      Kb = g.Kb;
  uint8_t* cw = (uint8_t*)calloc((size_t)(g.N_full * Z), 1);
  memcpy(cw, msg_bits, (size_t)(Kb * Z));

  /* lambda_m = sum over systematic columns of P^shift * s_col, for the four core rows. */
  uint8_t* lam = (uint8_t*)calloc((size_t)(4 * Z), 1);
  for (int m = 0; m != 4; ++m) {
    for (int e = g.row_start[m]; e != g.row_start[m + 1]; ++e) {
      int col = g.edge_col[e];
      if (col < Kb) {
        xor_rot(lam + m * Z, cw + col * Z, g.edge_shift[e] % Z, Z);
      }
    }
  }
  /* Core parity: column Kb carries three circulants of which two are equal, columns Kb+1..Kb+3 form the double
   * diagonal (rows m and m+1 hold the identity). Summing the four rows leaves P^b * p1. */
  uint8_t* sum = (uint8_t*)calloc((size_t)Z, 1);
  for (int m = 0; m != 4; ++m) {
    for (int i = 0; i != Z; ++i) {
      sum[i] ^= lam[m * Z + i];
    }
  }
  int c[4], b = -1;
  for (int m = 0; m != 4; ++m) {
    c[m] = edge_shift(&g, m, Kb, Z);
  }
  for (int m = 0; m != 4; ++m) {
    if (c[m] < 0) {
      continue;
    }
    int cnt = 0;
    for (int k = 0; k != 4; ++k) {
      cnt += (c[k] == c[m]);
    }
    if (cnt % 2 == 1) {
      b = c[m];
    }
  }
  uint8_t* p1 = cw + Kb * Z;
  for (int i = 0; i != Z; ++i) {
    p1[(i + b) % Z] = sum[i]; /* sum[i] = p1[(i+b) mod Z] */
  }
  /* Rows 0..2 give p2..p4 in turn: lambda_m + [P^c_m p1] + p_{m+1} (m>0) + p_{m+2} = 0. */
  for (int m = 0; m != 3; ++m) {
    uint8_t* pn = cw + (Kb + 1 + m) * Z;
    memcpy(pn, lam + m * Z, (size_t)Z);
    if (c[m] >= 0) {
      xor_rot(pn, p1, c[m], Z);
    }
    if (m > 0) {
      xor_rot(pn, cw + (Kb + m) * Z, 0, Z);
    }
  }
  /* Extension rows: the parity block of row m >= 4 is the sum of everything else on the row. */
  for (int m = 4; m != g.M; ++m) {
    uint8_t* pe = cw + (Kb + m) * Z;
    for (int e = g.row_start[m]; e != g.row_start[m + 1]; ++e) {
      int col = g.edge_col[e];
      if (col != Kb + m) {
        xor_rot(pe, cw + col * Z, g.edge_shift[e] % Z, Z);
      }
    }
  }
  memcpy(cb_bits, cw + 2 * Z, (size_t)((g.N_full - 2) * Z));
  free(cw);
  free(lam);
  free(sum);
  return 0;
}

/* Bit selection from the circular buffer (TS 38.212 5.4.2.1; ldpc_rate_matcher_impl.cpp select_bits) followed by the
 * row-column bit interleaver (5.4.2.2; interleave_bits): out[i*Qm + j] = sel[j*(E/Qm) + i]. */
int stim_rate_match(int bg, int Z, const uint8_t* cb_bits, int rv, int Qm, int Nref, int nof_filler, int E,
                   uint8_t* out_bits)
{
  static const int f_bg1[4] = {0, 17, 33, 56};
  static const int f_bg2[4] = {0, 13, 25, 43};
  const int*       f        = (bg == 1) ? f_bg1 : f_bg2;
  int              N        = ((bg == 1) ? 66 : 50) * Z;
  int              Kb       = (bg == 1) ? 22 : 10;
  int              Ncb      = (Nref > 0) ? ((Nref < N) ? Nref : N) : N;
  int              n_sys    = (Kb - 2) * Z;
  int              n_info   = n_sys - nof_filler;
  int              k0       = (int)floor(((double)f[rv] * (double)Ncb) / (double)N) * Z;
  if (E % Qm != 0) {
    return -1;
  }
  uint8_t* sel = (uint8_t*)malloc((size_t)E);
  int      e = 0, j = 0;
  while (e < E) {
    int idx = (k0 + j) % Ncb;
    if (!(idx >= n_info && idx < n_sys)) {
      sel[e++] = cb_bits[idx];
    }
    ++j;
  }
  int K = E / Qm;
  for (int i = 0; i != K; ++i) {
    for (int q = 0; q != Qm; ++q) {
      out_bits[i * Qm + q] = sel[q * K + i];
    }
  }
  free(sel);
  return 0;
}

/* CRC of nbits bits (one per byte), MSB first, init 0 (TS 38.212 5.1). crc_id: 1 = CRC16, 2 = CRC24A, 3 = CRC24B. */
uint32_t stim_crc(int crc_id, const uint8_t* bits, int nbits)
{
  uint32_t poly  = (crc_id == 1) ? 0x11021u : ((crc_id == 2) ? 0x1864CFBu : 0x1800063u);
  int      order = (crc_id == 1) ? 16 : 24;
  uint32_t top = 1u << order, reg = 0;
  for (int n = 0; n != nbits; ++n) {
    reg = (reg << 1) ^ ((uint32_t)(bits[n] & 1) << order);
    if (reg & top) {
      reg ^= poly;
    }
  }
  return reg & (top - 1);
}

typedef struct {
  int bg, Z, C, K, Kp, F, tb_crc_bits, cb_crc_bits, B;
} stim_seg;

/* TS 38.212 5.2.2 segmentation of a transport block of tbs bits on base graph bg (ldpc.h:128-214 in the reference). */
int stim_segmentation(int bg, int tbs, stim_seg* s)
{
  static const int sizes[51] = {2,   3,   4,   5,   6,   7,   8,   9,   10,  11,  12,  13,  14,  15,  16,  18,  20,
                                22,  24,  26,  28,  30,  32,  36,  40,  44,  48,  52,  56,  60,  64,  72,  80,  88,
                                96,  104, 112, 120, 128, 144, 160, 176, 192, 208, 224, 240, 256, 288, 320, 352, 384};
  s->bg          = bg;
  s->tb_crc_bits = (tbs > 3824) ? 24 : 16;
  s->B           = tbs + s->tb_crc_bits;
  int Kcb        = (bg == 1) ? 8448 : 3840;
  if (s->B <= Kcb) {
    s->C = 1, s->cb_crc_bits = 0;
  } else {
    s->C           = (s->B + (Kcb - 24) - 1) / (Kcb - 24);
    s->cb_crc_bits = 24;
  }
  int Bp = s->B + s->C * s->cb_crc_bits;
  s->Kp  = (Bp + s->C - 1) / s->C; /* the reference pads the TB so that every CB carries ceil(B'/C) bits */
  int Kb = 22;
  if (bg == 2) {
    Kb = (s->B > 640) ? 10 : ((s->B > 560) ? 9 : ((s->B > 192) ? 8 : 6));
  }
  s->Z = -1;
  for (int i = 0; i != 51; ++i) {
    if (Kb * sizes[i] >= s->Kp) {
      s->Z = sizes[i];
      break;
    }
  }
  if (s->Z < 0) {
    return -1;
  }
  s->K = ((bg == 1) ? 22 : 10) * s->Z;
  s->F = s->K - s->Kp;
  return 0;
}

/* Rate-matched length of code block r (TS 38.212 5.4.2.1; ldpc_segmenter_helpers.h:82-100). */
int stim_rm_length(int r, int C, int G, int nof_layers, int Qm)
{
  int sym = G / (nof_layers * Qm);
  int base = sym / C, rem = sym % C;
  return nof_layers * Qm * ((r < C - rem) ? base : base + 1);
}

/* Whole transmit chain of one transport block: tb_bits (tbs bits, one per byte) -> TB CRC -> segmentation -> CB CRC ->
 * fillers -> LDPC encoding -> rate matching. out_bits receives the G rate-matched bits of all code blocks, back to
 * back. msgs (optional) receives the C messages of K bits each (payload | CB CRC | fillers as 0). Returns C. */
int stim_tb_encode(int bg, int tbs, const uint8_t* tb_bits, int rv, int Qm, int Nref, int nof_layers, int G,
                   uint8_t* out_bits, uint8_t* msgs)
{
  stim_seg s;
  if (stim_segmentation(bg, tbs, &s) != 0) {
    return -1;
  }
  int      N    = ((bg == 1) ? 66 : 50) * s.Z;
  uint8_t* tb   = (uint8_t*)calloc((size_t)(s.C * s.Kp), 1); /* zero padding after the TB CRC, as the reference does */
  uint8_t* msg  = (uint8_t*)calloc((size_t)s.K, 1);
  uint8_t* cb   = (uint8_t*)malloc((size_t)N);
  memcpy(tb, tb_bits, (size_t)tbs);
  uint32_t c = stim_crc((s.tb_crc_bits == 24) ? 2 : 1, tb_bits, tbs);
  for (int k = 0; k != s.tb_crc_bits; ++k) {
    tb[tbs + k] = (uint8_t)((c >> (s.tb_crc_bits - 1 - k)) & 1u);
  }
  int off = 0, per_cb = s.Kp - s.cb_crc_bits, pos = 0;
  for (int r = 0; r != s.C; ++r) {
    memset(msg, 0, (size_t)s.K);
    memcpy(msg, tb + pos, (size_t)per_cb);
    pos += per_cb;
    if (s.cb_crc_bits) {
      uint32_t cc = stim_crc(3, msg, per_cb);
      for (int k = 0; k != 24; ++k) {
        msg[per_cb + k] = (uint8_t)((cc >> (23 - k)) & 1u);
      }
    }
    if (msgs) {
      memcpy(msgs + (size_t)r * s.K, msg, (size_t)s.K);
    }
    stim_ldpc_encode(bg, s.Z, msg, cb);
    int E = stim_rm_length(r, s.C, G, nof_layers, Qm);
    stim_rate_match(bg, s.Z, cb, rv, Qm, Nref, s.F, E, out_bits + off);
    off += E;
  }
  free(tb), free(msg), free(cb);
  return s.C;
}

/* LLR-domain AWGN channel: llr = clamp(round(amp * (1 - 2 b) + sigma * n), -120, 120), n ~ N(0,1) from a 64-bit
 * xorshift generator + Box-Muller (deterministic for a given seed, independent of the platform's libm to ~1 ulp; the
 * generated LLRs are inputs, so tiny libm differences would change the stimulus, never the parity of a comparison). */
void stim_awgn_llrs(const uint8_t* bits, int n, double amp, double sigma, uint64_t seed, int8_t* llrs)
{
  uint64_t s = seed * 0x9E3779B97F4A7C15ull + 0xD1B54A32D192ED03ull;
  for (int i = 0; i < n; i += 2) {
    double u[2];
    for (int k = 0; k != 2; ++k) {
      s ^= s << 13, s ^= s >> 7, s ^= s << 17;
      u[k] = ((double)(s >> 11) + 0.5) / 9007199254740992.0;
    }
    double r = sqrt(-2.0 * log(u[0])), a = 6.283185307179586 * u[1];
    double g[2] = {r * cos(a), r * sin(a)};
    for (int k = 0; k != 2 && i + k < n; ++k) {
      double x = amp * (1.0 - 2.0 * (double)bits[i + k]) + sigma * g[k];
      x        = floor(x + 0.5);
      x        = (x > 120.0) ? 120.0 : ((x < -120.0) ? -120.0 : x);
      llrs[i + k] = (int8_t)x;
    }
  }
}
