/* TEST INFRASTRUCTURE ONLY — CPU oracle, transmit-side stimulus (see pdc_oracle.c for the status header).
 *
 * The decode chain under test needs valid codewords as input. These two functions restate the transmit side of
 * TS 38.212 Sections 5.3.2 (LDPC encoding) and 5.4.2 (rate matching + bit interleaving); they are checked bit for bit
 * against the reference's ldpc_encoder_{generic,avx2} (lib/phy/upper/channel_coding/ldpc/ldpc_encoder_impl.cpp) and
 * ldpc_rate_matcher_impl (ldpc_rate_matcher_impl.cpp) in tests/test_oracle_vs_ref.py. Bits are one per byte (0/1).
 */
#include "pdc_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

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
int orc_ldpc_encode(int bg, int Z, const uint8_t* msg_bits, uint8_t* cb_bits)
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
int orc_rate_match(int bg, int Z, const uint8_t* cb_bits, int rv, int Qm, int Nref, int nof_filler, int E,
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
