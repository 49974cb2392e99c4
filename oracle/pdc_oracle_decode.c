/* TEST INFRASTRUCTURE ONLY — CPU oracle, LDPC decoder part (see pdc_oracle.c for the status header).
 *
 * Scalar restatement of ldpc_decoder_impl (lib/phy/upper/channel_coding/ldpc/ldpc_decoder_impl.cpp:57-350) with the
 * per-lane arithmetic of ldpc_decoder_avx512.cpp:86-270 (flavour AVX512/AVX2) or ldpc_decoder_generic.cpp:30-122
 * (flavour generic). Lanes are independent, so the SIMD node padding of the reference has no observable effect.
 */
#include "pdc_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define LLR_MAX 120
#define LLR_INF 127
#define MAX_Z 384
#define MAX_DEG 20

static int clampi(int x, int lo, int hi)
{
  return x < lo ? lo : (x > hi ? hi : x);
}

/* Min-sum normalisation. AVX2/AVX512: mm512::scale_epi8 (avx512_support.h:65-107) = (x*52428)>>16, where
 * 52428 = (uint16_t)(0.8f*65536). Generic: scale_llr (ldpc_decoder_generic.cpp:70-79) = round(0.8f*x). */
static int scale_min(int flavour, int x)
{
  if (flavour == ORC_FLAVOUR_GENERIC) {
    return (int)roundf((float)x * 0.8f);
  }
  return (x * 52428) >> 16;
}

/* compute_var_to_check_msgs (avx512 :86-126; generic :30-44 via log_likelihood_ratio::operator-). */
static int v2c_msg(int soft, int c2v)
{
  if (soft >= LLR_INF) {
    return LLR_INF;
  }
  if (soft <= -LLR_INF) {
    return -LLR_INF;
  }
  return clampi(soft - c2v, -LLR_MAX, LLR_MAX);
}

/* compute_soft_bits (avx512 :229-270; generic :110-122 via promotion_sum, log_likelihood_ratio.cpp:75-88). The
 * check-to-variable term is always finite here (|c2v| <= 96). */
static int soft_update(int v2c, int c2v)
{
  if (v2c >= LLR_INF) {
    return LLR_INF;
  }
  if (v2c <= -LLR_INF) {
    return -LLR_INF;
  }
  int t = v2c + c2v;
  if (t > LLR_MAX) {
    return LLR_INF;
  }
  if (t < -LLR_MAX) {
    return -LLR_INF;
  }
  return t;
}

/* hard_decision (log_likelihood_ratio.cpp:350-376) + get_hard_bits (ldpc_decoder_impl.cpp:331-350): bit = (soft <= 0),
 * MSB-first packing; returns 0 if any of the K soft bits is exactly zero. */
static int hard_bits(uint8_t* out_packed, const int8_t* soft, int K)
{
  int valid = 1;
  memset(out_packed, 0, (size_t)((K + 7) / 8));
  for (int n = 0; n != K; ++n) {
    if (soft[n] == 0) {
      valid = 0;
    }
    if (soft[n] <= 0) {
      out_packed[n >> 3] |= (uint8_t)(0x80u >> (n & 7));
    }
  }
  return valid;
}

int orc_ldpc_decode(int flavour, int bg, int Z, int nof_filler, int max_iter, int crc_id, int force_decoding,
                    const int8_t* llrs, int n_llrs, uint8_t* out_packed)
{
  orc_graph g;
  if (orc_graph_init(&g, bg, Z) != 0 || max_iter <= 0) {
    return -2;
  }
  int K = g.Kb * Z;
  /* decode :64-80 — input length contract. */
  if (n_llrs > g.N_short * Z || n_llrs < K + 2 * Z) {
    return -2;
  }
  /* :83-87 — index after the last non-zero LLR. */
  int input_size = n_llrs;
  while (input_size > 0 && llrs[input_size - 1] == 0) {
    --input_size;
  }
  /* :91-97 — NB the flag is inverted with respect to its documentation; restated as written. */
  if (input_size < K && force_decoding) {
    if (crc_id == ORC_CRC_NONE) {
      /* bit_buffer::one() sets whole bytes to 0xff, the last partial byte MSB-aligned. */
      memset(out_packed, 0, (size_t)((K + 7) / 8));
      for (int n = 0; n != K; ++n) {
        out_packed[n >> 3] |= (uint8_t)(0x80u >> (n & 7));
      }
    }
    return -1;
  }

  int     n_nodes_max = g.N_full;
  int8_t* soft        = (int8_t*)calloc((size_t)(n_nodes_max * Z), 1);
  int8_t* c2v         = (int8_t*)calloc((size_t)(g.nedges * Z), 1);
  uint8_t seen[46];
  memset(seen, 0, sizeof(seen));

  /* load_soft_bits :150-191 — two punctured nodes of zeros, then clamp(in, -64, 64) for every whole node that still
   * has LLRs (nof_llr != 0 is evaluated per node), zero afterwards; a non-multiple-of-Z tail is copied unclamped. */
  {
    int nof_full_nodes = n_llrs / Z + 2;
    int nof_llr        = input_size;
    int pos            = 0;
    for (int node = 2; node != nof_full_nodes; ++node) {
      if (nof_llr != 0) {
        for (int j = 0; j != Z; ++j) {
          soft[node * Z + j] = (int8_t)clampi(llrs[pos + j], -64, 64);
        }
      }
      pos += Z;
      nof_llr = (nof_llr >= Z) ? (nof_llr - Z) : 0;
    }
    int tail = n_llrs - pos;
    for (int j = 0; j != tail; ++j) {
      soft[nof_full_nodes * Z + j] = llrs[pos + j];
    }
  }

  /* :102-115 — number of layers actually processed. */
  int cb_len = input_size + 2 * Z;
  if (cb_len < K + 4 * Z) {
    cb_len = K + 4 * Z;
  }
  if (cb_len % Z != 0) {
    cb_len = (cb_len / Z + 1) * Z;
  }
  int nof_layers = cb_len / Z - g.Kb;
  int n_sig      = K - nof_filler;
  int result     = -1;

  int8_t v2c[MAX_DEG][MAX_Z];
  for (int it = 0; it != max_iter; ++it) {
    for (int m = 0; m != nof_layers; ++m) {
      int            deg   = g.row_deg[m];
      const uint8_t* cols  = g.edge_col + g.row_start[m];
      int8_t*        c2v_m = c2v + (size_t)g.row_start[m] * Z;
      /* update_variable_to_check_messages :193-236 (var-lane order). */
      for (int e = 0; e != deg; ++e) {
        const int8_t* s = soft + cols[e] * Z;
        for (int j = 0; j != Z; ++j) {
          v2c[e][j] = seen[m] ? (int8_t)v2c_msg(s[j], c2v_m[e * Z + j]) : s[j];
        }
      }
      /* update_check_to_variable_messages :253-329 (check lane i sees v2c[(i + shift) mod Z]). */
      for (int i = 0; i != Z; ++i) {
        int min1 = LLR_MAX, min2 = LLR_MAX, arg = 0, par = 0;
        for (int e = 0; e != deg; ++e) {
          int sh = g.edge_shift[g.row_start[m] + e] % Z;
          int r  = v2c[e][(i + sh) % Z];
          int a  = abs(r);
          /* analyze_var_to_check_msgs (avx512 :128-170 / generic :46-68). */
          int is_min  = a < min1;
          int new_2nd = is_min ? min1 : a;
          if (a < min2) {
            min2 = new_2nd;
          }
          if (is_min) {
            min1 = a, arg = e;
          }
          par ^= (r < 0);
        }
        int s1 = scale_min(flavour, min1), s2 = scale_min(flavour, min2);
        /* compute_check_to_var_msgs (avx512 :183-227 / generic :85-108), rotated back to var-lane order. */
        for (int e = 0; e != deg; ++e) {
          int sh             = g.edge_shift[g.row_start[m] + e] % Z;
          int j              = (i + sh) % Z;
          int mag            = (e == arg) ? s2 : s1;
          int neg            = par ^ (v2c[e][j] < 0);
          c2v_m[e * Z + j] = (int8_t)(neg ? -mag : mag);
        }
      }
      /* update_soft_bits :238-251. */
      for (int e = 0; e != deg; ++e) {
        int8_t* s = soft + cols[e] * Z;
        for (int j = 0; j != Z; ++j) {
          s[j] = (int8_t)soft_update(v2c[e][j], c2v_m[e * Z + j]);
        }
      }
      seen[m] = 1;
    }
    /* :127-135 — early stop. */
    if (crc_id != ORC_CRC_NONE) {
      int ok = hard_bits(out_packed, soft, K);
      if (ok && orc_crc_bits(crc_id, out_packed, n_sig) == 0) {
        result = it + 1;
        break;
      }
    }
  }
  if (crc_id == ORC_CRC_NONE) {
    hard_bits(out_packed, soft, K); /* :144-147 */
  }
  free(soft);
  free(c2v);
  return result;
}

/* pusch_codeblock_decoder::decode (lib/phy/upper/channel_processors/pusch/pusch_codeblock_decoder.cpp:35-74):
 * rate_match() then decode with the CRC calculator (early stop) or decode without and check the CRC once. */
int orc_codeblock_decode(int flavour, int bg, int Z, int8_t* buf, const int8_t* in, int E, int new_data, int rv,
                         int Qm, int Nref, int nof_filler, int crc_id, int max_iter, int use_early_stop,
                         int force_decoding, uint8_t* out_packed)
{
  int N = ((bg == 1) ? 66 : 50) * Z;
  int K = ((bg == 1) ? 22 : 10) * Z;
  if (orc_rate_dematch(flavour, buf, N, in, E, new_data, rv, Qm, Nref, nof_filler) != 0) {
    return -2;
  }
  if (use_early_stop) {
    return orc_ldpc_decode(flavour, bg, Z, nof_filler, max_iter, crc_id, force_decoding, buf, N, out_packed);
  }
  int r = orc_ldpc_decode(flavour, bg, Z, nof_filler, max_iter, ORC_CRC_NONE, force_decoding, buf, N, out_packed);
  if (r == -2) {
    return r;
  }
  /* :68-73 — without early stop the decoder returns nullopt; a single CRC decides and reports max_iter. */
  if (orc_crc_bits(crc_id, out_packed, K - nof_filler) == 0) {
    return max_iter;
  }
  return -1;
}
