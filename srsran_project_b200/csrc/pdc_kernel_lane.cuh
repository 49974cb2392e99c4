// Code-block kernel, one lifting-size lane per thread ("lane kernel").
//
// One CTA decodes one code block end to end: rate dematch + HARQ combine (prologue, global soft buffer) -> layered
// normalised min-sum over shared-memory soft bits and check-to-variable messages -> CB CRC / early stop (epilogue).
// This variant keeps one lane of the Z-circulant per thread and plain int32 arithmetic; it handles every lifting size
// (in particular the 15 sizes that are not multiples of 4 and all Z < 32) and both arithmetic flavours. The packed
// fp16x2 kernel in pdc_kernel_h2.cu takes over for the large lifting sizes.
//
// Differences in STRUCTURE from the reference (results are identical): check-to-variable messages are kept in check-lane
// order (never rotated back), variable-to-check messages are recomputed instead of stored twice, and "first visit of a
// layer" is simply "first iteration" because every decode processes the same layers in every iteration
// (ldpc_decoder_impl.cpp:100,117-124).
#pragma once
#include "pdc_device.cuh"

// One row (layer) for one lane. The edges of a row are independent until the two minima: with the degree as a template
// parameter every load of the row is issued before the first result is needed, and the variable-to-check messages stay in
// registers for the second pass. A code block of a small lifting size is ONE warp running up to 46 x 6 rows in sequence: its
// latency per row is what a mixed batch waits for. Row degrees: 3..10 and 19 (ldpc_bg_tables.h).
template <int DEG>
__device__ __forceinline__ void pdc_lane_row(const uint32_t* __restrict__ graph, int8_t* __restrict__ soft, int8_t* __restrict__ c2v,
                                             int Z, int tid, bool first, bool generic)
{
  int v[DEG], sidx[DEG];
#pragma unroll
  for (int e = 0; e < DEG; ++e) {
    const uint32_t ent = graph[e];
    int            j   = tid + (int)(ent >> 8);
    j -= (j >= Z) ? Z : 0;
    sidx[e]     = (int)(ent & 0xff) * Z + j;
    const int s = soft[sidx[e]];
    const int c = first ? 0 : c2v[e * Z + tid];
    v[e]        = pdc_v2c(s, c);
  }
  int min1 = PDC_LLR_MAX, min2 = PDC_LLR_MAX, arg = 0, par = 0;
#pragma unroll
  for (int e = 0; e < DEG; ++e) {
    const int  a      = abs(v[e]);
    const bool is_min = a < min1;
    const int  nsec   = is_min ? min1 : a;
    min2              = (a < min2) ? nsec : min2;
    arg               = is_min ? e : arg;
    min1              = is_min ? a : min1;
    par ^= (v[e] < 0);
  }
  const int s1 = pdc_scale(min1, generic), s2 = pdc_scale(min2, generic);
#pragma unroll
  for (int e = 0; e < DEG; ++e) {
    const int mag    = (e == arg) ? s2 : s1;
    const int nc     = (par ^ (v[e] < 0)) ? -mag : mag;
    c2v[e * Z + tid] = (int8_t)nc;
    soft[sidx[e]]    = (int8_t)pdc_soft_update(v[e], nc);
  }
}

__global__ void __launch_bounds__(384) pdc_cb_kernel_lane(pdc_launch_params prm)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint32_t s_graph[PDC_GRAPH_STRIDE];
  __shared__ uint32_t s_scratch[32];
  __shared__ int      s_nz;

  const int tid = threadIdx.x, nth = blockDim.x;
  const int cbi = prm.order ? prm.order[blockIdx.x] : blockIdx.x;
  const pdc_dev_cb cb = prm.cbs[cbi];

  pdc_cb_result res;
  res.iterations = -1, res.crc_pass = 0, res.status = 0, res.nof_layers = 0;

  pdc_geom g;
  pdc_geom_init(g, cb.bg, cb.Z);
  const int  Z       = g.Z;
  const bool generic = (cb.flags & PDC_F_GENERIC) != 0;
  const int  bgi     = cb.bg - 1;

  // ---- prologue: rate dematching + HARQ combining into the N-entry soft buffer in HBM --------------------------------
  const int8_t* dec_in;
  int           n_in;
  if (cb.flags & PDC_F_DECODE_ONLY) {
    dec_in = cb.llrs;
    n_in   = (int)cb.E;
  } else {
    pdc_dematch_plan pl;
    pdc_dematch_plan_init(pl, g, (int)cb.E, cb.Qm, cb.rv, (int)cb.Nref, (int)cb.nof_filler,
                          (cb.flags & PDC_F_NEW_DATA) != 0);
    pdc_dematch(pl, cb.llrs, cb.soft, generic, tid, nth);
    dec_in = cb.soft;
    n_in   = g.N;
    if (cb.flags & PDC_F_DEMATCH_ONLY) {
      if (tid == 0) {
        *cb.res = res;
      }
      return;
    }
  }
  for (int i = tid; i < PDC_GRAPH_STRIDE; i += nth) {
    s_graph[i] = prm.graphs[(size_t)cb.graph_idx * PDC_GRAPH_STRIDE + i];
  }
  if (tid == 0) {
    s_nz = 0;
  }
  __syncthreads(); // also orders the dematcher's global writes before the reads below (same CTA)

  // ---- index after the last non-zero soft bit (ldpc_decoder_impl.cpp:83-87) ------------------------------------------
  {
    int best = 0;
    for (int n = tid; n < n_in; n += nth) {
      if (dec_in[n] != 0) {
        best = n + 1;
      }
    }
    best = __reduce_max_sync(0xffffffffu, best);
    if ((tid & 31) == 0 && best > 0) {
      atomicMax(&s_nz, best);
    }
  }
  __syncthreads();
  const int nz = s_nz;
  const int K  = g.K;

  // ---- not enough soft bits (:91-97; the flag is inverted w.r.t. its documentation, as in the reference) ---------------
  if (nz < K && (cb.flags & PDC_F_FORCE_DECODING)) {
    if (cb.crc_type == PDC_CRC_NONE) {
      uint32_t* out32 = reinterpret_cast<uint32_t*>(cb.out);
      for (int w = tid; w < (K + 31) / 32; w += nth) {
        int      rem = K - 32 * w;
        uint32_t v   = 0xffffffffu;
        if (rem < 32) { // bytes are MSB-first, the 32-bit word is little-endian
          v = 0;
          for (int k = 0; k < rem; ++k) {
            v |= (0x80u >> (k & 7)) << (8 * (k >> 3));
          }
        }
        out32[w] = v;
      }
    } else {
      res.status = 2; // output left untouched, as the reference does
    }
    if (tid == 0) {
      *cb.res = res;
    }
    return;
  }

  const int L = pdc_layers_for_nz(g, nz);
  if (L > (int)cb.max_layers) { // cannot happen unless the host bound is wrong; fail loudly instead of overrunning
    res.status = 3;
    if (tid == 0) {
      *cb.res = res;
    }
    return;
  }
  res.nof_layers = (uint16_t)L;

  int8_t* soft = reinterpret_cast<int8_t*>(smem_raw);         // [(Kb + L) * Z]
  int8_t* c2v  = soft + ((g.Kb + (int)cb.max_layers) * Z + 15) / 16 * 16; // [edges(L) * Z], check-lane order

  // ---- load_soft_bits (:150-191): two punctured nodes, clamp to +-64, zero beyond the input ---------------------------
  {
    const int n_soft   = (g.Kb + L) * Z;
    const int n_clamp  = (n_in / Z) * Z; // a tail that does not fill a node is copied unclamped (:184-190)
    for (int n = tid; n < n_soft; n += nth) {
      int v = 0;
      int k = n - 2 * Z;
      if (k >= 0 && k < n_in) {
        v = dec_in[k];
        if (k < n_clamp) {
          v = pdc_clamp(v, -64, 64);
        }
      }
      soft[n] = (int8_t)v;
    }
  }
  __syncthreads();

  const uint16_t* row_start = c_bg_rows.start[bgi];
  const uint8_t*  row_deg   = c_bg_rows.deg[bgi];
  const uint32_t* crc_pow   = (cb.crc_type != PDC_CRC_NONE) ? prm.crc_pow + (cb.crc_type - 1) * PDC_CRC_TABLE_LEN : nullptr;
  const int       crc_order = (cb.crc_type == PDC_CRC16) ? 16 : 24;
  const int       n_sig     = K - (int)cb.nof_filler;
  const bool      early     = (cb.flags & PDC_F_EARLY_STOP) && cb.crc_type != PDC_CRC_NONE;
  const int       max_iter  = cb.max_iter;

  int  it_done = 0;
  bool success = false;
  for (int it = 0; it < max_iter && !success; ++it) {
    for (int m = 0; m < L; ++m) {
      if (tid < Z) {
        // The degree is uniform over the thread block: one fully unrolled body per degree (see pdc_lane_row).
        const int e0 = row_start[m];
        switch (row_deg[m]) {
#define PDC_LANE_CASE(D)                                                                                               \
  case D:                                                                                                              \
    pdc_lane_row<D>(s_graph + e0, soft, c2v + e0 * Z, Z, tid, it == 0, generic);                                          \
    break;
          PDC_LANE_CASE(3)
          PDC_LANE_CASE(4)
          PDC_LANE_CASE(5)
          PDC_LANE_CASE(6)
          PDC_LANE_CASE(7)
          PDC_LANE_CASE(8)
          PDC_LANE_CASE(9)
          PDC_LANE_CASE(10)
          default:
            pdc_lane_row<19>(s_graph + e0, soft, c2v + e0 * Z, Z, tid, it == 0, generic);
            break;
#undef PDC_LANE_CASE
        }
      }
      __syncthreads();
    }
    it_done = it + 1;
    if (early) {
      // get_hard_bits + CRC over the first K - F bits (:127-135): the CRC of a message is the XOR of x^(n_sig-1-n+order)
      // mod P over its set bits; a zero soft bit anywhere in the K systematic positions refuses the early stop.
      uint32_t acc = 0;
      int      bad = 0;
      for (int n = tid; n < K; n += nth) {
        int s = soft[n];
        bad |= (s == 0);
        if (s <= 0 && n < n_sig) {
          acc ^= crc_pow[n_sig - 1 - n + crc_order];
        }
      }
      bad          = __syncthreads_or(bad);
      uint32_t crc = pdc_block_xor(acc, s_scratch, tid, nth);
      success      = (!bad) && (crc == 0);
    }
  }

  // ---- hard decision, MSB-first packing (bit_buffer.h:106-136, log_likelihood_ratio.cpp:350-376) ----------------------
  {
    uint32_t* out32 = reinterpret_cast<uint32_t*>(cb.out);
    for (int w = tid; w < (K + 31) / 32; w += nth) {
      uint32_t v = 0;
      for (int k = 0; k < 32; ++k) {
        int n = 32 * w + k;
        if (n < K && soft[n] <= 0) {
          v |= (0x80u >> (k & 7)) << (8 * (k >> 3));
        }
      }
      out32[w] = v;
    }
  }
  if (success) {
    res.iterations = it_done;
    res.crc_pass   = 1;
  } else if (!early && cb.crc_type != PDC_CRC_NONE) {
    // pusch_codeblock_decoder.cpp:62-73: no early stop => one CRC after the last iteration, reports max_iter on success.
    uint32_t acc = 0;
    for (int n = tid; n < n_sig; n += nth) {
      if (soft[n] <= 0) {
        acc ^= crc_pow[n_sig - 1 - n + crc_order];
      }
    }
    uint32_t crc = pdc_block_xor(acc, s_scratch, tid, nth);
    if (crc == 0) {
      res.iterations = max_iter;
      res.crc_pass   = 1;
    }
  }
  if (tid == 0) {
    *cb.res = res;
  }
}
