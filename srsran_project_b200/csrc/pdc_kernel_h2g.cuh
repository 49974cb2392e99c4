// Packed half2 code-block kernel for ANY number of layers (4 < L <= 46 | 42): low and medium code rates, retransmissions
// that filled more of the circular buffer, the rate-1/3 case of ldpc_decoder_benchmark. Lifting sizes that are multiples
// of 2. Same arithmetic, lane mapping, lazily rotated doubled rows, CRC epilogue and output as
// pdc_kernel_h2.cuh (whose helpers it reuses); what differs:
//   * the number of layers is known only after the rate dematcher ran (ldpc_decoder_impl.cpp:83-115: it follows from the
//     index of the last non-zero soft bit), so the relative rotations of the columns — which depend on WHICH rows take
//     part — are derived per code block, in shared memory, by one thread per column;
//   * the four core rows keep their messages in registers; the extension rows (degree 3..10, each with one private
//     degree-1 parity column) keep theirs as the same int8 pairs in shared memory, [word][thread] so that every access is
//     conflict free, and their parity column as one private word per thread (never rotated, never shared);
//   * extension rows are a run-time loop over rows with the row degree as a template parameter (8 variants).
#pragma once
#include "pdc_kernel_h2.cuh"

#define PDC_H2G_MAX_EDGES 320
#if defined(PDC_KERNEL_TU)
// Static structure of the base graphs (filled by pdc_h2g_init_device): previous edge on the same column (0xffff: none) and,
// per number of layers L, the last edge on every core column among rows < L (0xffff: none).
__device__ uint16_t d_h2g_prev[2][PDC_H2G_MAX_EDGES];
__device__ uint16_t d_h2g_last[2][48][26];
#endif

template <int DEG>
__device__ __forceinline__ void h2g_ext_row(int e0, uint32_t cadr, uint32_t cstride, uint32_t padr,
                                            const uint32_t* __restrict__ xr, const uint32_t* __restrict__ xw, uint32_t soft_t,
                                            uint32_t half_row, bool active, bool generic)
{
  constexpr int  NW = (DEG + 1) / 2;
  uint32_t       u[DEG], cw[NW];
  uint32_t       par = 0;
  const uint32_t one = h2_opaque_one();
#pragma unroll
  for (int j = 0; j < NW; ++j) {
    cw[j] = lds32(cadr + j * cstride); // private to the thread: no ordering needed
  }
  const uint32_t Sp = lds32(padr);
  uint32_t       ents[DEG - 1];
#pragma unroll
  for (int k = 0; k < DEG - 1; ++k) {
    ents[k] = xr[e0 + k]; // table entries do not depend on the other warps either
  }
  h2_static_for<DEG>([&](auto k_) {
    constexpr int k = decltype(k_)::value;
    uint32_t      S;
    if constexpr (k + 1 < DEG) {
      const uint32_t ent = ents[k];
      const uint32_t adr = soft_t + (ent & 0xffffu);
      const uint32_t w0 = lds32(adr), w1 = lds32(adr + 4);
      S = __byte_perm(w0, w1, ent >> 16);
    } else {
      S = Sp; // the row's own parity column (last edge: columns ascend)
    }
    const uint32_t x = __byte_perm(cw[k / 2], 0x64646464u, (k & 1) ? 0x4342 : 0x4140);
    u[k]             = h2_sub(S, x);
    par ^= u[k];
  });
  uint32_t s1p, d, m1;
  h2_row_minima<DEG>(u, par, generic, m1, s1p, d);
  uint32_t y[DEG];
  h2_static_for<DEG>([&](auto k_) {
    constexpr int k = decltype(k_)::value;
    h2_edge_update(u[k], y[k], m1, s1p, d, one);
  });
  if (active) {
    h2_static_for<NW>([&](auto i_) {
      constexpr int i = decltype(i_)::value;
      if constexpr (2 * i + 1 < DEG) {
        sts32(cadr + i * cstride, __byte_perm(y[2 * i], y[2 * i + 1], 0x6420));
      } else {
        sts32(cadr + i * cstride, __byte_perm(y[2 * i], 0u, 0x4420));
      }
    });
    sts32(padr, u[DEG - 1]);
  }
  __syncthreads(); // every warp has read the row's soft bits
  if (active) {
    h2_static_for<DEG - 1>([&](auto k_) {
      constexpr int  k   = decltype(k_)::value;
      const uint32_t adr = soft_t + xw[e0 + k];
      sts32(adr, u[k]);
      sts32(adr + half_row, u[k]);
    });
  }
  __syncthreads(); // ... and written the new ones
}

// Device-side twin of pdc_h2g_smem_bytes (same formula; the row degrees come from constant memory).
template <int BG>
__device__ __forceinline__ size_t pdc_h2g_smem_bytes_dev(int Z, int max_layers)
{
  constexpr int  ncol = h2_graph<BG>::NCOL, kb = h2_graph<BG>::KB;
  constexpr int  M    = (BG == 1) ? PDC_BG1_ROWS : PDC_BG2_ROWS;
  const uint8_t* deg  = c_bg_rows.deg[BG - 1];
  const int      L    = max_layers > M ? M : max_layers;
  size_t         words = 0;
  for (int m = 4; m < L; ++m) {
    words += (deg[m] + 1) / 2;
  }
  const size_t work  = words * 2 * Z + (size_t)kb * 48;
  const size_t stage = (size_t)(kb + L - 2) * Z + 32;
  return (size_t)ncol * 4 * Z + (size_t)(L - 4) * 2 * Z + (work > stage ? work : stage);
}

template <int BG>
__global__ void __launch_bounds__(192, 1) pdc_cb_kernel_h2g(pdc_launch_params prm)
{
  using G = h2_graph<BG>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint32_t s_graph[PDC_H2G_MAX_EDGES]; // col | shift << 8 of every edge, rows concatenated
  __shared__ uint32_t s_tab[PDC_H2_TAB_STRIDE];   // core edges, column alignments, natural views: layout of c_h2_tab
  __shared__ uint32_t s_xr[PDC_H2G_MAX_EDGES];    // extension edges: read offset | PRMT selector << 16
  __shared__ uint32_t s_xw[PDC_H2G_MAX_EDGES];    // extension edges: byte offset of the column
  __shared__ uint16_t s_wstart[48];               // first message word of every extension row
  __shared__ uint32_t s_red[8];
  __shared__ int      s_nz;

  const int        tid = threadIdx.x, nth = blockDim.x;
  const int        cbi = prm.order ? prm.order[blockIdx.x] : blockIdx.x;
  const pdc_dev_cb cb  = prm.cbs[cbi];
  if (cb.bg != BG) {
    return;
  }
  pdc_geom g;
  pdc_geom_init(g, cb.bg, cb.Z);
  const int  Z = g.Z, W = Z >> 1, K = G::KB * Z, bgi = BG - 1;
  const bool active  = tid < W;
  const bool generic = (cb.flags & PDC_F_GENERIC) != 0;

  pdc_cb_result res;
  res.iterations = -1, res.crc_pass = 0, res.status = 0, res.nof_layers = 0;

  const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(smem_raw);
  const uint32_t row_bytes = 4u * (uint32_t)Z;
  const uint32_t soft_t    = smem_base + 4u * (uint32_t)(active ? tid : 0);
  if (tid == 0) {
    s_nz = 0;
  }
  const uint16_t* row_start = c_bg_rows.start[bgi];
  const uint8_t*  row_deg   = c_bg_rows.deg[bgi];
  const int       M_rows    = (BG == 1) ? PDC_BG1_ROWS : PDC_BG2_ROWS;
  for (int e = tid; e < row_start[M_rows]; e += nth) {
    s_graph[e] = prm.graphs[(size_t)cb.graph_idx * PDC_GRAPH_STRIDE + e];
  }

  // ---- prologue 1: rate dematching + HARQ combining into the soft buffer in HBM ------------------------------------------------
  const int8_t* dec_in;
  int           n_in;
  if (cb.flags & PDC_F_DECODE_ONLY) {
    dec_in = cb.llrs;
    n_in   = (int)cb.E;
  } else {
    pdc_dematch_plan pl;
    pdc_dematch_plan_init(pl, g, (int)cb.E, cb.Qm, cb.rv, (int)cb.Nref, (int)cb.nof_filler,
                          (cb.flags & PDC_F_NEW_DATA) != 0);
    if (pl.new_data && pl.rounds == 1) {
      pdc_dematch_first_tx(pl, cb.llrs, cb.soft, tid, nth);
    } else {
      pdc_dematch(pl, cb.llrs, cb.soft, generic, tid, nth);
    }
    dec_in = cb.soft;
    n_in   = g.N;
  }
  __syncthreads(); // orders the dematcher's writes before the reads below (same CTA); s_graph, s_nz, barriers visible

  // ---- the decoder input, staged at the tail of the dynamic shared memory (the host sized it: pdc_h2g_smem_bytes). Only the
  //      part the host's bound on the number of layers allows to be non-zero is needed. ----------------------------------------
  const int      n_stage  = min(n_in, (G::KB + (int)cb.max_layers - 2) * Z);
  const size_t   dyn_size = pdc_h2g_smem_bytes_dev<BG>(Z, (int)cb.max_layers);
  unsigned char* stage    = smem_raw + ((dyn_size - (size_t)((n_stage + 15) & ~15)) & ~(size_t)15);
  {
    const bool vec = ((reinterpret_cast<uintptr_t>(dec_in) & 15) == 0);
    const int  n16 = vec ? (n_stage >> 4) : 0;
    int        best = 0;
    for (int q = tid; q < n16; q += nth) {
      const int4 v = *reinterpret_cast<const int4*>(dec_in + 16 * q);
      reinterpret_cast<int4*>(stage)[q] = v;
      const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (w[j] != 0) {
          best = 16 * q + 4 * j + 4 - (__clz(w[j]) >> 3); // 1 + index of the last non-zero byte of the word
        }
      }
    }
    for (int n = 16 * n16 + tid; n < n_stage; n += nth) {
      const int8_t x = dec_in[n];
      stage[n]       = (unsigned char)x;
      if (x != 0) {
        best = n + 1;
      }
    }
    best = __reduce_max_sync(0xffffffffu, best);
    if ((tid & 31) == 0 && best > 0) {
      atomicMax(&s_nz, best);
    }
  }
  __syncthreads();
  const int L = pdc_layers_for_nz(g, s_nz);
  if (L > (int)cb.max_layers) { // cannot happen unless the host bound is wrong; fail loudly instead of overrunning
    res.status = 3;
    if (tid == 0) {
      *cb.res = res;
    }
    return;
  }
  res.nof_layers = (uint16_t)L;
  const int E_L  = row_start[L];

  // ---- relative rotations for THIS set of rows. The previous writer of a column is the nearest earlier edge on the same
  //      column (static: d_h2g_prev) or, for the first edge of a column, the last edge on it among the L rows that take part
  //      (d_h2g_last[L]): one thread per edge. -------------------------------------------------------------------------------
  for (int e = tid; e < E_L; e += nth) {
    const uint32_t ge = s_graph[e];
    const int      c  = (int)(ge & 0xffu);
    if (c < G::NCOL) {
      int pe = d_h2g_prev[bgi][e];
      pe     = (pe == 0xffff) ? (int)d_h2g_last[bgi][L][c] : pe;
      int rel = (int)(ge >> 8) - (int)(s_graph[pe] >> 8);
      rel += (rel < 0) ? Z : 0;
      const uint32_t off = (uint32_t)c * row_bytes + 4u * (uint32_t)(rel >> 1);
      const uint32_t val = off | (((rel & 1) ? 0x5432u : 0x3210u) << 16);
      if (e < G::NEDGE) {
        s_tab[e] = val;
      } else {
        s_xr[e] = val;
        s_xw[e] = (uint32_t)c * row_bytes;
      }
    }
  }
  if (tid < G::NCOL) {
    const int c     = tid;
    const int le    = d_h2g_last[bgi][L][c];
    const int sigma = (le == 0xffff) ? 0 : (int)(s_graph[le] >> 8); // alignment after a full iteration
    s_tab[PDC_H2_SIGMA_OFF + c] = (uint32_t)sigma;
    const int nat               = (Z - sigma) % Z;
    s_tab[PDC_H2_NAT_OFF + c]   = ((uint32_t)c * row_bytes + 4u * (uint32_t)(nat >> 1)) | (((nat & 1) ? 0x5432u : 0x3210u) << 16);
  }
  if (tid == nth - 1) { // a thread without a column to walk (nth >= 32 > NCOL)
    int acc = 0;
    for (int m = 4; m < L; ++m) {
      s_wstart[m] = (uint16_t)acc;
      acc += (row_deg[m] + 1) >> 1;
    }
    s_wstart[L] = (uint16_t)acc;
  }
  __syncthreads();

  // Shared-memory map: core columns (doubled rows) | parity columns of the extension rows (one word per thread) |
  // messages of the extension rows ([word][thread]) | ballot words of the output.
  const uint32_t ext_base = smem_base + G::NCOL * row_bytes;
  const uint32_t ext_row  = 4u * (uint32_t)W; // bytes per private word-row
  const uint32_t c2v_base = ext_base + (uint32_t)(L - 4) * ext_row;
  const uint32_t priv_t   = 4u * (uint32_t)(active ? tid : 0);
  unsigned char* aux      = smem_raw + G::NCOL * row_bytes + (size_t)(L - 4) * ext_row + (size_t)s_wstart[L] * ext_row;

  // ---- prologue 2: load_soft_bits (ldpc_decoder_impl.cpp:150-191) -------------------------------------------------------------
  if (active) {
    sts32(soft_t, H2_BIAS), sts32(soft_t + 2u * Z, H2_BIAS); // the two punctured columns: zero
    sts32(soft_t + row_bytes, H2_BIAS), sts32(soft_t + row_bytes + 2u * Z, H2_BIAS);
    for (int c = 2; c < G::KB + L; ++c) {
      const int sigma = (c < G::NCOL) ? (int)s_tab[PDC_H2_SIGMA_OFF + c] : 0;
      int       la    = 2 * tid + sigma;
      la -= (la >= Z) ? Z : 0;
      int lb = la + 1;
      lb -= (lb >= Z) ? Z : 0;
      const int      pa = (c - 2) * Z + la, pb = (c - 2) * Z + lb;
      const uint32_t xa = (pa < n_stage) ? stage[pa] : 0u;
      const uint32_t xb = (pb < n_stage) ? stage[pb] : 0u;
      const uint32_t S  = h2_soft_from_bytes(xa | (xb << 8));
      if (c < G::NCOL) {
        sts32(soft_t + c * row_bytes, S);
        sts32(soft_t + c * row_bytes + 2u * Z, S);
      } else {
        sts32(ext_base + (uint32_t)(c - G::NCOL) * ext_row + priv_t, S);
      }
    }
  }
  __syncthreads(); // the messages below may overlap the staged input
  if (active) {
    const int nwords = s_wstart[L];
    for (int w = 0; w < nwords; ++w) {
      sts32(c2v_base + (uint32_t)w * ext_row + priv_t, 0x80808080u); // all messages zero before the first iteration
    }
  }
  __syncthreads();

  const int  n_sig     = K - (int)cb.nof_filler;
  const bool has_crc   = cb.crc_type != PDC_CRC_NONE;
  const bool early     = (cb.flags & PDC_F_EARLY_STOP) && has_crc;
  const int  max_iter  = cb.max_iter;
  const int  crc_order = (cb.crc_type == PDC_CRC16) ? 16 : 24;
  const uint2* crc_pair = nullptr; // see pdc_kernel_h2.cuh
  if (has_crc) {
    const int X = PDC_CRC_BIT_PAD + n_sig - 1 + crc_order;
    const int v = (X - 1) & 1;
    crc_pair    = reinterpret_cast<const uint2*>(prm.crc_bit + (size_t)(2 * (cb.crc_type - 1) + v) * PDC_CRC_BIT_STRIDE +
                                              (v + X - 1 - 2 * (active ? tid : 0)));
  }

  uint32_t cp[h2_pack<BG>::NREG];
#pragma unroll
  for (int e = 0; e < h2_pack<BG>::NREG; ++e) {
    cp[e] = 0x80808080u;
  }

  int  it_done = 0;
  bool success = false;
  for (int it = 0; it < max_iter && !success; ++it) {
    h2_row<BG, 0, 0>(cp, s_tab, soft_t, row_bytes, active, generic);
    h2_row<BG, 0, 1>(cp, s_tab, soft_t, row_bytes, active, generic);
    h2_row<BG, 0, 2>(cp, s_tab, soft_t, row_bytes, active, generic);
    h2_row<BG, 0, 3>(cp, s_tab, soft_t, row_bytes, active, generic);
    for (int m = 4; m < L; ++m) {
      const int      e0   = row_start[m];
      const uint32_t cadr = c2v_base + (uint32_t)s_wstart[m] * ext_row + priv_t;
      const uint32_t padr = ext_base + (uint32_t)(m - 4) * ext_row + priv_t;
#define PDC_H2G_CASE(D)                                                                                                \
  case D:                                                                                                              \
    h2g_ext_row<D>(e0, cadr, ext_row, padr, s_xr, s_xw, soft_t, 2u * (uint32_t)Z, active, generic);                       \
    break;
      switch (row_deg[m]) {
        PDC_H2G_CASE(3)
        PDC_H2G_CASE(4)
        PDC_H2G_CASE(5)
        PDC_H2G_CASE(6)
        PDC_H2G_CASE(7)
        PDC_H2G_CASE(8)
        PDC_H2G_CASE(9)
        default:
          h2g_ext_row<10>(e0, cadr, ext_row, padr, s_xr, s_xw, soft_t, 2u * (uint32_t)Z, active, generic);
          break;
      }
#undef PDC_H2G_CASE
    }
    it_done = it + 1;
    const bool last = (it_done == max_iter);
    if (!has_crc || (!early && !last)) {
      continue;
    }
    // ---- hard decision + CRC, as in pdc_kernel_h2.cuh -------------------------------------------------------------------------
    uint32_t amin = H2_ONE, acc = 0;
    if (active) {
      h2_static_for<G::KB>([&](auto c_) {
        constexpr int  c = decltype(c_)::value;
        const uint32_t S = h2_read_natural<BG, 0, c>(s_tab, soft_t);
        amin             = h2_min(amin, h2_abs(h2_sub(S, H2_BIAS)));
        const uint2 tp   = __ldg(crc_pair - c * (Z >> 1));
        asm("{ .reg .pred p, q;\n"
            "setp.le.f16x2 p|q, %1, %2;\n"
            "@p xor.b32 %0, %0, %3;\n"
            "@q xor.b32 %0, %0, %4; }"
            : "+r"(acc)
            : "r"(S), "r"(H2_BIAS), "r"(tp.y), "r"(tp.x));
      });
    }
    acc = __reduce_xor_sync(0xffffffffu, acc);
    if ((tid & 31) == 0) {
      s_red[tid >> 5] = acc;
    }
    const int bad = __syncthreads_or(((amin & 0x7fffu) == 0) || ((amin & 0x7fff0000u) == 0));
    uint32_t  crc = 0;
    for (int w = 0; w < (nth >> 5); ++w) {
      crc ^= s_red[w];
    }
    success = (crc == 0u) && (early ? !bad : true);
  }

  // ---- output (as in pdc_kernel_h2.cuh) -----------------------------------------------------------------------------------------
  __syncthreads();
  h2_emit_bits<BG, 0>(s_tab, soft_t, reinterpret_cast<uint32_t*>(aux), active, tid, nth, Z, K, reinterpret_cast<uint32_t*>(cb.out));
  if (success) {
    res.iterations = early ? it_done : max_iter;
    res.crc_pass   = 1;
  }
  if (tid == 0) {
    *cb.res = res;
  }
}

// Dynamic shared memory of one code block with at most max_layers layers: core columns | parity columns of the extension
// rows | max(messages + ballot words, staged decoder input).
extern "C" size_t pdc_h2g_smem_bytes(int bg, int Z, int max_layers)
{
  const int      ncol = (bg == 1) ? 26 : 14, kb = (bg == 1) ? 22 : 10;
  const int      M    = (bg == 1) ? PDC_BG1_ROWS : PDC_BG2_ROWS;
  const uint8_t* deg  = (bg == 1) ? pdc_bg1_row_deg : pdc_bg2_row_deg;
  const int      L    = max_layers > M ? M : max_layers;
  size_t         words = 0;
  for (int m = 4; m < L; ++m) {
    words += (deg[m] + 1) / 2;
  }
  const size_t work  = words * 2 * Z + (size_t)kb * 48;
  const size_t stage = (size_t)(kb + L - 2) * Z + 32;
  return (size_t)ncol * 4 * Z + (size_t)(L - 4) * 2 * Z + (work > stage ? work : stage);
}
static cudaError_t pdc_h2g_init_device()
{
  static uint16_t prev[2][PDC_H2G_MAX_EDGES];
  static uint16_t last[2][48][26];
  for (int bg = 0; bg != 2; ++bg) {
    const uint8_t* deg = bg ? pdc_bg2_row_deg : pdc_bg1_row_deg;
    const uint8_t* col = bg ? pdc_bg2_edge_col : pdc_bg1_edge_col;
    const int      M = bg ? PDC_BG2_ROWS : PDC_BG1_ROWS, ncol = bg ? 14 : 26;
    int            cur[26];
    for (int c = 0; c != 26; ++c) {
      cur[c] = 0xffff;
    }
    int e = 0;
    for (int m = 0; m <= M; ++m) {
      for (int c = 0; c != 26; ++c) {
        last[bg][m][c] = (uint16_t)cur[c]; // rows < m
      }
      if (m == M) {
        break;
      }
      for (int k = 0; k != deg[m]; ++k, ++e) {
        if (col[e] < ncol) {
          prev[bg][e] = (uint16_t)cur[col[e]];
          cur[col[e]] = e;
        } else {
          prev[bg][e] = 0xffff;
        }
      }
    }
  }
  cudaError_t err = cudaMemcpyToSymbol(d_h2g_prev, prev, sizeof(prev));
  if (err != cudaSuccess) {
    return err;
  }
  return cudaMemcpyToSymbol(d_h2g_last, last, sizeof(last));
}
extern "C" cudaError_t pdc_launch_h2g(const pdc_launch_params& prm, int bg, int block, size_t smem, cudaStream_t stream)
{
  if (bg == 1) {
    pdc_cb_kernel_h2g<1><<<prm.n, block, smem, stream>>>(prm);
  } else {
    pdc_cb_kernel_h2g<2><<<prm.n, block, smem, stream>>>(prm);
  }
  return cudaGetLastError();
}
