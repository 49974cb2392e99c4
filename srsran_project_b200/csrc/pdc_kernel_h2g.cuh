// Packed half2 code-block kernel for ANY number of layers (4 < L <= 46 | 42): low and medium code rates, retransmissions
// that filled more of the circular buffer, the rate-1/3 case of ldpc_decoder_benchmark. Lifting sizes that are multiples
// of 2. Same arithmetic, lane mapping, lazily rotated doubled rows, CRC epilogue and output as
// pdc_kernel_h2.cuh (whose helpers it reuses); what differs:
//   * the number of layers is known only after the rate dematcher ran (ldpc_decoder_impl.cpp:83-115: it follows from the
//     index of the last non-zero soft bit), so the relative rotations of the columns — which depend on WHICH rows take
//     part — are derived per code block, in shared memory, by one thread per edge;
//   * the four core rows keep their messages in registers; the extension rows (degree 3..10, each with one private
//     degree-1 parity column) keep theirs, the same int8 pairs, in TENSOR MEMORY: a thread's messages are only ever
//     touched by that thread, and the 32x32b shape of tcgen05.ld / tcgen05.st gives every thread of a warp its own TMEM
//     lane, so the 256 KB of TMEM per SM serve as a private scratch that costs neither registers nor shared memory. At
//     rate 1/3 and Z = 384 a code block's 107 message words per thread would otherwise need 80 KB of shared memory: with
//     them in TMEM two code blocks are resident per SM instead of one (the kernel is latency-bound: profiles/);
//   * the degree-1 parity column of an extension row is read by that row alone, so instead of its soft value S and the
//     row's message c to it only u = S - c is kept (one private shared-memory word per thread): u never changes until the
//     update promotes S to +-infinity, and then stays infinite (pdc_kernel_h2.cuh: h2_edge_update; here h2g_parity_update);
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

// ---- tensor memory as per-thread scratch -----------------------------------------------------------------------------------
// 32x32b: thread i of warp w addresses TMEM lane 32 (w % 4) + i; the address (lane << 16 | column) is warp-uniform.
template <int N, bool WAIT = true>
__device__ __forceinline__ void h2g_tm_ld(uint32_t taddr, uint32_t (&v)[N])
{
  static_assert(N >= 1 && N <= 5, "message words of one extension row");
  if constexpr (N == 1) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v[0]) : "r"(taddr));
  } else if constexpr (N == 2 || N == 3) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(v[0]), "=r"(v[1]) : "r"(taddr));
    if constexpr (N == 3) {
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v[2]) : "r"(taddr + 2));
    }
  } else {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
                 : "r"(taddr));
    if constexpr (N == 5) {
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v[4]) : "r"(taddr + 4));
    }
  }
  if constexpr (WAIT) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  }
}
template <int N>
__device__ __forceinline__ void h2g_tm_st(uint32_t taddr, const uint32_t (&v)[N])
{
  if constexpr (N == 1) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr), "r"(v[0]));
  } else if constexpr (N == 2 || N == 3) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(taddr), "r"(v[0]), "r"(v[1]));
    if constexpr (N == 3) {
      asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr + 2), "r"(v[2]));
    }
  } else {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(v[0]), "r"(v[1]),
                 "r"(v[2]), "r"(v[3]));
    if constexpr (N == 5) {
      asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr + 4), "r"(v[4]));
    }
  }
}

// Pass 2 for the row's own degree-1 parity column: only u = S - c is kept. Not promoted: S' = clamp(u) + c', so
// u' = S' - c' = clamp(u, +-120) = u (|u| <= 64 until it is infinite). Promoted (u infinite, or min(|u|, 120) + ms > 120):
// S' = sign(u) inf, and so is u' from then on.
__device__ __forceinline__ void h2g_parity_update(uint32_t& u, uint32_t m1, uint32_t s1p, uint32_t d, uint32_t one)
{
  const uint32_t a  = h2_abs(u);
  const uint32_t eq = h2_eq_mask(a, m1);
  const uint32_t ms = s1p ^ (d & eq);
  const uint32_t sg = h2_copysign_one(u, one);
  const uint32_t w  = h2_add(h2_min(a, H2_120), ms);
  const uint32_t qa = h2_add_sat(a, H2_N336);
  const uint32_t qb = h2_fma_sat(w, H2_ONE, H2_N120);
  const uint32_t G  = h2_fma(qa, H2_65504, h2_mul(qb, H2_65504)); // 0 | 65504 | inf
  u                 = h2_fma(sg, h2_add(G, G), u);                 // sg * (0 | inf) + u
}

// One extension row. taddr: TMEM address of the row's message words of this warp; padr: shared-memory address of the
// thread's parity word (u of the degree-1 column); xe: shared-memory address of the row's first edge entry
// {read offset, PRMT selector, column offset, -} (16 bytes per edge: one load, no unpacking).
__device__ __forceinline__ uint4 lds128(uint32_t addr)
{
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t h2g_prmt(uint32_t a, uint32_t b, uint32_t sel)
{
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel)); // sel is a clean 4-nibble selector: no masking
  return d;
}
// The work of one extension row in four pieces, so that two rows that share no column can go through them side by side
// (h2g_ext_row2): what the row reads (its own words first: nobody else touches them), the arithmetic, what only this thread ever
// sees again (messages, parity word), and the soft bits the other threads read in later rows.
template <int DEG>
struct h2g_ext_regs {
  uint32_t u[DEG];        // pass 1: v2c messages; after pass 2: the new soft bits (the last entry: the parity column's u)
  uint32_t cw[DEG / 2];   // message words (DEG - 1 messages, two per word)
  uint32_t wadr[DEG - 1]; // where the new soft bits go
  uint32_t w0[DEG - 1], w1[DEG - 1], sel[DEG - 1];
};
template <int DEG, bool WAIT_LD>
__device__ __forceinline__ void h2g_ext_load(h2g_ext_regs<DEG>& r, uint32_t xe, uint32_t taddr, uint32_t padr, uint32_t soft_t)
{
  h2g_tm_ld<DEG / 2, false>(taddr, r.cw); // private to the thread: no ordering needed; on its way while shared memory is read
  uint4 ents[DEG - 1];
#pragma unroll
  for (int k = 0; k < DEG - 1; ++k) {
    ents[k] = lds128(xe + 16u * k); // table entries do not depend on the other warps
  }
  r.u[DEG - 1] = lds32(padr); // private as well
#pragma unroll
  for (int k = 0; k < DEG - 1; ++k) {
    const uint32_t adr = soft_t + ents[k].x;
    r.w0[k] = lds32(adr), r.w1[k] = lds32(adr + 4);
    r.sel[k]  = ents[k].y;
    r.wadr[k] = soft_t + ents[k].z;
  }
  if constexpr (WAIT_LD) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  }
}
template <int DEG>
__device__ __forceinline__ void h2g_ext_compute(h2g_ext_regs<DEG>& r, bool generic, uint32_t one)
{
  uint32_t par = 0;
  h2_static_for<DEG - 1>([&](auto k_) {
    constexpr int  k = decltype(k_)::value;
    const uint32_t S = h2g_prmt(r.w0[k], r.w1[k], r.sel[k]);
    const uint32_t x = __byte_perm(r.cw[k / 2], 0x64646464u, (k & 1) ? 0x4342 : 0x4140);
    r.u[k]           = h2_sub(S, x);
    par ^= r.u[k];
  });
  par ^= r.u[DEG - 1];
  uint32_t s1p, d, m1;
  h2_row_minima<DEG>(r.u, par, generic, m1, s1p, d);
  uint32_t y[DEG - 1];
  h2_static_for<DEG - 1>([&](auto k_) {
    constexpr int k = decltype(k_)::value;
    h2_edge_update(r.u[k], y[k], m1, s1p, d, one);
  });
  h2g_parity_update(r.u[DEG - 1], m1, s1p, d, one);
  h2_static_for<DEG / 2>([&](auto i_) {
    constexpr int i = decltype(i_)::value;
    if constexpr (2 * i + 1 < DEG - 1) {
      r.cw[i] = __byte_perm(y[2 * i], y[2 * i + 1], 0x6420);
    } else {
      r.cw[i] = __byte_perm(y[2 * i], 0u, 0x4420);
    }
  });
}
template <int DEG>
__device__ __forceinline__ void h2g_ext_store_private(const h2g_ext_regs<DEG>& r, uint32_t taddr, uint32_t padr, bool active)
{
  h2g_tm_st<DEG / 2>(taddr, r.cw); // idle lanes of a partly filled warp write their own (unused) TMEM lanes
  if (active) {
    sts32(padr, r.u[DEG - 1]);
  }
}
template <int DEG>
__device__ __forceinline__ void h2g_ext_store_shared(const h2g_ext_regs<DEG>& r, uint32_t half_row, bool active)
{
  if (active) {
    h2_static_for<DEG - 1>([&](auto k_) {
      constexpr int k = decltype(k_)::value;
      sts32(r.wadr[k], r.u[k]);
      sts32(r.wadr[k] + half_row, r.u[k]);
    });
  }
}
template <int DEG>
__device__ __forceinline__ void h2g_ext_row(uint32_t xe, uint32_t taddr, uint32_t padr, uint32_t soft_t, uint32_t half_row,
                                            bool active, bool generic)
{
  h2g_ext_regs<DEG> r;
  h2g_ext_load<DEG, true>(r, xe, taddr, padr, soft_t);
  h2g_ext_compute<DEG>(r, generic, h2_opaque_one());
  h2g_ext_store_private<DEG>(r, taddr, padr, active);
  __syncthreads(); // every warp has read the row's soft bits
  h2g_ext_store_shared<DEG>(r, half_row, active);
  __syncthreads(); // ... and written the new ones
}
// Two consecutive extension rows that share no column (h2g_pair_mask): the layered schedule gives the same result whichever of
// the two runs first, so they run together — one pair of barriers instead of two, and two independent dependency chains per
// thread in a kernel that is latency-bound (two code blocks, twelve warps per SM).
template <int DA, int DB>
__device__ __forceinline__ void h2g_ext_row2(uint32_t xe_a, uint32_t taddr_a, uint32_t padr_a, uint32_t xe_b, uint32_t taddr_b,
                                             uint32_t padr_b, uint32_t soft_t, uint32_t half_row, bool active, bool generic)
{
  h2g_ext_regs<DA> ra;
  h2g_ext_regs<DB> rb;
  h2g_ext_load<DA, false>(ra, xe_a, taddr_a, padr_a, soft_t);
  h2g_ext_load<DB, true>(rb, xe_b, taddr_b, padr_b, soft_t);
  const uint32_t one = h2_opaque_one();
  h2g_ext_compute<DA>(ra, generic, one);
  h2g_ext_compute<DB>(rb, generic, one);
  h2g_ext_store_private<DA>(ra, taddr_a, padr_a, active);
  h2g_ext_store_private<DB>(rb, taddr_b, padr_b, active);
  __syncthreads();
  h2g_ext_store_shared<DA>(ra, half_row, active);
  h2g_ext_store_shared<DB>(rb, half_row, active);
  __syncthreads();
}
// Bit m set: extension rows m and m + 1 share no column of the core part (their own parity columns are private anyway) and form
// a pair; pairs are formed greedily from row 4 on. BG1: (16,17) and every row from 20 on; BG2: (11,12), (17,18) and from 20 on.
template <int BG>
__host__ __device__ constexpr uint64_t h2g_pair_mask()
{
  constexpr int M     = (BG == 1) ? PDC_BG1_ROWS : PDC_BG2_ROWS;
  constexpr int ncore = (BG == 1) ? 26 : 14;
  uint64_t      mask  = 0;
  int           start[48] = {};
  int           acc       = 0;
  for (int m = 0; m != M; ++m) {
    start[m] = acc;
    acc += (BG == 1) ? pdc_bg1_row_deg[m] : pdc_bg2_row_deg[m];
  }
  start[M] = acc;
  int m    = 4;
  while (m + 1 < M) {
    bool disjoint = true;
    for (int e = start[m]; e != start[m + 1]; ++e) {
      for (int f = start[m + 1]; f != start[m + 2]; ++f) {
        const int ce = (BG == 1) ? pdc_bg1_edge_col[e] : pdc_bg2_edge_col[e];
        const int cf = (BG == 1) ? pdc_bg1_edge_col[f] : pdc_bg2_edge_col[f];
        disjoint     = disjoint && !(ce == cf && ce < ncore);
      }
    }
    if (disjoint) {
      mask |= 1ull << m;
      m += 2;
    } else {
      m += 1;
    }
  }
  return mask;
}
static_assert(h2g_pair_mask<1>() == 0x155555510000ull, "BG1: (16,17), (20,21), (22,23), ... (44,45)");
static_assert(h2g_pair_mask<2>() == 0x15555520800ull, "BG2: (11,12), (17,18), (20,21), (22,23), ... (40,41)");

// Message words (int8 pairs, two edges per word) of the extension rows 4 .. L - 1: the parity edge of a row carries none.
template <int BG>
__host__ __device__ inline int pdc_h2g_ext_words(const uint8_t* deg, int L)
{
  int words = 0;
  for (int m = 4; m < L; ++m) {
    words += deg[m] / 2;
  }
  return words;
}
// TMEM columns of one code block: every warp owns the TMEM lane quarter warp % 4, so warps 4, 5 (Z > 256) take a second
// bank of columns. Allocations are powers of two >= 32.
__host__ __device__ inline uint32_t pdc_h2g_tmem_cols(int words, int nth)
{
  const uint32_t need = (uint32_t)words * (uint32_t)((nth + 127) / 128);
  uint32_t       c    = 32;
  while (c < need) {
    c <<= 1;
  }
  return c;
}

template <int BG>
__global__ void __launch_bounds__(192, 2) pdc_cb_kernel_h2g(pdc_launch_params prm)
{
  using G = h2_graph<BG>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint32_t s_graph[PDC_H2G_MAX_EDGES]; // col | shift << 8 of every edge, rows concatenated
  __shared__ uint32_t s_tab[PDC_H2_TAB_STRIDE];   // core edges, column alignments, natural views: layout of c_h2_tab
  __shared__ __align__(16) uint4 s_xe[PDC_H2G_MAX_EDGES - 76 + 8]; // extension edges, from edge NEDGE on: {read offset, PRMT
                                                                   // selector, byte offset of the column, -}
  __shared__ __align__(16) uint4 s_row[48]; // extension rows: {address of the row's first s_xe entry, first message word,
                                            // degree, -}; entry L: degree 0
  __shared__ uint32_t s_red[8];
  __shared__ int      s_nz;
  __shared__ uint32_t s_tmem; // base address of this code block's tensor-memory allocation

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

  // Tensor memory for the messages of the extension rows, sized by the host's bound on the number of layers. Warp 0 allocates
  // (and frees at the end); the allocation of a CTA that finds the SM's TMEM taken waits for a resident CTA to finish.
  const int      L_cap     = min((int)cb.max_layers, (BG == 1) ? PDC_BG1_ROWS : PDC_BG2_ROWS);
  const int      tm_words  = pdc_h2g_ext_words<BG>(c_bg_rows.deg[bgi], L_cap);
  const uint32_t tm_cols   = pdc_h2g_tmem_cols(tm_words, nth);
  if (tid < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(&s_tmem)),
                 "r"(tm_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");

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
  __syncthreads(); // orders the dematcher's writes before the reads below (same CTA); s_graph, s_nz, s_tmem visible
  asm volatile("tcgen05.fence::after_thread_sync;");
  // This warp's TMEM lane quarter; warps 4, 5 use a second bank of columns.
  const uint32_t tm_warp = s_tmem + ((uint32_t)((tid >> 5) & 3) << 21) + (uint32_t)(tid >> 7) * (uint32_t)tm_words;

  // ---- the decoder input, staged at the tail of the dynamic shared memory (the host sized it: pdc_h2g_smem_bytes). Only the
  //      part the host's bound on the number of layers allows to be non-zero is needed. ----------------------------------------
  // Shared-memory map: core columns (doubled rows) | parity words of the extension rows (one word per thread and row, sized
  // by the host's bound L_cap) | staged decoder input, later the ballot words of the output.
  const int      n_stage  = min(n_in, (G::KB + L_cap - 2) * Z);
  unsigned char* stage    = smem_raw + (((size_t)G::NCOL * 4 * Z + (size_t)(L_cap - 4) * 2 * Z + 15) & ~(size_t)15);
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
  if (L > L_cap) { // cannot happen unless the host bound is wrong; fail loudly instead of overrunning
    res.status = 3;
    if (tid == 0) {
      *cb.res = res;
    }
    if (tid < 32) {
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(s_tmem), "r"(tm_cols));
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
        s_xe[e - G::NEDGE] = make_uint4(off, (rel & 1) ? 0x5432u : 0x3210u, (uint32_t)c * row_bytes, 0u);
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
    const uint32_t xe0 = (uint32_t)__cvta_generic_to_shared(s_xe);
    int            acc = 0;
    for (int m = 4; m < L; ++m) {
      const uint32_t paired = (m + 1 < L) ? (uint32_t)((h2g_pair_mask<BG>() >> m) & 1ull) : 0u; // with row m + 1
      s_row[m] = make_uint4(xe0 + 16u * (uint32_t)(row_start[m] - G::NEDGE), (uint32_t)acc, row_deg[m], paired);
      acc += row_deg[m] >> 1; // the parity edge carries no message
    }
    s_row[L] = make_uint4(0u, (uint32_t)acc, 0u, 0u);
  }
  __syncthreads();

  const uint32_t ext_base = smem_base + G::NCOL * row_bytes;
  const uint32_t ext_row  = 4u * (uint32_t)W; // bytes per private word-row
  const uint32_t priv_t   = 4u * (uint32_t)(active ? tid : 0);
  unsigned char* aux      = stage; // free once the soft bits have been loaded

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
        // degree-1 parity column: u = S - c with c = 0 before the first iteration, unbiased
        sts32(ext_base + (uint32_t)(c - G::NCOL) * ext_row + priv_t, h2_sub(S, H2_BIAS));
      }
    }
  }
  {
    const int nwords = (int)s_row[L].y;
    for (int w = 0; w < nwords; ++w) { // all messages zero (byte 128) before the first iteration
      asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(tm_warp + (uint32_t)w), "r"(0x80808080u));
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  __syncthreads(); // the ballot words of the output overlap the staged input

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
    // the previous iteration's message stores have landed before this one's loads of the same words
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    h2_row<BG, 0, 0>(cp, s_tab, soft_t, row_bytes, active, generic);
    h2_row<BG, 0, 1>(cp, s_tab, soft_t, row_bytes, active, generic);
    h2_row<BG, 0, 2>(cp, s_tab, soft_t, row_bytes, active, generic);
    h2_row<BG, 0, 3>(cp, s_tab, soft_t, row_bytes, active, generic);
    uint4 hdr = s_row[4];
    for (int m = 4; m < L;) {
      const uint4    nxt  = s_row[m + 1]; // the next rows' headers arrive while this row computes
      const uint4    nxt2 = s_row[min(m + 2, 47)];
      const uint32_t tadr = tm_warp + hdr.y;
      const uint32_t padr = ext_base + (uint32_t)(m - 4) * ext_row + priv_t;
      const uint32_t hr   = 2u * (uint32_t)Z;
      if (hdr.w != 0) {
        // rows m and m + 1 together (h2g_ext_row2); the larger degree first, the order within a pair does not matter
        const uint32_t tadr2 = tm_warp + nxt.y, padr2 = padr + ext_row;
        const uint32_t key   = hdr.z * 16u + nxt.z;
        bool           handled = true;
        if (key == 0x55) {
          h2g_ext_row2<5, 5>(hdr.x, tadr, padr, nxt.x, tadr2, padr2, soft_t, hr, active, generic);
        } else if (key == 0x44) {
          h2g_ext_row2<4, 4>(hdr.x, tadr, padr, nxt.x, tadr2, padr2, soft_t, hr, active, generic);
        } else if (key == 0x54) {
          h2g_ext_row2<5, 4>(hdr.x, tadr, padr, nxt.x, tadr2, padr2, soft_t, hr, active, generic);
        } else if (key == 0x45) {
          h2g_ext_row2<5, 4>(nxt.x, tadr2, padr2, hdr.x, tadr, padr, soft_t, hr, active, generic);
        } else if (key == 0x43) {
          h2g_ext_row2<4, 3>(hdr.x, tadr, padr, nxt.x, tadr2, padr2, soft_t, hr, active, generic);
        } else if (key == 0x34) {
          h2g_ext_row2<4, 3>(nxt.x, tadr2, padr2, hdr.x, tadr, padr, soft_t, hr, active, generic);
        } else if (key == 0x66) {
          h2g_ext_row2<6, 6>(hdr.x, tadr, padr, nxt.x, tadr2, padr2, soft_t, hr, active, generic);
        } else if (key == 0x65) {
          h2g_ext_row2<6, 5>(hdr.x, tadr, padr, nxt.x, tadr2, padr2, soft_t, hr, active, generic);
        } else if (key == 0x56) {
          h2g_ext_row2<6, 5>(nxt.x, tadr2, padr2, hdr.x, tadr, padr, soft_t, hr, active, generic);
        } else if (key == 0x53) {
          h2g_ext_row2<5, 3>(hdr.x, tadr, padr, nxt.x, tadr2, padr2, soft_t, hr, active, generic);
        } else if (key == 0x35) {
          h2g_ext_row2<5, 3>(nxt.x, tadr2, padr2, hdr.x, tadr, padr, soft_t, hr, active, generic);
        } else {
          handled = false; // a combination the base graphs do not contain: one after the other, below
        }
        if (handled) {
          hdr = nxt2;
          m += 2;
          continue;
        }
      }
      // most frequent degrees first (BG1: 18 rows of degree 5, 8 of 6; BG2: 22 of 4, 7 of 5, 7 of 3)
      if (hdr.z == 5) {
        h2g_ext_row<5>(hdr.x, tadr, padr, soft_t, hr, active, generic);
      } else if (hdr.z == 4) {
        h2g_ext_row<4>(hdr.x, tadr, padr, soft_t, hr, active, generic);
      } else if (hdr.z == 6) {
        h2g_ext_row<6>(hdr.x, tadr, padr, soft_t, hr, active, generic);
      } else if (hdr.z == 7) {
        h2g_ext_row<7>(hdr.x, tadr, padr, soft_t, hr, active, generic);
      } else if (hdr.z == 3) {
        h2g_ext_row<3>(hdr.x, tadr, padr, soft_t, hr, active, generic);
      } else if (hdr.z == 8) {
        h2g_ext_row<8>(hdr.x, tadr, padr, soft_t, hr, active, generic);
      } else if (hdr.z == 9) {
        h2g_ext_row<9>(hdr.x, tadr, padr, soft_t, hr, active, generic);
      } else {
        h2g_ext_row<10>(hdr.x, tadr, padr, soft_t, hr, active, generic);
      }
      hdr = nxt;
      ++m;
    }
    it_done = it + 1;
    const bool last = (it_done == max_iter);
    if (!has_crc || (!early && !last)) {
      continue;
    }
    // ---- hard decision + CRC, as in pdc_kernel_h2.cuh -------------------------------------------------------------------------
    uint32_t acc = 0;
    if (active) {
      h2_static_for<G::KB>([&](auto c_) {
        constexpr int  c = decltype(c_)::value;
        const uint32_t S = h2_read_natural<BG, 0, c>(s_tab, soft_t);
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
    __syncthreads();
    uint32_t crc = 0;
    for (int w = 0; w < (nth >> 5); ++w) {
      crc ^= s_red[w];
    }
    success = (crc == 0u);
    if (success && early) { // the zero-soft-bit sweep only when the checksum matches (see pdc_kernel_h2.cuh)
      success = !h2_any_zero_soft<BG, 0>(s_tab, soft_t, active);
    }
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
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (tid < 32) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(s_tmem), "r"(tm_cols));
  }
}

// Dynamic shared memory of one code block with at most max_layers layers: core columns | parity words of the extension
// rows | max(staged decoder input, ballot words of the output). The messages of the extension rows live in tensor memory.
extern "C" size_t pdc_h2g_smem_bytes(int bg, int Z, int max_layers)
{
  const int    ncol = (bg == 1) ? 26 : 14, kb = (bg == 1) ? 22 : 10;
  const int    M    = (bg == 1) ? PDC_BG1_ROWS : PDC_BG2_ROWS;
  const int    L    = max_layers > M ? M : max_layers;
  const size_t out  = (size_t)kb * 48;
  size_t       stage = (size_t)(kb + L - 2) * Z + 32;
  stage              = stage > out ? stage : out;
  return (((size_t)ncol * 4 * Z + (size_t)(L - 4) * 2 * Z + 15) & ~(size_t)15) + stage;
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
