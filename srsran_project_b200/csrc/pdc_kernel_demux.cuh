// UL-SCH demultiplexer on the device: ulsch_demultiplex_impl.cpp. Included by pdc_demod.cu (it shares the
// scrambling-sequence generator, needed to undo the scrambling of the placeholder bits).
//
// One CTA per (codeword, OFDM symbol). The host has reduced the reference's per-symbol set construction (:330-463) to a few
// integers: every set is "every d-th candidate, the first m of them" (re_set_select, :77-100). The CTA rebuilds the sets:
//   reserved  = select(all REs)                                      rank closed-form (i / d)
//   HARQ-ACK  = select(all REs)           (more than 2 bits)         rank closed-form
//             = select(reserved)          (1 or 2 bits)              rank closed-form ((i / rvd_d) / d)
//   CSI 1     = select(not HARQ-ACK, not reserved)                   candidate rank by a block scan
//   CSI 2     = select(not HARQ-ACK (> 2 bits), not CSI 1)           candidate rank by a block scan
//   UL-SCH    = not HARQ-ACK (> 2 bits), not CSI 1, not CSI 2        rank by a block scan
// then moves the soft bits: every RE goes to its stream at rank * nbpr; REs punctured by a 1- or 2-bit HARQ-ACK go to the
// HARQ-ACK stream as well and reach their own stream as zeros (:487-500); placeholder REs get the scrambling of their x / y
// bits undone (:106-196).
#pragma once

#define PDC_DEMUX_THREADS 256
#define PDC_DEMUX_MAX_RE 3300 /* MAX_RB * NRE */

// Exclusive prefix of `count` over the CTA (contiguous REs per thread, so thread order is RE order).
__device__ __forceinline__ uint32_t demux_block_scan(uint32_t count, uint32_t* s_warp)
{
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t  incl = count;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) {
      incl += v;
    }
  }
  __syncthreads(); // s_warp may still be read from the previous scan
  if (lane == 31) {
    s_warp[warp] = incl;
  }
  __syncthreads();
  uint32_t base = 0;
#pragma unroll
  for (int w = 0; w < PDC_DEMUX_THREADS / 32; ++w) {
    base += (w < warp) ? s_warp[w] : 0u;
  }
  return base + incl - count;
}

__device__ __forceinline__ bool demux_selected(uint32_t rank, uint32_t d, uint32_t m)
{
  return m != 0 && (rank % d) == 0 && (rank / d) < m;
}

// Scrambling words of one OFDM symbol from the register states the host jumped to: 32 words of each LFSR by the bootstrap
// matrices, the rest by the word-level recurrence, xor-ed in place into w1. One warp; w1 / w2 hold nwords + 28 words.
__device__ __forceinline__ void demux_sequence(uint32_t* w1, uint32_t* w2, int nwords, uint32_t s1, uint32_t s2, int lane)
{
  w1[lane] = prg_word(d_prg_word, s1, 0, lane);
  w2[lane] = prg_word(d_prg_word, s2, 1, lane);
  __syncwarp();
  for (int base = 32; base < nwords; base += 28) {
    if (lane < 28) {
      const int j     = base + lane - 31;
      w1[base + lane] = w1[j + 3] ^ w1[j];
      w2[base + lane] = w2[j + 3] ^ w2[j + 2] ^ w2[j + 1] ^ w2[j];
    }
    __syncwarp();
  }
  for (int j = lane; j < nwords; j += 32) {
    w1[j] ^= w2[j];
  }
}

// "Every d-th candidate, the first m of them" (re_set_select) along consecutive candidate ranks: rank = q * d + r.
struct demux_walk {
  uint32_t q, r, d, m;
  __device__ __forceinline__ demux_walk(uint32_t d_, uint32_t m_, uint32_t rank) : q(0), r(0), d(d_), m(m_)
  {
    if (m != 0) {
      q = rank / d, r = rank - q * d;
    }
  }
  __device__ __forceinline__ bool selected() const { return m != 0 && r == 0 && q < m; }
  __device__ __forceinline__ void next()
  {
    if (++r == d) {
      r = 0, ++q;
    }
  }
};

// Placeholder handling of one soft bit: b = index within its modulation symbol, mask / mask_prev = scrambling bits of this soft
// bit and of the one before it. 1 bit (:106-150): bit 0 kept, bit 1 re-scrambled with the mask of bit 0, others unscrambled.
// 2 bits (:152-196): bits 0 and 1 kept, others unscrambled.
__device__ __forceinline__ int demux_placeholder(int v, uint32_t b, uint32_t mask, uint32_t mask_prev, int ph)
{
  if (b == 0 || (b == 1 && ph == 2)) {
    return v;
  }
  const uint32_t flip = (b == 1) ? (mask ^ mask_prev) : mask;
  return flip ? -v : v;
}

// Per-RE word in s_cls: [2:0] stream (0 UL-SCH, 1 HARQ-ACK, 2 CSI Part 1, 3 CSI Part 2), [3] punctured by a 1- / 2-bit
// HARQ-ACK, [15:4] rank in its stream, [27:16] rank in the HARQ-ACK stream when punctured.
template <int W>
struct demux_vec;
template <>
struct demux_vec<1> {
  using type = uint8_t;
};
template <>
struct demux_vec<4> {
  using type = uint32_t;
};
template <>
struct demux_vec<16> {
  using type = uint4;
};

// W bytes per thread and step (W divides nbpr and every pointer is W-aligned, so a step never straddles two REs).
template <int W>
__device__ __forceinline__ void demux_move(const pdc_dev_demux_sym& sy, const uint32_t* s_cls, const uint32_t* s_seq, bool any_ph)
{
  using vec_t          = typename demux_vec<W>::type;
  // This CTA's share of the symbol's bytes (the sets are rebuilt by every CTA of the symbol, the bytes are moved once).
  const uint32_t total = sy.nof_re * sy.nbpr;
  const uint32_t chunk = ((total + gridDim.y - 1) / gridDim.y + 15u) & ~15u;
  const uint32_t begin = min(total, blockIdx.y * chunk), end = min(total, begin + chunk);
  const uint32_t qm    = sy.Qm;
  for (uint32_t j = begin + threadIdx.x * W; j < end; j += PDC_DEMUX_THREADS * W) {
    const uint32_t re = j / sy.nbpr, k = j - re * sy.nbpr;
    const uint32_t c  = s_cls[re];
    const uint32_t stream = c & 7u, rank = (c >> 4) & 0xfffu;
    const bool     punct  = (c & 8u) != 0;
    const vec_t    raw    = *reinterpret_cast<const vec_t*>(sy.in + j);
    int8_t*        dst    = (stream == 0) ? sy.sch : (stream == 1) ? sy.harq : (stream == 2) ? sy.csi1 : sy.csi2;
    const int      ph     = (stream == 1) ? sy.harq_ph : (stream == 2) ? sy.csi1_ph : (stream == 3) ? sy.csi2_ph : 0;
    dst += rank * sy.nbpr + k;
    if (!punct && (ph == 0 || !any_ph)) {
      *reinterpret_cast<vec_t*>(dst) = raw; // the bulk: UL-SCH and long UCI fields are plain copies
      continue;
    }
    union {
      vec_t   v;
      uint8_t b[W];
    } in_u, own, hq;
    in_u.v = raw;
#pragma unroll
    for (int q = 0; q < W; ++q) {
      int v = (int8_t)in_u.b[q];
      int h = v;
      if (any_ph && qm > 1) {
        const uint32_t p = j + q, b = (k + q) % qm;
        const uint32_t mk = (s_seq[p >> 5] >> (p & 31u)) & 1u;
        const uint32_t mp = (p == 0) ? 0u : (s_seq[(p - 1) >> 5] >> ((p - 1) & 31u)) & 1u;
        if (punct) {
          h = demux_placeholder(v, b, mk, mp, sy.harq_ph);
        }
        if (ph != 0) {
          v = demux_placeholder(v, b, mk, mp, ph);
        }
      }
      hq.b[q]  = (uint8_t)h;
      own.b[q] = punct ? 0 : (uint8_t)v; // srsvec::zero(re_data) after the HARQ-ACK extraction
    }
    if (punct) {
      const uint32_t rh = (c >> 16) & 0xfffu;
      *reinterpret_cast<vec_t*>(sy.harq + rh * sy.nbpr + k) = hq.v;
    }
    *reinterpret_cast<vec_t*>(dst) = own.v;
  }
}

// A symbol without UCI: its soft bits are the next stretch of the UL-SCH stream.
__device__ __forceinline__ void demux_copy(const int8_t* __restrict__ src, int8_t* __restrict__ dst, uint32_t n)
{
  const uintptr_t a = reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst);
  uint32_t        done = 0;
  if ((a & 15u) == 0) {
    const uint32_t nv = n >> 4;
#pragma unroll 4
    for (uint32_t j = threadIdx.x; j < nv; j += PDC_DEMUX_THREADS) {
      reinterpret_cast<uint4*>(dst)[j] = __ldcs(reinterpret_cast<const uint4*>(src) + j);
    }
    done = nv << 4;
  } else if ((a & 3u) == 0) {
    const uint32_t nv = n >> 2;
    for (uint32_t j = threadIdx.x; j < nv; j += PDC_DEMUX_THREADS) {
      reinterpret_cast<uint32_t*>(dst)[j] = reinterpret_cast<const uint32_t*>(src)[j];
    }
    done = nv << 2;
  }
  for (uint32_t j = done + threadIdx.x; j < n; j += PDC_DEMUX_THREADS) {
    dst[j] = src[j];
  }
}

// OFDM symbols without UCI (the host lists them first): plain stretches of the UL-SCH stream, PDC_DEMUX_COPY_PARTS CTAs each, no
// shared memory, so the bulk of a slot moves at copy speed beside the few CTAs that build RE sets.
#define PDC_DEMUX_COPY_PARTS 4
#define PDC_DEMUX_UCI_PARTS 4 /* CTAs per symbol with UCI: a slot has few such symbols, one CTA each would leave the device idle */
__global__ void __launch_bounds__(PDC_DEMUX_THREADS) pdc_ulsch_copy_kernel(const pdc_dev_demux_sym* __restrict__ syms)
{
  const pdc_dev_demux_sym& sy = syms[blockIdx.x];
  const uint32_t           n  = sy.nof_re * sy.nbpr;
  const uint32_t           chunk = ((n + gridDim.y - 1) / gridDim.y + 15u) & ~15u;
  const uint32_t           begin = min(n, blockIdx.y * chunk), end = min(n, begin + chunk);
  demux_copy(sy.in + begin, sy.sch + begin, end - begin);
}

__global__ void __launch_bounds__(PDC_DEMUX_THREADS) pdc_ulsch_demux_kernel(const pdc_dev_demux_sym* __restrict__ syms)
{
  extern __shared__ uint32_t s_dyn[];
  __shared__ uint32_t        s_warp[PDC_DEMUX_THREADS / 32];
  const pdc_dev_demux_sym    sy    = syms[blockIdx.x];
  uint32_t*                  s_cls = s_dyn;                    // nof_re words
  uint32_t*                  s_w1  = s_dyn + PDC_DEMUX_MAX_RE; // sequence words (and the x1 history), then the x2 history
  const uint32_t             M     = sy.nof_re;
  const int                  tid   = threadIdx.x;
  // (Symbols without UCI never get here: the host hands them to pdc_ulsch_copy_kernel. Reserved REs without HARQ-ACK on them
  // change nothing, so such a symbol counts as one without UCI.)
  const bool any_ph = sy.Qm > 1 && ((sy.harq_m != 0 && sy.harq_ph != 0) || (sy.csi1_m != 0 && sy.csi1_ph != 0) ||
                                    (sy.csi2_m != 0 && sy.csi2_ph != 0));

  // ---- scrambling bits of the symbol (only when a placeholder RE lies in it): the last warp ------------------------------------
  const uint32_t nwords = (M * sy.nbpr + 31) / 32 + 1;
  if (any_ph && tid >= PDC_DEMUX_THREADS - 32) {
    demux_sequence(s_w1, s_w1 + nwords + 64, (int)nwords, sy.x1_state, sy.x2_state, tid & 31);
  }

  // ---- sets: contiguous REs per thread, up to three scans. Every "every d-th, the first m" test walks its candidates with a
  //      (quotient, remainder) pair: one division per thread and set instead of two per RE. -----------------------------------------
  const uint32_t per = (M + PDC_DEMUX_THREADS - 1) / PDC_DEMUX_THREADS;
  const uint32_t i0 = min(M, (uint32_t)tid * per), i1 = min(M, i0 + per);
  const bool     hq_all = !sy.harq_of_rvd && sy.harq_m != 0; // HARQ-ACK of more than two bits: selected among all REs
  // CSI Part 1 among the REs that are neither HARQ-ACK nor reserved.
  {
    demux_walk rv(sy.rvd_d, sy.rvd_m, i0), hq(sy.harq_d, hq_all ? sy.harq_m : 0u, i0);
    uint32_t   n = 0;
    for (uint32_t i = i0; i < i1; ++i) {
      n += (!hq.selected() && !rv.selected()) ? 1u : 0u;
      rv.next(), hq.next();
    }
    const uint32_t r0 = (sy.csi1_m != 0) ? demux_block_scan(n, s_warp) : 0u;
    demux_walk     rv2(sy.rvd_d, sy.rvd_m, i0), hq2(sy.harq_d, hq_all ? sy.harq_m : 0u, i0), c1(sy.csi1_d, sy.csi1_m, r0);
    for (uint32_t i = i0; i < i1; ++i) {
      uint32_t c = 0;
      if (hq2.selected()) {
        c = 1u | (hq2.q << 4);
      } else if (!rv2.selected()) {
        if (c1.selected()) {
          c = 2u | (c1.q << 4);
        }
        c1.next();
      }
      s_cls[i] = c;
      rv2.next(), hq2.next();
    }
  }
  // CSI Part 2 among what is left of the UCI candidates (reserved REs included).
  if (sy.csi2_m != 0) {
    uint32_t n = 0;
    for (uint32_t i = i0; i < i1; ++i) {
      n += (s_cls[i] == 0) ? 1u : 0u;
    }
    demux_walk c2(sy.csi2_d, sy.csi2_m, demux_block_scan(n, s_warp));
    for (uint32_t i = i0; i < i1; ++i) {
      if (s_cls[i] == 0) {
        if (c2.selected()) {
          s_cls[i] = 3u | (c2.q << 4);
        }
        c2.next();
      }
    }
  }
  // UL-SCH: everything else; HARQ-ACK of 1 or 2 bits punctures its reserved REs whatever they carry.
  {
    uint32_t n = 0;
    for (uint32_t i = i0; i < i1; ++i) {
      n += (s_cls[i] == 0) ? 1u : 0u;
    }
    uint32_t   r = demux_block_scan(n, s_warp);
    demux_walk rv(sy.rvd_d, sy.harq_of_rvd ? sy.rvd_m : 0u, i0);
    // HARQ-ACK among the reserved REs: walks the rank of the reserved RE (rv.q), which advances when rv wraps.
    demux_walk hq(sy.harq_d, (sy.harq_of_rvd && sy.rvd_m != 0) ? sy.harq_m : 0u, rv.r == 0 ? rv.q : rv.q + 1);
    for (uint32_t i = i0; i < i1; ++i) {
      uint32_t c = s_cls[i];
      if (c == 0) {
        c = r << 4;
        ++r;
      }
      if (rv.r == 0 && rv.m != 0) {
        if (rv.q < rv.m && hq.selected()) {
          c |= 8u | (hq.q << 16);
        }
        hq.next();
      }
      s_cls[i] = c;
      rv.next();
    }
  }
  __syncthreads();

  // ---- soft bits ----------------------------------------------------------------------------------------------------------------
  const uintptr_t align = reinterpret_cast<uintptr_t>(sy.in) | reinterpret_cast<uintptr_t>(sy.sch) |
                          reinterpret_cast<uintptr_t>(sy.harq) | reinterpret_cast<uintptr_t>(sy.csi1) |
                          reinterpret_cast<uintptr_t>(sy.csi2) | sy.nbpr;
  if ((align & 15u) == 0) {
    demux_move<16>(sy, s_cls, s_w1, any_ph);
  } else if ((align & 3u) == 0) {
    demux_move<4>(sy, s_cls, s_w1, any_ph);
  } else {
    demux_move<1>(sy, s_cls, s_w1, any_ph);
  }
}

// d_syms: nplain symbols without UCI, then nuci symbols with some.
extern "C" cudaError_t pdc_demux_launch(const pdc_dev_demux_sym* d_syms, uint32_t nplain, uint32_t nuci, bool placeholders,
                                        cudaStream_t stream)
{
  if (nplain != 0) {
    pdc_ulsch_copy_kernel<<<dim3(nplain, PDC_DEMUX_COPY_PARTS), PDC_DEMUX_THREADS, 0, stream>>>(d_syms);
  }
  if (nuci != 0) {
    // Set words; with placeholders also the sequence words of the largest symbol (3300 REs x 32 soft bits) and the two histories.
    const size_t smem = sizeof(uint32_t) * (PDC_DEMUX_MAX_RE + (placeholders ? 2 * (PDC_DEMUX_MAX_RE + 2 + 64) : 0));
    pdc_ulsch_demux_kernel<<<dim3(nuci, PDC_DEMUX_UCI_PARTS), PDC_DEMUX_THREADS, smem, stream>>>(d_syms + nplain);
  }
  return cudaGetLastError();
}
