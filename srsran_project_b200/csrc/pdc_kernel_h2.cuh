// Code-block kernel for high-rate code blocks: packed half2 arithmetic, check-to-variable messages in registers.
//
// Scope: code blocks whose decode touches only the four core rows of the base graph (nof_layers == 4, i.e. the code rate
// is high enough that no extension row carries received LLRs: ldpc_decoder_impl.cpp:106-115) and whose lifting size is
// even (44 of the 51 sizes). The operating point of the metric is 100 MHz / 256QAM PUSCH: BG1, Z = 384, E = 8960 | 8992.
// More layers: pdc_kernel_h2g.cuh (which reuses this file's row code); odd lifting sizes: pdc_kernel_lane.cuh.
//
// Mapping. One CTA per code block, Z/2 threads; thread t owns check lanes (2t, 2t+1) of EVERY row, packed as one half2.
//   * Soft bits live in shared memory as IEEE halves. Integers up to 2048 are exact in a half, and the reference's
//     "+-127 means infinity" (log_likelihood_ratio.h:250-255) becomes a real +-inf, so saturation, promotion and the
//     sticky infinities of ldpc_decoder_avx512.cpp:86-126,229-270 each cost one or two half2 instructions instead of
//     byte-wise emulation (sm_100a has no native byte SIMD except VABSDIFF4).
//   * The 76 (BG1) / 36 (BG2) check-to-variable messages of a lane pair stay in REGISTERS for the whole decode: they are
//     only ever touched by the thread that owns the check lane, so they never travel through shared memory.
//   * Circulant shifts: every column is stored in the lane order of the row that wrote it last ("lazy rotation"), twice
//     back to back, so that the next row reads its rotated view as a plain offset (the same idea as the reference's
//     doubled node buffer, ldpc_decoder_impl.h:95-98) and always writes whole aligned words. An odd relative shift costs
//     one PRMT to realign the pair. The relative shifts are computed on the host per (base graph, Z) and sit in constant
//     memory.
//   * The instruction mix is split over both issue pipes on purpose: min / max / logic on the ALU pipe, add / promotion /
//     scaling as HFMA2 on the FMA pipe (see the pipe utilisation in profiles/).
//
// Two instantiations per base graph: ZC == 0 takes the lifting size at run time (relative shifts from the constant-memory
// table below); ZC == 384 is the metric's operating point with every shift, offset and PRMT selector folded into the
// instruction stream as an immediate (no table loads, no uniform-datapath address arithmetic, and no second load + PRMT
// for the edges whose relative shift is even).
//
// Arithmetic (bit-exact with ldpc_decoder_avx512 / ldpc_decoder_generic; model and proof-by-test in tests/h2_model.py):
//   u   = S - c_old                                     v2c, unclamped; +-inf stays +-inf
//   m1, m2 = two smallest |u| capped at 120; par = xor of signs
//   s1, s2 = scale(m1), scale(m2)                       HFMA2 tricks, exact for 0..120 (tests/test_h2_model.py)
//   c_new = +-( |u| == m1 ? s2 : s1 ), sign = par ^ sign(u)   (ties make s1 == s2, so no argmin is needed)
//   t   = clamp(u, +-120) + c_new
//   S'  = t * (1 + 65504 * sat(max(|t|, |u| - 216) - 120))    |t| > 120 or u infinite  ->  overflow to +-inf
#pragma once
#include "ldpc_bg_tables.h"
#include "pdc_device.cuh"
#include <utility>

#define PDC_H2_TAB_STRIDE 128 /* u32 per (bg, Z): 76 edge entries (offset | PRMT selector << 16), 26 column alignments,
                                 26 natural-order views (offset | selector) */
#define PDC_H2_SIGMA_OFF 76
#define PDC_H2_NAT_OFF 102
#if defined(PDC_KERNEL_TU)
__constant__ uint32_t c_h2_tab[2 * PDC_NOF_LIFTING_SIZES][PDC_H2_TAB_STRIDE];
#endif

// ---- half2 instructions on raw 32-bit registers ------------------------------------------------------------------------
#define PDC_H2_OP2(name, ptx)                                                                                          \
  __device__ __forceinline__ uint32_t name(uint32_t a, uint32_t b)                                                     \
  {                                                                                                                    \
    uint32_t d;                                                                                                        \
    asm(ptx " %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));                                                                \
    return d;                                                                                                          \
  }
PDC_H2_OP2(h2_add, "add.f16x2")
PDC_H2_OP2(h2_sub, "sub.f16x2")
PDC_H2_OP2(h2_mul, "mul.f16x2")
PDC_H2_OP2(h2_min, "min.f16x2")
PDC_H2_OP2(h2_max, "max.f16x2")
PDC_H2_OP2(h2_eq_mask, "set.eq.u32.f16x2")
PDC_H2_OP2(h2_le_mask, "set.le.u32.f16x2")
__device__ __forceinline__ uint32_t h2_abs(uint32_t a)
{
  uint32_t d;
  asm("abs.f16x2 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}
__device__ __forceinline__ uint32_t h2_fma(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t d;
  asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ uint32_t h2_fma_sat(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t d;
  asm("fma.rn.sat.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ uint32_t h2_add_sat(uint32_t a, uint32_t b)
{
  uint32_t d;
  asm("add.rn.sat.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
// (a & 0x80008000) | one as ONE LOP3: the 1.0 pair has to live in a register the compiler cannot fold (an instruction takes
// a single immediate), hence the opaque move.
__device__ __forceinline__ uint32_t h2_opaque_one()
{
  uint32_t r;
  asm volatile("mov.b32 %0, 0x3C003C00;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t h2_copysign_one(uint32_t a, uint32_t one)
{
  uint32_t d;
  asm("lop3.b32 %0, %1, 0x80008000, %2, 0xEA;" : "=r"(d) : "r"(a), "r"(one));
  return d;
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr)
{
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v)
{
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v));
}
// Same with a compile-time byte offset that ends up in the instruction's immediate field.
template <int OFF>
__device__ __forceinline__ uint32_t lds32i(uint32_t addr)
{
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF));
  return v;
}
template <int OFF>
__device__ __forceinline__ void sts32i(uint32_t addr, uint32_t v)
{
  asm volatile("st.shared.u32 [%0+%1], %2;" ::"r"(addr), "n"(OFF), "r"(v));
}
__device__ __forceinline__ uint32_t lds16(uint32_t addr)
{
  uint32_t v;
  asm volatile("{ .reg .b16 t; ld.shared.u16 t, [%1]; cvt.u32.u16 %0, t; }" : "=r"(v) : "r"(addr));
  return v;
}

static constexpr uint32_t H2_120 = 0x57805780u, H2_N120 = 0xD780D780u, H2_64 = 0x54005400u, H2_N64 = 0xD400D400u;
static constexpr uint32_t H2_N216 = 0xDAC0DAC0u, H2_65504 = 0x7BFF7BFFu, H2_ONE = 0x3C003C00u, H2_1024 = 0x64006400u;
static constexpr uint32_t H2_N1024 = 0xE400E400u, H2_1023_5 = 0x63FF63FFu, H2_0_8 = 0x3A663A66u, H2_N1152 = 0xE480E480u;
static constexpr uint32_t H2_SIGN = 0x80008000u;
// Soft bits sit in shared memory with a bias of 1152 = 1024 + 128: the half 1152 + x is the bit pattern 0x6400 | (x + 128)
// for every int8 x, which is what makes packing / unpacking int8 <-> half a single PRMT (see h2_row).
static constexpr uint32_t H2_BIAS = 0x64806480u, H2_BIAS_P64 = 0x64C064C0u, H2_BIAS_N64 = 0x64406440u;
static constexpr uint32_t H2_N336 = 0xDD40DD40u;

// Normalisation of a non-negative integer <= 120 held in both halves: (x * 52428) >> 16 (AVX2 / AVX512) or
// round(0.8f * x) (generic). One fused multiply-add whose single rounding lands on the integer we want.
__device__ __forceinline__ uint32_t h2_scale(uint32_t m, bool generic)
{
  if (generic) {
    return h2_add(h2_fma(m, H2_0_8, H2_1024), H2_N1024);
  }
  return h2_add(h2_max(h2_fma(m, H2_0_8, H2_1023_5), H2_1024), H2_N1024);
}

// ---- compile-time structure of the four core rows -------------------------------------------------------------------------
template <int BG>
struct h2_graph;
template <>
struct h2_graph<1> {
  static constexpr int KB = 22, NCOL = 26, NEDGE = 76;
  __host__ __device__ static constexpr int deg(int m) { return pdc_bg1_row_deg[m]; }
  __host__ __device__ static constexpr int col(int e) { return pdc_bg1_edge_col[e]; }
};
template <>
struct h2_graph<2> {
  static constexpr int KB = 10, NCOL = 14, NEDGE = 36;
  __host__ __device__ static constexpr int deg(int m) { return pdc_bg2_row_deg[m]; }
  __host__ __device__ static constexpr int col(int e) { return pdc_bg2_edge_col[e]; }
};
template <int BG>
__host__ __device__ constexpr int h2_shift(int ils, int e)
{
  return (BG == 1) ? pdc_bg1_edge_shift[ils][e] : pdc_bg2_edge_shift[ils][e];
}
__host__ __device__ constexpr int h2_set_index(int Z)
{
  const int base[8] = {2, 3, 5, 7, 9, 11, 13, 15};
  int       ils     = 0;
  for (int i = 0; i != 8; ++i) {
    int z = base[i];
    while (z < Z) {
      z *= 2;
    }
    ils = (z == Z) ? i : ils;
  }
  return ils;
}
// Alignment of column c after a full iteration = shift of the last core-row edge that holds it (0 if none does).
template <int BG>
__host__ __device__ constexpr int h2_sigma(int Z, int c)
{
  const int ils = h2_set_index(Z);
  int       sg  = 0;
  for (int e = 0; e != h2_graph<BG>::NEDGE; ++e) {
    if (h2_graph<BG>::col(e) == c) {
      sg = h2_shift<BG>(ils, e) % Z;
    }
  }
  return sg;
}
// Rotation of edge e's view relative to how its column was stored by the previous writer (the nearest earlier edge on the
// same column, cyclically: the previous iteration's last writer for the first edge of a column).
template <int BG>
__host__ __device__ constexpr int h2_rel(int Z, int e)
{
  const int ils  = h2_set_index(Z);
  int       prev = h2_sigma<BG>(Z, h2_graph<BG>::col(e));
  for (int e2 = 0; e2 != e; ++e2) {
    if (h2_graph<BG>::col(e2) == h2_graph<BG>::col(e)) {
      prev = h2_shift<BG>(ils, e2) % Z; // edges are listed row by row: a later match is a later row
    }
  }
  return ((h2_shift<BG>(ils, e) % Z) - prev + Z) % Z;
}
template <int BG>
__host__ __device__ constexpr int h2_row_start(int m)
{
  int s = 0;
  for (int i = 0; i < m; ++i) {
    s += h2_graph<BG>::deg(i);
  }
  return s;
}

template <class F, int... I>
__device__ __forceinline__ void h2_static_for_impl(F&& f, std::integer_sequence<int, I...>)
{
  (f(std::integral_constant<int, I>{}), ...);
}
template <int N, class F>
__device__ __forceinline__ void h2_static_for(F&& f)
{
  h2_static_for_impl(static_cast<F&&>(f), std::make_integer_sequence<int, N>{});
}

// 16 bits -> 32 bits with a zero between neighbours.
__device__ __forceinline__ uint32_t h2_spread16(uint32_t x)
{
  x = (x | (x << 8)) & 0x00ff00ffu;
  x = (x | (x << 4)) & 0x0f0f0f0fu;
  x = (x | (x << 2)) & 0x33333333u;
  x = (x | (x << 1)) & 0x55555555u;
  return x;
}

// The rotated view of edge E (absolute edge index) of the thread's lane pair: compile-time offsets for ZC > 0, else the
// (offset | PRMT selector << 16) entry of the constant-memory table.
template <int BG, int ZC, int E>
__device__ __forceinline__ uint32_t h2_read_edge(const uint32_t* __restrict__ tab, uint32_t soft_t)
{
  if constexpr (ZC > 0) {
    constexpr int rel = h2_rel<BG>(ZC, E);
    constexpr int off = h2_graph<BG>::col(E) * 4 * ZC + 4 * (rel >> 1);
    if constexpr ((rel & 1) == 0) {
      return lds32i<off>(soft_t);
    } else {
      const uint32_t w0 = lds32i<off>(soft_t), w1 = lds32i<off + 4>(soft_t);
      return __byte_perm(w0, w1, 0x5432);
    }
  } else {
    const uint32_t ent = tab[E];
    const uint32_t adr = soft_t + (ent & 0xffffu);
    const uint32_t w0 = lds32(adr), w1 = lds32(adr + 4);
    return __byte_perm(w0, w1, ent >> 16);
  }
}
// Natural-order view (lanes 2 tid, 2 tid + 1) of column C.
template <int BG, int ZC, int C>
__device__ __forceinline__ uint32_t h2_read_natural(const uint32_t* __restrict__ tab, uint32_t soft_t)
{
  if constexpr (ZC > 0) {
    constexpr int nat = (ZC - h2_sigma<BG>(ZC, C)) % ZC;
    constexpr int off = C * 4 * ZC + 4 * (nat >> 1);
    if constexpr ((nat & 1) == 0) {
      return lds32i<off>(soft_t);
    } else {
      const uint32_t w0 = lds32i<off>(soft_t), w1 = lds32i<off + 4>(soft_t);
      return __byte_perm(w0, w1, 0x5432);
    }
  } else {
    const uint32_t ent = tab[PDC_H2_NAT_OFF + C];
    const uint32_t adr = soft_t + (ent & 0xffffu);
    const uint32_t w0 = lds32(adr), w1 = lds32(adr + 4);
    return __byte_perm(w0, w1, ent >> 16);
  }
}
// Both copies of the thread's word of column C.
template <int ZC, int C>
__device__ __forceinline__ void h2_write_col(uint32_t soft_t, uint32_t row_bytes, uint32_t v)
{
  if constexpr (ZC > 0) {
    sts32i<C * 4 * ZC>(soft_t, v);
    sts32i<C * 4 * ZC + 2 * ZC>(soft_t, v);
  } else {
    const uint32_t adr = soft_t + C * row_bytes;
    sts32(adr, v);
    sts32(adr + (row_bytes >> 1), v);
  }
}

// ---- check-to-variable messages: int8 pairs, two edges per register ---------------------------------------------------------
// A message is an integer in [-96, 96]; stored as the byte c + 128, edge k of a row in register k / 2, bytes (2 (k % 2),
// 2 (k % 2) + 1) = (lane 2 tid, lane 2 tid + 1). PRMT with the constant 0x64 turns the two bytes into the half2
// (1152 + c, 1152 + c) without arithmetic; because the soft bits carry the same bias, S_biased - that = S - c.
template <int BG>
__host__ __device__ constexpr int h2_pack_start(int m)
{
  int s = 0;
  for (int i = 0; i < m; ++i) {
    s += (h2_graph<BG>::deg(i) + 1) / 2;
  }
  return s;
}
template <int BG>
struct h2_pack {
  static constexpr int NREG = h2_pack_start<BG>(4);
};

// Two smallest of the node list (lo[i] <= hi[i]); pairs of nodes merge with min / max / min3.
template <int N>
__device__ __forceinline__ void h2_min2_tree(const uint32_t (&lo)[N], const uint32_t (&hi)[N], uint32_t& m1, uint32_t& m2)
{
  if constexpr (N == 1) {
    m1 = lo[0], m2 = hi[0];
  } else {
    constexpr int H = (N + 1) / 2;
    uint32_t      l2[H], h2[H];
#pragma unroll
    for (int i = 0; i < N / 2; ++i) {
      l2[i] = h2_min(lo[2 * i], lo[2 * i + 1]);
      h2[i] = h2_min(h2_min(h2_max(lo[2 * i], lo[2 * i + 1]), hi[2 * i]), hi[2 * i + 1]);
    }
    if constexpr ((N & 1) != 0) {
      l2[H - 1] = lo[N - 1], h2[H - 1] = hi[N - 1];
    }
    h2_min2_tree<H>(l2, h2, m1, m2);
  }
}

// ---- output: K hard bits in natural order, MSB first (bit_buffer.h:106-136) -----------------------------------------------
// s_nat[(c * 6 + w) * 2 + {0, 1}]: the 64 hard bits of lanes 64 w .. 64 w + 63 of column c, natural order, LSB first (built
// from the two warp ballots of the even and the odd lanes). Word idx of the output holds bits 32 idx .. 32 idx + 31; when Z is
// a multiple of 32 that is half of one entry, otherwise it is stitched from pieces of up to 32 / Z + 1 columns.
__device__ __forceinline__ uint32_t h2_output_word(const uint32_t* s_nat, int idx, int Z, int K)
{
  uint32_t nat;
  if ((Z & 31) == 0) {
    const int nw = Z >> 5, c = idx / nw, k = idx - c * nw;
    nat          = s_nat[(c * 6 + (k >> 1)) * 2 + (k & 1)];
  } else {
    nat        = 0;
    int n      = 32 * idx;
    int filled = 0;
    while (filled < 32 && n < K) {
      const int c = n / Z, l = n - c * Z;
      const int cnt = min(32 - filled, Z - l);
      const int w = l >> 6, o = l & 63;
      unsigned long long v = ((unsigned long long)s_nat[(c * 6 + w) * 2 + 1] << 32) | s_nat[(c * 6 + w) * 2];
      v >>= o;
      if (o != 0 && w + 1 < 6) {
        const unsigned long long nx = ((unsigned long long)s_nat[(c * 6 + w + 1) * 2 + 1] << 32) | s_nat[(c * 6 + w + 1) * 2];
        v |= nx << (64 - o);
      }
      const uint32_t piece = (uint32_t)v & ((cnt == 32) ? 0xffffffffu : ((1u << cnt) - 1u));
      nat |= piece << filled;
      filled += cnt;
      n += cnt;
    }
  }
  return __byte_perm(__brev(nat), 0u, 0x0123);
}

// The output phase of a packed kernel. Every thread reads its natural lane pair of each systematic column (as the CRC pass
// does); two warp ballots give the even and the odd lanes of 64 consecutive bits, which lane 0 stores as they are. Then all
// threads interleave them into natural-order words (in parallel: a warp instruction with one active lane costs a full issue
// slot) and store the output with coalesced words.
template <int BG, int ZC, typename TAB>
__device__ __forceinline__ void
h2_emit_bits(TAB tab, uint32_t soft_t, uint32_t* s_nat, bool active, int tid, int nth, int Z, int K, uint32_t* out32)
{
  using G = h2_graph<BG>;
  if (nth < 192) { // warps that do not exist (small Z) hold no bits
    for (int q = tid; q < 2 * 6 * G::KB; q += nth) {
      s_nat[q] = 0u;
    }
    __syncthreads();
  }
  h2_static_for<G::KB>([&](auto c_) {
    constexpr int  c  = decltype(c_)::value;
    const uint32_t S  = h2_read_natural<BG, ZC, c>(tab, soft_t);
    const uint32_t le = active ? h2_le_mask(S, H2_BIAS) : 0u;
    const uint32_t be = __ballot_sync(0xffffffffu, (le & 0x0000ffffu) != 0);
    const uint32_t bo = __ballot_sync(0xffffffffu, (le & 0xffff0000u) != 0);
    if ((tid & 31) == 0) {
      s_nat[(c * 6 + (tid >> 5)) * 2 + 0] = be; // 4-byte accesses: the scratch area is only word aligned for small Z
      s_nat[(c * 6 + (tid >> 5)) * 2 + 1] = bo;
    }
  });
  __syncthreads();
  for (int q = tid; q < 6 * G::KB; q += nth) { // one (column, warp) pair of words per task, converted in place
    const uint32_t be = s_nat[2 * q], bo = s_nat[2 * q + 1];
    s_nat[2 * q]     = h2_spread16(be & 0xffffu) | (h2_spread16(bo & 0xffffu) << 1);
    s_nat[2 * q + 1] = h2_spread16(be >> 16) | (h2_spread16(bo >> 16) << 1);
  }
  __syncthreads();
  for (int idx = tid; idx < ((K + 31) >> 5); idx += nth) {
    out32[idx] = h2_output_word(s_nat, idx, Z, K);
  }
}

// Pass 1, second half: the two smallest magnitudes of u[0..DEG) capped at 120, scaled, and the row's sign parity.
// Out: m1 (smallest magnitude), s1p (scaled m1 with the parity sign), d (scaled m1 xor scaled m2).
template <int DEG>
__device__ __forceinline__ void h2_row_minima(const uint32_t (&u)[DEG], uint32_t par, bool generic, uint32_t& m1,
                                              uint32_t& s1p, uint32_t& d)
{
  constexpr int NLEAF = DEG / 2 + 1; // pairs of edges, plus (last edge | 120, 120): caps both minima at 120
  uint32_t      lo[NLEAF], hi[NLEAF];
  h2_static_for<DEG / 2>([&](auto i_) {
    constexpr int  i = decltype(i_)::value;
    const uint32_t a = h2_abs(u[2 * i]), b = h2_abs(u[2 * i + 1]);
    lo[i] = h2_min(a, b), hi[i] = h2_max(a, b);
  });
  if constexpr ((DEG & 1) != 0) {
    lo[NLEAF - 1] = h2_min(h2_abs(u[DEG - 1]), H2_120), hi[NLEAF - 1] = H2_120;
  } else {
    lo[NLEAF - 1] = H2_120, hi[NLEAF - 1] = H2_120;
  }
  uint32_t m2;
  h2_min2_tree<NLEAF>(lo, hi, m1, m2);
  const uint32_t s1 = h2_scale(m1, generic);
  const uint32_t s2 = h2_scale(m2, generic);
  d                 = s1 ^ s2;
  s1p               = s1 | (par & H2_SIGN);
}

// Pass 2 for one edge: in u = v2c message (S - c_old); out u = new biased soft value, y = new message as 1152 + c.
__device__ __forceinline__ void h2_edge_update(uint32_t& u, uint32_t& y, uint32_t m1, uint32_t s1p, uint32_t d, uint32_t one)
{
  const uint32_t a  = h2_abs(u);
  const uint32_t eq = h2_eq_mask(a, m1);
  const uint32_t ms = s1p ^ (d & eq);
  const uint32_t sg = h2_copysign_one(u, one);
  const uint32_t qa = h2_add_sat(a, H2_N336);                    // 1 iff u is infinite (finite |u| <= 216)
  const uint32_t w  = h2_add(h2_fma(qa, H2_65504, h2_min(a, H2_120)), ms); // min(|u|, 120) + ms, infinite iff u is
  const uint32_t qb = h2_fma_sat(w, H2_ONE, H2_N120);            // 1 iff w > 120 (w >= -96: no negative overflow)
  const uint32_t sm = h2_fma(qb, H2_65504, w);                   // w, or +inf when promoted (121 + 65504 rounds to inf)
  y                 = h2_fma(ms, sg, H2_BIAS);
  u                 = h2_fma(sm, sg, H2_BIAS);                   // the bias also turns a -0 product into the reference's +0
}

// One row (layer) of the decode for the lane pair of this thread. cp: the thread's packed message registers.
//
// Arithmetic of pass 2 in sign-normalised form (sg = sign of u as +-1, everything else is computed on magnitudes, so the
// work is spread over both issue pipes: 4 ALU-pipe and 7 FMA-pipe instructions per edge):
//   ms = (|u| == m1 ? s2 : s1) with the sign of the row parity          (magnitude of the new message, sign-normalised)
//   w  = min(|u|, 120) + ms                                             (sign-normalised soft value: t = sg w)
//   w  becomes +inf when u is infinite (qa 65504 + 120 overflows), sm = w + qb 65504 becomes +inf when w > 120  (qa, qb)
//   S' = sg sm + bias              overflow = promotion to +-inf; the bias turns a -0 product into the reference's +0
//   c' = sg ms                     stored as the byte c' + 128
template <int BG, int ZC, int M>
__device__ __forceinline__ void h2_row(uint32_t (&cp)[h2_pack<BG>::NREG], const uint32_t* __restrict__ tab,
                                       uint32_t soft_t, uint32_t row_bytes, bool active, bool generic)
{
  using G             = h2_graph<BG>;
  constexpr int DEG   = G::deg(M);
  constexpr int E0    = h2_row_start<BG>(M);
  constexpr int P0    = h2_pack_start<BG>(M);
  uint32_t      u[DEG];
  uint32_t      par = 0;
  const uint32_t one = h2_opaque_one();
  // ---- pass 1: variable-to-check messages, two smallest magnitudes, sign parity ------------------------------------------
  h2_static_for<DEG>([&](auto k_) {
    constexpr int  k = decltype(k_)::value;
    const uint32_t S = h2_read_edge<BG, ZC, E0 + k>(tab, soft_t);
    const uint32_t x = __byte_perm(cp[P0 + k / 2], 0x64646464u, (k & 1) ? 0x4342 : 0x4140);
    u[k]             = h2_sub(S, x);
    par ^= u[k];
  });
  uint32_t s1p, d, m1;
  h2_row_minima<DEG>(u, par, generic, m1, s1p, d);
  // ---- pass 2: new check-to-variable messages and soft bits (kept in u[] until every thread has read the old ones) ---------
  uint32_t y[DEG];
  h2_static_for<DEG>([&](auto k_) {
    constexpr int k = decltype(k_)::value;
    h2_edge_update(u[k], y[k], m1, s1p, d, one);
  });
  h2_static_for<(DEG + 1) / 2>([&](auto i_) {
    constexpr int i = decltype(i_)::value;
    if constexpr (2 * i + 1 < DEG) {
      cp[P0 + i] = __byte_perm(y[2 * i], y[2 * i + 1], 0x6420);
    } else {
      cp[P0 + i] = __byte_perm(y[2 * i], 0u, 0x4420);
    }
  });
  __syncthreads(); // every warp has consumed the row's soft bits: they may be overwritten
  if (ZC > 0 || active) {
    h2_static_for<DEG>([&](auto k_) {
      constexpr int k = decltype(k_)::value;
      h2_write_col<ZC, G::col(E0 + k)>(soft_t, row_bytes, u[k]);
    });
  }
  __syncthreads(); // ... and the next row sees the new ones
}

// ---- prologue of a first transmission: deinterleaver into a natural-order int8 staging tile in shared memory ------------
// deinterleave (ldpc_rate_dematcher_impl.cpp:203-257): x[(E/Qm) j + i] = in[i Qm + j]; the walk over the circular buffer
// then skips the filler window (allot_llrs :128-201), i.e. stream index g lands at g (g < n_info) or g + F.
//
// Vector variant: one task = 4 consecutive symbols = 4 Qm bytes loaded with 64/128-bit loads; a byte transposition in
// registers (PRMT) turns them into Qm words of 4 consecutive stream positions each, stored with 32-bit stores.
// Needs Kq % 4 == 0, F % 4 == 0 and an 8-byte (Qm = 8, 4: 16-byte) aligned input; the caller checks.
__device__ __forceinline__ void h2_transpose4(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t (&o)[4])
{
  const uint32_t t0 = __byte_perm(a, b, 0x5140), t1 = __byte_perm(a, b, 0x7362);
  const uint32_t t2 = __byte_perm(c, d, 0x5140), t3 = __byte_perm(c, d, 0x7362);
  o[0] = __byte_perm(t0, t2, 0x5410), o[1] = __byte_perm(t0, t2, 0x7632);
  o[2] = __byte_perm(t1, t3, 0x5410), o[3] = __byte_perm(t1, t3, 0x7632);
}
template <int QM>
__device__ __forceinline__ void h2_task_words(const int8_t* __restrict__ src, uint32_t (&w)[QM])
{
  if constexpr (QM == 8 || QM == 4) {
#pragma unroll
    for (int r = 0; r < QM / 4; ++r) {
      const int4 v = __ldcs(reinterpret_cast<const int4*>(src) + r); // read once: do not displace the CRC tables in L1
      w[4 * r] = (uint32_t)v.x, w[4 * r + 1] = (uint32_t)v.y, w[4 * r + 2] = (uint32_t)v.z, w[4 * r + 3] = (uint32_t)v.w;
    }
  } else {
#pragma unroll
    for (int r = 0; r < QM / 2; ++r) {
      const int2 v = __ldcs(reinterpret_cast<const int2*>(src) + r);
      w[2 * r] = (uint32_t)v.x, w[2 * r + 1] = (uint32_t)v.y;
    }
  }
}
// Byte k of the 4 Qm-byte chunk held in w[].
template <int QM, int K0, int K1>
__device__ __forceinline__ uint32_t h2_pick2(const uint32_t (&w)[QM])
{
  return __byte_perm(w[K0 / 4], w[K1 / 4], (K0 % 4) | ((4 + K1 % 4) << 4));
}
template <int QM>
__device__ __forceinline__ void h2_task_transpose(const uint32_t (&w)[QM], uint32_t (&o)[QM])
{
  if constexpr (QM == 8) {
    uint32_t lo[4], hi[4];
    h2_transpose4(w[0], w[2], w[4], w[6], lo);
    h2_transpose4(w[1], w[3], w[5], w[7], hi);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      o[j] = lo[j], o[4 + j] = hi[j];
    }
  } else if constexpr (QM == 4) {
    h2_transpose4(w[0], w[1], w[2], w[3], o);
  } else if constexpr (QM == 2) {
    o[0] = __byte_perm(w[0], w[1], 0x6420), o[1] = __byte_perm(w[0], w[1], 0x7531);
  } else {
    h2_static_for<QM>([&](auto j_) {
      constexpr int j = decltype(j_)::value;
      o[j] = __byte_perm(h2_pick2<QM, j, QM + j>(w), h2_pick2<QM, 2 * QM + j, 3 * QM + j>(w), 0x5410);
    });
  }
}
template <int QM>
__device__ __forceinline__ void h2_deinterleave_vec(const pdc_dematch_plan& pl, const int8_t* __restrict__ in,
                                                    uint32_t stg_addr, int tid, int nth)
{
  const int F = pl.n_sys - pl.n_info, Kq = pl.E / QM, ntask = Kq >> 2;
  for (int k0 = tid; k0 < ntask; k0 += 2 * nth) { // two tasks in flight per thread
    const int k1   = k0 + nth;
    const bool has1 = k1 < ntask;
    uint32_t   w0[QM], w1[QM], o[QM];
    h2_task_words<QM>(in + (size_t)k0 * 4 * QM, w0);
    if (has1) {
      h2_task_words<QM>(in + (size_t)k1 * 4 * QM, w1);
    }
    h2_task_transpose<QM>(w0, o);
#pragma unroll
    for (int j = 0; j < QM; ++j) {
      const int g = j * Kq + 4 * k0;
      sts32(stg_addr + (uint32_t)((g < pl.n_info) ? g : (g + F)), o[j]);
    }
    if (has1) {
      h2_task_transpose<QM>(w1, o);
#pragma unroll
      for (int j = 0; j < QM; ++j) {
        const int g = j * Kq + 4 * k1;
        sts32(stg_addr + (uint32_t)((g < pl.n_info) ? g : (g + F)), o[j]);
      }
    }
  }
}
// Byte variant for inputs that miss the alignment conditions (odd fillers, Kq not a multiple of 4, Qm = 1).
template <int QM>
__device__ __forceinline__ void h2_deinterleave_bytes(const pdc_dematch_plan& pl, const int8_t* __restrict__ in,
                                                      int8_t* __restrict__ stg, int tid, int nth)
{
  const int F = pl.n_sys - pl.n_info, Kq = pl.E / QM;
  for (int i = tid; i < Kq; i += nth) {
#pragma unroll
    for (int j = 0; j < QM; ++j) {
      const int g                        = j * Kq + i;
      stg[(g < pl.n_info) ? g : (g + F)] = in[(size_t)i * QM + j];
    }
  }
}
template <int QM>
__device__ __forceinline__ void h2_deinterleave(const pdc_dematch_plan& pl, const int8_t* __restrict__ in, int8_t* stg,
                                                uint32_t stg_addr, int tid, int nth)
{
  if constexpr (QM == 1) {
    h2_deinterleave_bytes<1>(pl, in, stg, tid, nth);
  } else {
    const int       F     = pl.n_sys - pl.n_info, Kq = pl.E / QM;
    const uintptr_t align = (QM == 8 || QM == 4) ? 15 : 7;
    if (((Kq | F) & 3) == 0 && (reinterpret_cast<uintptr_t>(in) & align) == 0) {
      h2_deinterleave_vec<QM>(pl, in, stg_addr, tid, nth);
    } else {
      h2_deinterleave_bytes<QM>(pl, in, stg, tid, nth);
    }
  }
}

// int8 pair (low two bytes of w) -> clamped, biased half2 soft bits: 0x6400 | (x + 128) is the half 1152 + x; the clamp to
// +-64 of load_soft_bits (ldpc_decoder_impl.cpp:162-181) is applied on the biased value.
__device__ __forceinline__ uint32_t h2_soft_from_bytes(uint32_t w)
{
  const uint32_t p = __byte_perm(w ^ 0x8080u, 0x64646464u, 0x4140);
  return h2_max(h2_min(p, H2_BIAS_P64), H2_BIAS_N64);
}

// True if any of the K systematic soft bits is zero (hard_decision returns false then: log_likelihood_ratio.cpp:350-376).
// Uniform over the CTA (contains a barrier): every thread calls it.
template <int BG, int ZC, typename TAB>
__device__ __forceinline__ bool h2_any_zero_soft(TAB tab, uint32_t soft_t, bool active)
{
  uint32_t amin = H2_ONE;
  if (active) {
    h2_static_for<h2_graph<BG>::KB>([&](auto c_) {
      constexpr int  c = decltype(c_)::value;
      const uint32_t S = h2_read_natural<BG, ZC, c>(tab, soft_t);
      amin             = h2_min(amin, h2_abs(h2_sub(S, H2_BIAS)));
    });
  }
  return __syncthreads_or(((amin & 0x7fffu) == 0) || ((amin & 0x7fff0000u) == 0)) != 0;
}

template <int BG, int ZC>
__global__ void __launch_bounds__(192, 4) pdc_cb_kernel_h2(pdc_launch_params prm)
{
  using G = h2_graph<BG>;
  extern __shared__ __align__(16) unsigned char smem_raw[];

  const int        tid = threadIdx.x, nth = blockDim.x;
  const int        cbi = prm.order ? prm.order[blockIdx.x] : blockIdx.x;
  const pdc_dev_cb cb  = prm.cbs[cbi];
  if (cb.bg != BG || (ZC > 0 && (cb.Z != ZC || nth != ZC / 2))) {
    return; // the host launches one grid per (base graph, specialisation); a mismatch is a host bug: status stays untouched
  }
  pdc_geom g;
  pdc_geom_init(g, cb.bg, cb.Z);
  const int       Z = (ZC > 0) ? ZC : g.Z, W = Z >> 1, K = G::KB * Z;
  const bool      active  = (ZC > 0) || (tid < W);
  const bool      generic = (cb.flags & PDC_F_GENERIC) != 0;
  const uint32_t* tab     = c_h2_tab[cb.graph_idx];

  pdc_cb_result res;
  res.iterations = -1, res.crc_pass = 0, res.status = 0, res.nof_layers = 4;

  const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(smem_raw);
  const uint32_t row_bytes = 4u * (uint32_t)Z; // one column: 2 copies x Z halves
  const uint32_t soft_t    = smem_base + 4u * (uint32_t)(active ? tid : 0); // idle lanes read lane 0's words
  // After the soft rows: the natural-order int8 staging copy (prologue) / the ballot words (output).
  unsigned char* aux       = smem_raw + G::NCOL * row_bytes;
  const uint32_t aux_addr  = smem_base + G::NCOL * row_bytes;
  // Per-warp CRC partials live in the scratch area too (it is idle between the prologue and the output phase): with no static
  // shared memory four CTAs of Z = 384 need exactly 196 KB, which leaves the SM the larger L1 for the CRC tables.
  uint32_t* s_red = reinterpret_cast<uint32_t*>(aux);
  const int      nstage    = (G::NCOL - 2) * Z;


  // ---- prologue 1: rate dematching + HARQ combining into the soft buffer in HBM ------------------------------------------------
  const int8_t* dec_in = nullptr; // where load_soft_bits reads from (global memory), unless the staging tile holds it
  int           n_in   = 0;
  bool          from_smem = false;
  if (cb.flags & PDC_F_DECODE_ONLY) {
    dec_in = cb.llrs;
    n_in   = (int)cb.E;
    __syncthreads();
  } else {
    n_in = g.N;
    if (cb.flags & PDC_DEVF_FIRST_TX_TILE) {
      // The host planner (classify) established: new data, rv 0, no LBRM, one lap, n_info <= E and E + F <= nstage.
      pdc_dematch_plan pl;
      pl.N = ((BG == 1) ? 66 : 50) * Z, pl.n_sys = (G::KB - 2) * Z, pl.n_info = pl.n_sys - (int)cb.nof_filler;
      pl.E = (int)cb.E, pl.Qm = cb.Qm;
      const int end_pos = (int)cb.E + (int)cb.nof_filler;
      // First transmission from the start of the buffer, one lap, reaching past the systematic part: every position of the
      // N-entry soft buffer is a zero, a filler or exactly one LLR (a shorter one leaves [E, n_info) untouched, i.e. holding
      // whatever the buffer held before: allot_llrs :134-195), and everything that is not zero lies in the first nstage
      // entries. Build them in the natural-order staging tile, then stream tile + zeros to HBM (the HARQ state) with
      // 16-byte stores while the soft rows are loaded from the tile.
      int8_t* stg = reinterpret_cast<int8_t*>(aux);
      for (int q = pl.n_info + tid; q < pl.n_sys; q += nth) {
        if (q < end_pos) {
          stg[q] = PDC_LLR_INF; // filler bits (:162-166)
        }
      }
      for (int q = end_pos + tid; q < nstage; q += nth) {
        stg[q] = 0; // final zero (:198-200)
      }
      switch (pl.Qm) {
        case 1:
          h2_deinterleave<1>(pl, cb.llrs, stg, aux_addr, tid, nth);
          break;
        case 2:
          h2_deinterleave<2>(pl, cb.llrs, stg, aux_addr, tid, nth);
          break;
        case 4:
          h2_deinterleave<4>(pl, cb.llrs, stg, aux_addr, tid, nth);
          break;
        case 6:
          h2_deinterleave<6>(pl, cb.llrs, stg, aux_addr, tid, nth);
          break;
        default:
          h2_deinterleave<8>(pl, cb.llrs, stg, aux_addr, tid, nth);
          break;
      }
      __syncthreads();
      if ((reinterpret_cast<uintptr_t>(cb.soft) & 15) == 0 && ((nstage | pl.N) & 15) == 0) {
        int4*     dst = reinterpret_cast<int4*>(cb.soft);
        const int n16 = nstage >> 4, N16 = pl.N >> 4;
        // Streaming stores: the soft buffer is the HARQ state for a later slot, nothing in this launch reads it back.
        for (int q = tid; q < n16; q += nth) {
          __stcs(dst + q, reinterpret_cast<const int4*>(stg)[q]);
        }
        for (int q = n16 + tid; q < N16; q += nth) {
          __stcs(dst + q, make_int4(0, 0, 0, 0));
        }
      } else {
        for (int q = tid; q < nstage; q += nth) {
          cb.soft[q] = stg[q];
        }
        for (int q = nstage + tid; q < pl.N; q += nth) {
          cb.soft[q] = 0;
        }
      }
      from_smem = true;
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
      __syncthreads(); // orders the dematcher's writes before the reads below (same CTA)
    }
  }

  // ---- prologue 2: load_soft_bits (ldpc_decoder_impl.cpp:150-191) into the rotated, doubled half rows --------------------------
  if (active) {
    sts32(soft_t, H2_BIAS), sts32(soft_t + 2u * Z, H2_BIAS); // the two punctured columns: zero
    sts32(soft_t + row_bytes, H2_BIAS), sts32(soft_t + row_bytes + 2u * Z, H2_BIAS);
    if (ZC > 0 && from_smem) {
      // Staging tile -> rows with every rotation known at compile time: lanes (2 tid + sigma, + 1) mod Z of column c.
      h2_static_for<G::NCOL - 2>([&](auto c_) {
        constexpr int c     = decltype(c_)::value + 2;
        constexpr int sigma = (ZC > 0) ? h2_sigma<BG>(ZC > 0 ? ZC : 32, c) : 0;
        constexpr int Zc    = (ZC > 0) ? ZC : 32;
        const uint32_t base = aux_addr + (uint32_t)((c - 2) * Zc);
        uint32_t       w;
        if constexpr ((sigma & 1) == 0) {
          int la = 2 * tid + sigma;
          la -= (la >= Zc) ? Zc : 0;
          w = lds16(base + la);
        } else {
          int la = 2 * tid + sigma; // odd
          la -= (la >= Zc) ? Zc : 0;
          int lb = la + 1;          // even
          lb -= (lb >= Zc) ? Zc : 0;
          w = __byte_perm(lds16(base + la - 1), lds16(base + lb), 0x0041);
        }
        h2_write_col<ZC, c>(soft_t, row_bytes, h2_soft_from_bytes(w));
      });
    } else {
#pragma unroll 4
      for (int c = 2; c < G::NCOL; ++c) {
        const int sigma = (int)tab[PDC_H2_SIGMA_OFF + c];
        int       la    = 2 * tid + sigma;
        la -= (la >= Z) ? Z : 0;
        int lb = la + 1;
        lb -= (lb >= Z) ? Z : 0;
        const int pa = (c - 2) * Z + la, pb = (c - 2) * Z + lb;
        uint32_t  xa, xb;
        if (from_smem) {
          xa = aux[pa], xb = aux[pb];
        } else {
          xa = (pa < n_in) ? (uint8_t)dec_in[pa] : 0u;
          xb = (pb < n_in) ? (uint8_t)dec_in[pb] : 0u;
        }
        const uint32_t S = h2_soft_from_bytes(xa | (xb << 8));
        sts32(soft_t + c * row_bytes, S);
        sts32(soft_t + c * row_bytes + 2u * Z, S);
      }
    }
  }
  __syncthreads();

  const int       n_sig    = K - (int)cb.nof_filler;
  const bool      has_crc  = cb.crc_type != PDC_CRC_NONE;
  const bool      early    = (cb.flags & PDC_F_EARLY_STOP) && has_crc;
  const int       max_iter = cb.max_iter;
  const int       crc_order = (cb.crc_type == PDC_CRC16) ? 16 : 24;
  // x^j mod P table, zero for exponents below the CRC order (filler positions), padded in front: see build_crc_bit_tables().
  // Bit n contributes entry X - n, X = PAD + n_sig - 1 + order; this thread's lanes (2 tid, 2 tid + 1) of column c need
  // entries X - c Z - 2 tid - 1 and X - c Z - 2 tid: one aligned 64-bit load from the copy whose shift makes X - 1 even.
  const uint2* crc_pair = nullptr;
  if (has_crc) {
    const int X = PDC_CRC_BIT_PAD + n_sig - 1 + crc_order;
    const int v = (X - 1) & 1;
    crc_pair    = reinterpret_cast<const uint2*>(prm.crc_bit + (size_t)(2 * (cb.crc_type - 1) + v) * PDC_CRC_BIT_STRIDE +
                                              (v + X - 1 - 2 * tid));
  }

  uint32_t cp[h2_pack<BG>::NREG]; // check-to-variable messages, all zero (byte 128) before the first iteration
#pragma unroll
  for (int e = 0; e < h2_pack<BG>::NREG; ++e) {
    cp[e] = 0x80808080u;
  }

  int  it_done = 0;
  bool success = false;
  for (int it = 0; it < max_iter && !success; ++it) {
    h2_row<BG, ZC, 0>(cp, tab, soft_t, row_bytes, active, generic);
    h2_row<BG, ZC, 1>(cp, tab, soft_t, row_bytes, active, generic);
    h2_row<BG, ZC, 2>(cp, tab, soft_t, row_bytes, active, generic);
    h2_row<BG, ZC, 3>(cp, tab, soft_t, row_bytes, active, generic);
    it_done = it + 1;
    const bool last = (it_done == max_iter);
    if (!has_crc || (!early && !last)) {
      continue;
    }
    // ---- hard decision + CRC (ldpc_decoder_impl.cpp:127-135, log_likelihood_ratio.cpp:350-376). The CRC is linear: it is
    //      the xor of x^(n_sig - 1 - n + order) mod P over the set bits n, so every thread folds its own lane pairs, in
    //      natural order, straight from the soft rows; a zero soft bit among the K systematic ones refuses the early stop.
    uint32_t acc = 0;
    if (active) {
      h2_static_for<G::KB>([&](auto c_) {
        constexpr int  c  = decltype(c_)::value;
        const uint32_t S  = h2_read_natural<BG, ZC, c>(tab, soft_t); // natural lanes (2 tid, 2 tid + 1) of column c
        const uint2 tp = __ldg(crc_pair - c * (Z >> 1)); // .x: lane 2 tid + 1, .y: lane 2 tid
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
    // pusch_codeblock_decoder.cpp:62-73: without early stop there is one CRC after the last iteration and zero soft bits
    // do not matter. With early stop a zero among the K systematic soft bits refuses the stop: that second sweep runs only
    // when the checksum matches, i.e. once per code block instead of once per iteration.
    success = (crc == 0u);
    if (success && early) {
      success = !h2_any_zero_soft<BG, ZC>(tab, soft_t, active);
    }
  }

  // ---- output ------------------------------------------------------------------------------------------------------------------
  __syncthreads();
  h2_emit_bits<BG, ZC>(tab, soft_t, reinterpret_cast<uint32_t*>(aux), active, tid, nth, Z, K, reinterpret_cast<uint32_t*>(cb.out));
  if (success) {
    res.iterations = early ? it_done : max_iter;
    res.crc_pass   = 1;
  }
  if (tid == 0) {
    *cb.res = res;
  }
}

// Every even lifting size: a thread owns the lane pair (2 t, 2 t + 1). The 7 odd sizes (3 .. 15) stay with the lane kernel.
extern "C" int pdc_h2_supported(int bg, int Z)
{
  return (bg == 1 || bg == 2) && (Z % 2 == 0) && Z >= 2 && Z <= PDC_MAX_Z;
}
extern "C" size_t pdc_h2_smem_bytes(int bg, int Z, int)
{
  const int    ncol = (bg == 1) ? 26 : 14, kb = (bg == 1) ? 22 : 10;
  const size_t aux  = (size_t)(ncol - 2) * Z > (size_t)kb * 48 ? (size_t)(ncol - 2) * Z : (size_t)kb * 48;
  return (size_t)ncol * 4 * Z + aux; // soft rows + (staging copy of the prologue | ballot words of the output)
}
// zc: 0 = lifting sizes taken at run time (mixed list), 384 = every code block of the list has Z = 384.
extern "C" cudaError_t pdc_launch_h2(const pdc_launch_params& prm, int bg, int zc, int block, size_t smem,
                                     cudaStream_t stream)
{
  if (zc == 384 && block != 192) {
    return cudaErrorInvalidValue;
  }
  if (bg == 1) {
    if (zc == 384) {
      pdc_cb_kernel_h2<1, 384><<<prm.n, block, smem, stream>>>(prm);
    } else {
      pdc_cb_kernel_h2<1, 0><<<prm.n, block, smem, stream>>>(prm);
    }
  } else {
    if (zc == 384) {
      pdc_cb_kernel_h2<2, 384><<<prm.n, block, smem, stream>>>(prm);
    } else {
      pdc_cb_kernel_h2<2, 0><<<prm.n, block, smem, stream>>>(prm);
    }
  }
  return cudaGetLastError();
}

// Relative shifts of the lazily rotated columns, for all 2 x 51 (bg, Z); uploaded once per device.
static cudaError_t pdc_h2_init_device()
{
  static uint32_t host_tab[2 * PDC_NOF_LIFTING_SIZES][PDC_H2_TAB_STRIDE];
  static const int sizes[PDC_NOF_LIFTING_SIZES] = {2,   3,   4,   5,   6,   7,   8,   9,   10,  11,  12,  13,  14,
                                                   15,  16,  18,  20,  22,  24,  26,  28,  30,  32,  36,  40,  44,
                                                   48,  52,  56,  60,  64,  72,  80,  88,  96,  104, 112, 120, 128,
                                                   144, 160, 176, 192, 208, 224, 240, 256, 288, 320, 352, 384};
  static const int base[8] = {2, 3, 5, 7, 9, 11, 13, 15};
  for (int bg = 0; bg != 2; ++bg) {
    const uint8_t* deg  = bg ? pdc_bg2_row_deg : pdc_bg1_row_deg;
    const uint8_t* cols = bg ? pdc_bg2_edge_col : pdc_bg1_edge_col;
    const int      ncol = bg ? 14 : 26;
    for (int zi = 0; zi != PDC_NOF_LIFTING_SIZES; ++zi) {
      const int Z   = sizes[zi];
      int       ils = 0;
      for (int i = 0; i != 8; ++i) {
        int z = base[i];
        while (z < Z) {
          z *= 2;
        }
        ils = (z == Z) ? i : ils;
      }
      const uint16_t* sh  = bg ? pdc_bg2_edge_shift[ils] : pdc_bg1_edge_shift[ils];
      uint32_t*       dst = host_tab[bg * PDC_NOF_LIFTING_SIZES + zi];
      int             row_of[80], ne = 0;
      for (int m = 0; m != 4; ++m) {
        for (int k = 0; k != deg[m]; ++k) {
          row_of[ne++] = m;
        }
      }
      // Alignment of a column after a full iteration = shift of the last core row that holds it.
      int sigma[26];
      for (int c = 0; c != ncol; ++c) {
        sigma[c] = 0;
        for (int e = 0; e != ne; ++e) {
          if (cols[e] == c) {
            sigma[c] = sh[e] % Z;
          }
        }
        dst[PDC_H2_SIGMA_OFF + c] = (uint32_t)sigma[c];
        // Natural order seen through the doubled row: lane l sits at half index l + (Z - sigma) mod Z.
        const int      nat = (Z - sigma[c]) % Z;
        const uint32_t no  = (uint32_t)c * 4u * (uint32_t)Z + 4u * (uint32_t)(nat >> 1);
        dst[PDC_H2_NAT_OFF + c] = no | (((nat & 1) ? 0x5432u : 0x3210u) << 16);
      }
      for (int e = 0; e != ne; ++e) {
        // Previous writer of this column: the nearest earlier row holding it, cyclically (previous iteration included).
        int prev = sigma[cols[e]];
        for (int e2 = 0; e2 != e; ++e2) {
          if (cols[e2] == cols[e] && row_of[e2] < row_of[e]) {
            prev = sh[e2] % Z;
          }
        }
        const int rel = ((sh[e] % Z) - prev + Z) % Z;
        const uint32_t off = (uint32_t)cols[e] * 4u * (uint32_t)Z + 4u * (uint32_t)(rel >> 1);
        const uint32_t sel = (rel & 1) ? 0x5432u : 0x3210u;
        dst[e]             = off | (sel << 16);
      }
    }
  }
  return cudaMemcpyToSymbol(c_h2_tab, host_tab, sizeof(host_tab));
}
