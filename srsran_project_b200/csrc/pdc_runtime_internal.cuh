// Host runtime behind the C ABI of include/pusch_dec_cuda.h: one context per GPU with a small ring of "tickets"
// (stream + event + pinned staging + device staging), the device-resident HARQ soft-buffer arena, and the constant
// tables (lifted base graphs, CRC powers). No CPU fallback: without a CUDA device pdc_create fails.
//
// Role in the reference's terms: this is what a bbdev-style accelerator driver is to
// lib/hal/phy/upper/channel_processors/pusch (enqueue = pdc_submit, dequeue = pdc_poll/pdc_fetch, HARQ context =
// arena slots; hw_accelerator_pusch_dec.h:83-115, ext_harq_buffer_context_repository.h:36-107).
#include "ldpc_bg_tables.h"
#include "pdc_device.cuh"
#include <chrono>
#include <nvtx3/nvToolsExt.h> // header-only: ranges cost nothing unless a profiler is attached
#include <atomic>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

extern "C" cudaError_t pdc_launch_lane(const pdc_launch_params& prm, int block, size_t smem, cudaStream_t stream);
extern "C" cudaError_t pdc_launch_h2(const pdc_launch_params& prm, int bg, int zc, int block, size_t smem,
                                     cudaStream_t stream);
extern "C" cudaError_t pdc_kernels_init_device(const pdc_bg_rows* rows);
extern "C" cudaError_t pdc_probe_run(int mode, int sm_count, cudaStream_t stream, double* thread_inst_per_s);
extern "C" int         pdc_h2_supported(int bg, int Z);
extern "C" size_t      pdc_h2_smem_bytes(int bg, int Z, int max_layers);
extern "C" size_t      pdc_h2g_smem_bytes(int bg, int Z, int max_layers);
extern "C" cudaError_t pdc_launch_h2g(const pdc_launch_params& prm, int bg, int block, size_t smem, cudaStream_t stream);
extern "C" cudaError_t pdc_enc_init_device();
extern "C" cudaError_t pdc_enc_launch(const pdc_dev_enc_tb* d_tbs, uint32_t ntb, uint32_t max_cbs, const uint32_t* graphs,
                                      const uint32_t* crc_pow, size_t smem, cudaStream_t stream);
extern "C" cudaError_t pdc_demod_init_device();
extern "C" bool        pdc_prg_states(uint32_t c_init, uint32_t pos, uint32_t* s1, uint32_t* s2);
extern "C" cudaError_t pdc_demux_launch(const pdc_dev_demux_sym* d_syms, uint32_t nplain, uint32_t nuci, bool placeholders,
                                        cudaStream_t stream);
extern "C" cudaError_t pdc_demod_launch(const pdc_demod_seg* d_segs, const void* d_chunks, uint32_t nchunks,
                                        const float* d_symbols, const float* d_noise, int8_t* d_llrs, cudaStream_t stream);
extern "C" cudaError_t pdc_launch_upload(void* dst0, const void* src0, size_t bytes0, void* dst1, const void* src1, size_t bytes1,
                                         cudaStream_t stream);
extern "C" cudaError_t pdc_launch_tb(const pdc_dev_tb* d_tbs, const pdc_dev_cb* d_cbs, const uint32_t* crc_pow, int ntb,
                                     int max_cbs, cudaStream_t stream);

namespace {

// NVTX range for the lifetime of a scope: the CUDA counterpart of the trace events the reference emits around its code-block
// tasks (pusch_decoder_impl.cpp:332,381 "cb_decode"); visible in Nsight Systems timelines next to the kernels and copies.
struct nvtx_scope {
  explicit nvtx_scope(const char* name) { nvtxRangePushA(name); }
  ~nvtx_scope() { nvtxRangePop(); }
};

struct ticket_t {
  cudaStream_t stream   = nullptr;
  cudaEvent_t  done     = nullptr;
  cudaEvent_t  staged   = nullptr; // pdc_decode_device*: the descriptor copies out of the pinned staging have been consumed
  bool         in_use   = false;
  int64_t      id       = -1;
  uint32_t     n        = 0;
  // pinned host staging
  pdc_dev_cb*    h_cbs  = nullptr;
  int8_t*        h_llrs = nullptr;
  uint8_t*       h_bits = nullptr;
  pdc_cb_result* h_res  = nullptr;
  int8_t*        h_soft = nullptr;
  uint32_t*      h_order = nullptr;
  size_t         soft_cap = 0;
  // device staging
  pdc_dev_cb*    d_cbs  = nullptr;
  int8_t*        d_llrs = nullptr;
  uint8_t*       d_bits = nullptr;
  pdc_cb_result* d_res  = nullptr;
  int8_t*        d_soft = nullptr;
  uint32_t*      d_order = nullptr;
  // transport-block path (pdc_submit_tbs / pdc_decode_device_tbs)
  pdc_dev_tb*    h_tbs = nullptr, *d_tbs = nullptr;
  uint32_t*      d_terms = nullptr;
  uint8_t*       h_tbout = nullptr, *d_tbout = nullptr;
  pdc_tb_result* h_tbres = nullptr, *d_tbres = nullptr;
  uint32_t       ntb = 0;
  std::vector<pdc_tb_desc> tbs;
  // what to hand back
  std::vector<pdc_cb_desc> descs;
  std::vector<uint8_t>     lst; // kernel list of every code block (planner scratch)
  std::vector<int32_t>     nz;  // planner scratch: new non-zero extents of the arena slots (batch_buffers::nz_scratch)
  std::vector<int32_t>     zero_slots; // planner scratch: released arena slots to clear before the kernels
  uint32_t                 nof_zero_slots = 0;
  size_t                   bits_bytes = 0, soft_bytes = 0;
  bool                     soft_back  = false, bits_dense = false;
  bool                     borrowed   = false; // staging lent to pdc_decode_device: its copy may still be in flight
  // PDC_TRACE=1: device-side timeline of a submission (events with timing), host clock at the end of the submit call
  cudaEvent_t                           tr[4] = {nullptr, nullptr, nullptr, nullptr};
  std::chrono::steady_clock::time_point submitted_at;
};

} // namespace

// A staging arena (pdc_stage_*): page-locked host memory mirrored 1:1 by device memory, with its own copy stream. Regions of it
// are pushed to the device as soon as their owner has filled them, long before the submission that uses them is formed.
struct pdc_stage {
  pdc_ctx*     ctx    = nullptr;
  int8_t*      h_base = nullptr;
  int8_t*      d_base = nullptr;
  size_t       bytes  = 0;
  cudaStream_t copy   = nullptr;
  cudaEvent_t  pushed = nullptr; // recorded behind the latest push at submission time
  std::mutex   mtx;
};

struct pdc_ctx {
  pdc_config            cfg;
  std::mutex            mtx;
  std::mutex            err_mtx; // submit / fetch run outside mtx: the error string has its own lock
  std::string           err;
  bool                  zero_copy = false; // pdc_set_zero_copy
  bool                  trace     = false; // PDC_TRACE=1 in the environment at pdc_create: timeline summary on stderr at pdc_destroy
  double                trace_ms[5] = {0, 0, 0, 0, 0}; // descs H2D, kernels, D2H, submit -> done seen by the host, count
  std::vector<ticket_t> tickets;
  int64_t               next_id = 1;
  unsigned              next_ticket = 0; // reserve_ticket starts its search here (round robin)
  // A mixed batch runs one kernel per list (launch_batch): the lists are independent, so all but the first go to side streams
  // that fork from and join the caller's stream through events (irregular batches are bound by the slowest code block of each
  // kernel, not by the machine: in sequence those critical paths add up).
  static constexpr int  NOF_SIDE_STREAMS = 6;
  std::mutex            side_mtx;
  cudaStream_t          side_stream[NOF_SIDE_STREAMS] = {};
  cudaEvent_t           side_fork = nullptr, side_join[NOF_SIDE_STREAMS] = {};
  uint32_t*             d_graphs = nullptr;
  uint32_t*             d_crc    = nullptr;
  uint32_t*             d_crc_bit = nullptr;
  int8_t*               d_harq   = nullptr;
  std::vector<int32_t>  harq_nz;            // per arena slot: bound on the extent of non-zero soft bits (batch_buffers::harq_nz)
  std::vector<uint8_t>  harq_released;      // per arena slot: released, reads as zeros, memory not cleared yet (pdc_harq_release)
  uint8_t*              d_harq_bits = nullptr; // decoded bits kept with every HARQ slot (transport-block path), PDC_MAX_K / 8 each
  std::atomic<uint64_t> launches{0};
  size_t                bits_stride_total = 0;
  // LLR producers (pdc_demodulate*): a small ring of descriptor staging (segments + chunk table), stream-ordered through one
  // event per entry, and grow-only device buffers for the host-buffer entry point.
  struct demod_slot {
    void*       d_desc = nullptr;
    size_t      cap    = 0;
    void*       h_desc = nullptr; // pinned image of d_desc: a pageable source would make the upload wait for the stream
    size_t      h_cap  = 0;
    cudaEvent_t done   = nullptr;
    bool        used   = false;
  };
  demod_slot   demod_ring[8];
  unsigned     demod_next = 0;
  std::mutex   demod_mtx, demod_call_mtx;
  float*       d_demod_in = nullptr; // symbols, then noise variances
  int8_t*      d_demod_out = nullptr;
  size_t       demod_in_cap = 0, demod_out_cap = 0;
  cudaStream_t demod_stream = nullptr;
  // UL-SCH demultiplexer (pdc_ulsch_demultiplex*): same arrangement.
  demod_slot   demux_ring[8];
  unsigned     demux_next = 0;
  int8_t*      d_demux_in = nullptr, *d_demux_out = nullptr;
  size_t       demux_in_cap = 0, demux_out_cap = 0;
  // PDSCH encoder (pdc_pdsch_encode*): same arrangement.
  demod_slot   enc_ring[8];
  unsigned     enc_next = 0;
  uint8_t*     d_enc_in = nullptr, *d_enc_out = nullptr;
  size_t       enc_in_cap = 0, enc_out_cap = 0;
};

namespace {

std::mutex  g_static_err_mtx;
std::string g_static_err; // last error of a call without a context (pdc_create)

bool cuda_ok(pdc_ctx* ctx, cudaError_t e, const char* what)
{
  if (e == cudaSuccess) {
    return true;
  }
  char buf[256];
  snprintf(buf, sizeof(buf), "%s: %s", what, cudaGetErrorString(e));
  if (ctx) {
    std::lock_guard<std::mutex> lock(ctx->err_mtx);
    ctx->err = buf;
  }
  return false;
}

#define CK(call)                                                                                                       \
  do {                                                                                                                 \
    if (!cuda_ok(ctx, (call), #call)) {                                                                                \
      return PDC_ERR_CUDA;                                                                                             \
    }                                                                                                                  \
  } while (0)

int lifting_set_index(int Z)
{
  static const int base[8] = {2, 3, 5, 7, 9, 11, 13, 15};
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

// Lifted graphs for all 2 x 51 (bg, Z): entry = col | (V mod Z) << 8 (ldpc_graph_impl.cpp:29-57 does the same reduction).
void build_graph_tables(std::vector<uint32_t>& out, pdc_bg_rows& rows)
{
  out.assign((size_t)2 * PDC_NOF_LIFTING_SIZES * PDC_GRAPH_STRIDE, 0);
  memset(&rows, 0, sizeof(rows));
  static const int sizes[PDC_NOF_LIFTING_SIZES] = {2,   3,   4,   5,   6,   7,   8,   9,   10,  11,  12,  13,  14,
                                                   15,  16,  18,  20,  22,  24,  26,  28,  30,  32,  36,  40,  44,
                                                   48,  52,  56,  60,  64,  72,  80,  88,  96,  104, 112, 120, 128,
                                                   144, 160, 176, 192, 208, 224, 240, 256, 288, 320, 352, 384};
  for (int bg = 0; bg != 2; ++bg) {
    const uint8_t* deg   = bg ? pdc_bg2_row_deg : pdc_bg1_row_deg;
    const uint8_t* col   = bg ? pdc_bg2_edge_col : pdc_bg1_edge_col;
    int            M     = bg ? PDC_BG2_ROWS : PDC_BG1_ROWS;
    int            E     = bg ? PDC_BG2_EDGES : PDC_BG1_EDGES;
    int            acc   = 0;
    for (int m = 0; m != M; ++m) {
      rows.start[bg][m] = (uint16_t)acc;
      rows.deg[bg][m]   = deg[m];
      acc += deg[m];
    }
    rows.start[bg][M] = (uint16_t)acc;
    for (int zi = 0; zi != PDC_NOF_LIFTING_SIZES; ++zi) {
      int             Z   = sizes[zi];
      int             ils = lifting_set_index(Z);
      const uint16_t* sh  = bg ? pdc_bg2_edge_shift[ils] : pdc_bg1_edge_shift[ils];
      uint32_t*       dst = &out[((size_t)bg * PDC_NOF_LIFTING_SIZES + zi) * PDC_GRAPH_STRIDE];
      for (int e = 0; e != E; ++e) {
        dst[e] = (uint32_t)col[e] | ((uint32_t)(sh[e] % Z) << 8);
      }
    }
  }
}

// x^j mod P for j < PDC_CRC_TABLE_LEN, for CRC16 / CRC24A / CRC24B (crc_calculator_generic_impl.cpp:31-56).
void build_crc_tables(std::vector<uint32_t>& out)
{
  static const uint32_t poly[3]  = {0x11021, 0x1864CFB, 0x1800063};
  static const int      order[3] = {16, 24, 24};
  out.assign((size_t)3 * PDC_CRC_TABLE_LEN, 0);
  for (int c = 0; c != 3; ++c) {
    uint32_t top = 1u << order[c];
    uint32_t v   = 1;
    for (int j = 0; j != PDC_CRC_TABLE_LEN; ++j) {
      out[(size_t)c * PDC_CRC_TABLE_LEN + j] = v;
      v <<= 1;
      if (v & top) {
        v ^= poly[c];
      }
    }
  }
}

// Per-bit CRC table of the packed kernel: entry PAD + j = x^j mod P for j >= order and 0 below, so that the bit at
// position n of an n_sig-bit message adds entry PAD + (n_sig - 1 - n + order) and filler positions (n >= n_sig) add 0.
void build_crc_bit_tables(const std::vector<uint32_t>& pow, std::vector<uint32_t>& out)
{
  static const int order[3] = {16, 24, 24};
  out.assign((size_t)3 * 2 * PDC_CRC_BIT_STRIDE, 0);
  for (int c = 0; c != 3; ++c) {
    for (int v = 0; v != 2; ++v) {
      for (int j = order[c]; j < PDC_CRC_TABLE_LEN; ++j) {
        out[(size_t)(2 * c + v) * PDC_CRC_BIT_STRIDE + v + PDC_CRC_BIT_PAD + j] = pow[(size_t)c * PDC_CRC_TABLE_LEN + j];
      }
    }
  }
}

// pdc_lifting_pos through a table (the planner calls it for every class of a batch).
inline int lifting_pos_of(unsigned Z)
{
  struct table {
    int8_t pos[PDC_MAX_Z + 1];
    table()
    {
      for (int z = 0; z <= PDC_MAX_Z; ++z) {
        pos[z] = (int8_t)pdc_lifting_pos(z);
      }
    }
  };
  static const table t;
  return (Z <= PDC_MAX_Z) ? t.pos[Z] : -1;
}

// Validates one descriptor the way the reference's asserts would and fills the derived fields. pl_out / zpos_out: the
// dematcher's plan (valid unless the descriptor is decode-only) and the position of Z in the list of lifting sizes, for callers
// that go on with them.
bool validate_desc(const pdc_cb_desc& d, pdc_geom& g, int& max_layers, pdc_dematch_plan* pl_out = nullptr, int* zpos_out = nullptr)
{
  const int zpos = lifting_pos_of(d.Z);
  if (!pdc_geom_init(g, d.bg, d.Z) || zpos < 0) {
    return false;
  }
  if (zpos_out) {
    *zpos_out = zpos;
  }
  if (d.max_iter == 0 || d.crc_type > PDC_CRC24B) {
    return false;
  }
  if ((uint64_t)d.nof_filler >= (uint64_t)g.K) {
    return false;
  }
  if (d.flags & PDC_F_DECODE_ONLY) {
    // ldpc_decoder_impl.cpp:64-80: (Kb + 2) Z <= input <= N_short Z.
    if ((int)d.E < g.K + 2 * g.Z || (int)d.E > g.N) {
      return false;
    }
    max_layers = pdc_layers_for_nz(g, (int)d.E);
    return true;
  }
  pdc_dematch_plan pl;
  if (d.E > 35u * PDC_MAX_K || d.Nref > (uint32_t)PDC_MAX_N ||
      !pdc_dematch_plan_init(pl, g, (int)d.E, d.Qm, d.rv, (int)d.Nref, (int)d.nof_filler,
                             (d.flags & PDC_F_NEW_DATA) != 0)) {
    return false;
  }
  max_layers = pdc_layers_for_nz(g, pdc_dematch_nz_bound(pl));
  if (pl_out) {
    *pl_out = pl;
  }
  return true;
}

size_t lane_smem_bytes(const pdc_geom& g, int max_layers)
{
  int bgi   = (g.Kb == 22) ? 0 : 1;
  int M     = bgi ? PDC_BG2_ROWS : PDC_BG1_ROWS;
  int L     = max_layers > M ? M : max_layers;
  int edges = 0;
  const uint8_t* deg = bgi ? pdc_bg2_row_deg : pdc_bg1_row_deg;
  for (int m = 0; m != L; ++m) {
    edges += deg[m];
  }
  return (size_t)(((g.Kb + L) * g.Z + 15) / 16 * 16) + (size_t)edges * g.Z;
}

#define PDC_NOF_PACKED_LISTS 6
#define PDC_NOF_LISTS (PDC_NOF_PACKED_LISTS + 1)
#define PDC_LIST_SKIP 255 /* PDC_F_DECODED: no kernel touches the code block */
#define PDC_SLOT_BITS (PDC_MAX_K / 8)

struct batch_plan {
  // Packed-kernel lists, four core rows: [0] BG1 any Z, [1] BG2 any Z, [2] BG1 Z = 384, [3] BG2 Z = 384 (compile-time
  // specialisations); any number of rows (pdc_kernel_h2g.cuh): [4] BG1, [5] BG2. Then the lane kernel.
  size_t   smem_lane = 0, smem_h2[PDC_NOF_PACKED_LISTS] = {0, 0, 0, 0, 0, 0};
  int      block_lane = 32, block_h2[PDC_NOF_PACKED_LISTS] = {32, 32, 32, 32, 32, 32};
  uint32_t n_lane = 0, n_h2[PDC_NOF_PACKED_LISTS] = {0, 0, 0, 0, 0, 0}; // order array: packed lists, then lane-kernel CBs
  bool     identity = false;                   // one list holds every code block: launch without the order array
  size_t   bits_end = 0, soft_end = 0;         // extent of the packed bits / host-exchanged soft buffers
  bool     bits_dense = false, any_host_soft = false;
};

// Which code blocks the packed half2 kernels take: Z even; the four core rows only (pdc_kernel_h2.cuh) or any
// number of rows as long as the shared memory fits (pdc_kernel_h2g.cuh).
bool takes_h2(const pdc_cb_desc& d, int max_layers, bool general)
{
  if (!pdc_h2_supported(d.bg, d.Z) || d.nof_filler > PDC_CRC_BIT_PAD - 32) {
    return false;
  }
  if (general ? (max_layers <= 4 || pdc_h2g_smem_bytes(d.bg, d.Z, max_layers) > (size_t)PDC_MAX_DYN_SMEM) : (max_layers != 4)) {
    return false;
  }
  if (d.flags & (PDC_F_DEMATCH_ONLY | PDC_F_FORCE_DECODING)) {
    return false;
  }
  if ((d.flags & PDC_F_DECODE_ONLY) && (d.E % d.Z) != 0) {
    return false;
  }
  return true;
}

// Everything that is derived from the first 20 bytes of a descriptor (bg .. nof_filler). The code blocks of a transport
// block differ only in E (two values), their offsets and their HARQ slot, so the planner recomputes a class only when those
// 20 bytes change: the per-code-block host cost is a compare and a 56-byte copy.
struct desc_class {
  pdc_cb_desc key;
  pdc_dev_cb  proto; // pointers zero
  size_t      K_bytes4 = 0; // packed message size rounded up to words
  uint32_t    N        = 0;
  int         list     = 0; // packed-kernel lists, then PDC_NOF_PACKED_LISTS = lane kernel
  bool        valid    = false;
  // What the dematcher of this class does to the extent of non-zero soft bits of a HARQ arena slot (see plan_batch):
  // track 0: nothing (decode only / already decoded); 1: new data, one lap: [0, end_idx) written, [tail_start, N) zeroed,
  // [end_idx, tail_start) left as it was (empty without limited-buffer rate matching); 2: anything up to N may be non-zero.
  int  track = 0, end_idx = 0, tail_start = 0;
  // Limited-buffer first transmission that fits the packed kernel's staging tile whenever the stretch it leaves untouched is
  // known to hold zeros: then it IS a first transmission into an all-zero buffer and takes the four-row kernel (list_fast).
  bool   tile_if_clean = false;
  int    list_fast = 0, block_fast = 0;
  size_t smem_fast = 0;
};

inline bool same_class(const pdc_cb_desc& a, const pdc_cb_desc& b)
{
  static_assert(offsetof(pdc_cb_desc, harq_slot) == 20, "class key = the first 20 bytes of pdc_cb_desc");
  return memcmp(&a, &b, 20) == 0;
}

bool classify(const pdc_cb_desc& d, desc_class& cls, batch_plan& bp)
{
  pdc_geom         g;
  pdc_dematch_plan pl;
  int              ml = 0, zpos = -1;
  cls.valid           = false;
  if (d.flags & PDC_F_DECODED) {
    // Only the geometry matters: the code block is not processed again.
    zpos = lifting_pos_of(d.Z);
    if (!pdc_geom_init(g, d.bg, d.Z) || zpos < 0 || d.crc_type == PDC_CRC_NONE || d.crc_type > PDC_CRC24B ||
        (uint64_t)d.nof_filler >= (uint64_t)g.K) {
      return false;
    }
  } else if (!validate_desc(d, g, ml, &pl, &zpos)) {
    return false;
  }
  cls.key       = d;
  pdc_dev_cb& c = cls.proto;
  memset(&c, 0, sizeof(c));
  c.E = d.E, c.Nref = d.Nref, c.nof_filler = d.nof_filler, c.Z = d.Z;
  c.bg = d.bg, c.Qm = d.Qm, c.rv = d.rv, c.crc_type = d.crc_type, c.max_iter = d.max_iter;
  c.flags = d.flags & (uint8_t)~PDC_DEVF_FIRST_TX_TILE;
  cls.track = 0, cls.tile_if_clean = false;
  if (!(d.flags & (PDC_F_DECODED | PDC_F_DECODE_ONLY))) {
    { // pl: the plan validate_desc derived
      cls.track = 2;
      if (pl.final_zero) { // new data, one lap, ends before Ncb
        cls.track      = 1;
        cls.end_idx    = pl.end_idx;
        cls.tail_start = pl.N - (pl.Ncb - pl.end_idx);
      }
      const int  end_pos = pl.final_zero ? pl.end_idx : pl.N;
      const bool fits    = pl.new_data && pl.rounds == 1 && pl.k0 == 0 && pl.E >= pl.n_info && end_pos <= (g.Kb + 2) * g.Z &&
                        !(d.flags & PDC_F_DEMATCH_ONLY);
      if (fits && pl.Ncb == pl.N) {
        c.flags |= PDC_DEVF_FIRST_TX_TILE; // then end_pos == E + F (see pdc_dematch_plan_init with k0 = 0, one lap)
      } else if (fits && pl.final_zero && takes_h2(d, 4, false)) {
        // Ncb < N (limited-buffer rate matching, what the live stack always asks for: pusch_processor_impl.cpp:327-334): the
        // reference leaves [end_idx, N - (Ncb - end_idx)) as it was (allot_llrs :198-200 zeroes the LAST Ncb - end_idx entries
        // of the N-long buffer). The decoder sees that stretch, so in general up to 37 layers may take part; when the arena
        // slot is known to hold zeros there, the result is the one of a first transmission into an empty buffer.
        cls.tile_if_clean = true;
        cls.list_fast     = (d.bg - 1) + ((d.Z == 384) ? 2 : 0);
        cls.smem_fast     = pdc_h2_smem_bytes(d.bg, d.Z, 4);
        cls.block_fast    = (d.Z / 2 + 31) / 32 * 32;
      }
    }
  }
  c.graph_idx  = (uint16_t)((d.bg - 1) * PDC_NOF_LIFTING_SIZES + zpos);
  c.max_layers = (uint16_t)ml;
  cls.K_bytes4 = (size_t)((g.K + 31) / 32) * 4;
  cls.N        = (uint32_t)g.N;
  const bool core_only = takes_h2(d, ml, false), general = !core_only && takes_h2(d, ml, true);
  if (d.flags & PDC_F_DECODED) {
    cls.list = PDC_LIST_SKIP;
  } else if (core_only || general) {
    int k    = general ? (4 + (d.bg - 1)) : ((d.bg - 1) + ((d.Z == 384) ? 2 : 0));
    cls.list = k;
    size_t s = general ? pdc_h2g_smem_bytes(d.bg, d.Z, ml) : pdc_h2_smem_bytes(d.bg, d.Z, ml);
    bp.smem_h2[k]  = s > bp.smem_h2[k] ? s : bp.smem_h2[k];
    int blk        = (d.Z / 2 + 31) / 32 * 32;
    bp.block_h2[k] = blk > bp.block_h2[k] ? blk : bp.block_h2[k];
  } else {
    cls.list      = PDC_NOF_PACKED_LISTS;
    size_t s      = (d.flags & PDC_F_DEMATCH_ONLY) ? 0 : lane_smem_bytes(g, ml);
    bp.smem_lane  = s > bp.smem_lane ? s : bp.smem_lane;
    int blk       = (d.Z + 31) / 32 * 32;
    bp.block_lane = blk > bp.block_lane ? blk : bp.block_lane;
  }
  cls.valid = true;
  return true;
}

// Where the buffers of a batch live on the device (ticket staging for pdc_submit, caller's pointers for pdc_decode_device).
struct batch_buffers {
  const int8_t*  d_llrs = nullptr;
  int8_t*        d_soft = nullptr; // soft buffers exchanged with the host (harq_slot < 0); may be null if none is needed
  uint8_t*       d_bits = nullptr;
  pdc_cb_result* d_res  = nullptr;
  int8_t*        d_harq = nullptr;
  uint32_t       harq_slots = 0;
  size_t         llr_bytes  = (size_t)-1; // bound for llr_offset + E
  uint8_t*       d_slot_bits = nullptr;   // transport-block path: decoded bits of CBs with a HARQ slot stay with the slot
  // Host-side bookkeeping of the arena: harq_nz[slot] = upper bound on 1 + index of the last non-zero soft bit of the slot
  // once everything planned so far has run (0: the slot holds zeros). nz_scratch: n entries, holds the new values until the
  // whole batch has been planned (a rejected batch changes nothing). Both null: no bookkeeping, no limited-buffer fast path.
  int32_t* harq_nz    = nullptr;
  int32_t* nz_scratch = nullptr;
  // pdc_harq_release: harq_released[slot] != 0 = the slot reads as zeros although its memory has not been cleared yet. A first
  // transmission that rewrites the whole buffer needs no clearing; any other operation on such a slot is listed in
  // zero_slots[0 .. *nof_zero_slots) (room for n entries) and the caller clears those slots on the stream before the kernels.
  uint8_t*  harq_released  = nullptr;
  int32_t*  zero_slots     = nullptr;
  uint32_t* nof_zero_slots = nullptr;
};

// One pass over the descriptors: validation, device descriptors (dev[i], pointers included), kernel list of every code
// block (lst[i]) and the launch order. Reports the extent of the bits / host soft buffers the batch touches.
pdc_status plan_batch(const pdc_cb_desc* descs, uint32_t n, const batch_buffers& buf, pdc_dev_cb* dev, uint32_t* order,
                      uint8_t* lst, batch_plan& bp)
{
  desc_class cls;
  uint32_t   cnt[PDC_NOF_LISTS] = {0}, skipped = 0;
  size_t     bits_end = 0, soft_end = 0;
  bool       any_host_soft = false, dense = true;
  const bool track = buf.harq_nz != nullptr && buf.nz_scratch != nullptr;
  for (uint32_t i = 0; i != n; ++i) {
    const pdc_cb_desc& d = descs[i];
    if (!cls.valid || !same_class(d, cls.key)) {
      if (!classify(d, cls, bp)) {
        return PDC_ERR_INVALID_ARG;
      }
    }
    pdc_dev_cb& c = dev[i];
    c             = cls.proto;
    c.res         = buf.d_res + i;
    if (track) {
      buf.nz_scratch[i] = -1; // no change
    }
    const bool slot_bits = buf.d_slot_bits != nullptr && d.harq_slot >= 0 && !(d.flags & PDC_F_DECODE_ONLY);
    if (cls.list == PDC_LIST_SKIP) {
      if (!slot_bits || (uint32_t)d.harq_slot >= buf.harq_slots) {
        return PDC_ERR_INVALID_ARG; // PDC_F_DECODED needs the transport-block path and a HARQ slot that holds the bits
      }
      c.out  = buf.d_slot_bits + (size_t)d.harq_slot * PDC_SLOT_BITS;
      c.soft = buf.d_harq + (size_t)d.harq_slot * PDC_MAX_N;
      lst[i] = PDC_LIST_SKIP;
      ++skipped;
      continue;
    }
    if ((d.out_offset & 3) != 0 || d.llr_offset + d.E > buf.llr_bytes) {
      return PDC_ERR_INVALID_ARG;
    }
    c.llrs = buf.d_llrs + d.llr_offset;
    if (slot_bits) {
      c.out = buf.d_slot_bits + (size_t)d.harq_slot * PDC_SLOT_BITS; // range checked with the soft buffer below
      dense = false;
    } else {
      if (!buf.d_bits) {
        return PDC_ERR_INVALID_ARG;
      }
      c.out = buf.d_bits + d.out_offset;
      dense &= (d.out_offset == bits_end) && !(d.flags & PDC_F_DEMATCH_ONLY); // outputs back to back from offset 0, in
                                                                             // order, every one written by a kernel
      size_t be = d.out_offset + cls.K_bytes4;
      bits_end  = be > bits_end ? be : bits_end;
    }
    if (!(d.flags & PDC_F_DECODE_ONLY)) {
      if (d.harq_slot >= 0) {
        if ((uint32_t)d.harq_slot >= buf.harq_slots) {
          return PDC_ERR_INVALID_ARG;
        }
        c.soft = buf.d_harq + (size_t)d.harq_slot * PDC_MAX_N;
      } else {
        any_host_soft = true;
        c.soft        = buf.d_soft + d.soft_offset; // checked by the caller through bp.soft_end
        size_t se     = d.soft_offset + (size_t)cls.N;
        soft_end      = se > soft_end ? se : soft_end;
      }
    }
    int list = cls.list;
    if (track) {
      if (cls.track != 0 && d.harq_slot >= 0) {
        const bool released = buf.harq_released != nullptr && buf.harq_released[d.harq_slot] != 0;
        const int  old      = released ? 0 : buf.harq_nz[d.harq_slot];
        int        nz       = cls.N;
        // Everything of the buffer is rewritten by a one-lap first transmission without a buffer limit (and by the staging
        // tile of the four-row kernel, below); otherwise a released slot has to be cleared for real first.
        bool rewrites_all = (cls.track == 1 && cls.tail_start == cls.end_idx);
        if (cls.track == 1) {
          if (cls.tile_if_clean && old <= cls.end_idx) {
            list = cls.list_fast;
            c.flags |= PDC_DEVF_FIRST_TX_TILE;
            c.max_layers       = 4;
            bp.smem_h2[list]   = cls.smem_fast > bp.smem_h2[list] ? cls.smem_fast : bp.smem_h2[list];
            bp.block_h2[list]  = cls.block_fast > bp.block_h2[list] ? cls.block_fast : bp.block_h2[list];
            rewrites_all       = true;
          }
          const int kept = (old > cls.end_idx) ? (old < cls.tail_start ? old : cls.tail_start) : 0;
          nz             = kept > cls.end_idx ? kept : cls.end_idx;
        } else {
          nz = old > nz ? old : nz;
        }
        buf.nz_scratch[i] = nz;
        if (released && !rewrites_all && buf.zero_slots != nullptr) {
          buf.zero_slots[(*buf.nof_zero_slots)++] = d.harq_slot;
        }
      }
    }
    lst[i] = (uint8_t)list;
    cnt[list]++;
  }
  if (track) {
    for (uint32_t i = 0; i != n; ++i) {
      if (buf.nz_scratch[i] >= 0) {
        buf.harq_nz[descs[i].harq_slot] = buf.nz_scratch[i];
        if (buf.harq_released != nullptr) {
          buf.harq_released[descs[i].harq_slot] = 0;
        }
      }
    }
  }
  for (int k = 0; k != PDC_NOF_PACKED_LISTS; ++k) {
    bp.n_h2[k] = cnt[k];
  }
  bp.n_lane        = cnt[PDC_NOF_PACKED_LISTS];
  bp.bits_end      = bits_end;
  bp.bits_dense    = dense; // the packed outputs tile [0, bits_end): one copy hands them back
  bp.soft_end      = soft_end;
  bp.any_host_soft = any_host_soft;
  int used         = 0;
  for (int k = 0; k != PDC_NOF_LISTS; ++k) {
    used += (cnt[k] != 0);
  }
  bp.identity = (used == 1) && skipped == 0; // a single kernel takes the whole batch in the given order: no indirection
  if (!bp.identity) {
    uint32_t off[PDC_NOF_LISTS], acc = 0;
    for (int k = 0; k != PDC_NOF_LISTS; ++k) {
      off[k] = acc;
      acc += cnt[k];
    }
    for (uint32_t i = 0; i != n; ++i) {
      if (lst[i] != PDC_LIST_SKIP) {
        order[off[lst[i]]++] = i;
      }
    }
  }
  return PDC_OK;
}

// Transport blocks over a planned batch: validation (pdc_tb_desc contract) and the device descriptors.
struct tb_plan {
  uint32_t max_cbs  = 0;
  size_t   bytes_end = 0; // extent of the TB payload buffer
};

pdc_status plan_tbs(const pdc_tb_desc* tbs, uint32_t ntb, const pdc_cb_desc* descs, uint32_t n, uint8_t* d_tb_out,
                    pdc_tb_result* d_tb_res, uint32_t* d_terms, pdc_dev_tb* dev, tb_plan& tp)
{
  for (uint32_t t = 0; t != ntb; ++t) {
    const pdc_tb_desc& tb = tbs[t];
    if (tb.nof_cbs == 0 || tb.first_cb >= n || tb.nof_cbs > n - tb.first_cb || (tb.tbs_bits & 7) != 0 || tb.tbs_bits == 0 ||
        (tb.tb_offset & 3) != 0) {
      return PDC_ERR_INVALID_ARG;
    }
    const pdc_cb_desc& d0 = descs[tb.first_cb];
    pdc_geom           g;
    if (!pdc_geom_init(g, d0.bg, d0.Z)) {
      return PDC_ERR_INVALID_ARG;
    }
    for (uint32_t c = 0; c != tb.nof_cbs; ++c) {
      const pdc_cb_desc& d = descs[tb.first_cb + c];
      if (d.bg != d0.bg || d.Z != d0.Z || d.nof_filler != d0.nof_filler || d.crc_type != d0.crc_type ||
          (d.flags & PDC_F_DECODE_ONLY)) {
        return PDC_ERR_INVALID_ARG;
      }
    }
    uint32_t cb_bits = 0, last_bits = 0;
    if (tb.nof_cbs == 1) {
      // pusch_decoder_impl.cpp:35-46: CRC16 up to 3824 bits, CRC24A above; the message is payload || CRC || fillers.
      const uint32_t ncrc = (d0.crc_type == PDC_CRC16) ? 16 : 24;
      if ((d0.crc_type != PDC_CRC16 && d0.crc_type != PDC_CRC24A) || tb.tbs_bits + ncrc + d0.nof_filler != (uint32_t)g.K) {
        return PDC_ERR_INVALID_ARG;
      }
    } else {
      if (d0.crc_type != PDC_CRC24B || (uint32_t)g.K < d0.nof_filler + 24 + 32) {
        return PDC_ERR_INVALID_ARG;
      }
      // The payload and its CRC fill the code blocks in order; what is left of the last one is zero padding
      // (ldpc_segmenter_rx_impl.cpp:52-106: K' = ceil(B' / C); pusch_decoder_impl.cpp:485-487).
      cb_bits              = (uint32_t)g.K - d0.nof_filler - 24;
      const uint64_t total = (uint64_t)tb.tbs_bits + 24, before = (uint64_t)cb_bits * (tb.nof_cbs - 1);
      if (total <= before + 24 || total > before + cb_bits) {
        return PDC_ERR_INVALID_ARG;
      }
      last_bits = (uint32_t)(total - before);
    }
    pdc_dev_tb& o = dev[t];
    o.out = d_tb_out + tb.tb_offset, o.res = d_tb_res + t, o.terms = d_terms + tb.first_cb;
    o.first_cb = tb.first_cb, o.nof_cbs = tb.nof_cbs, o.tbs_bits = tb.tbs_bits, o.cb_bits = cb_bits, o.last_bits = last_bits;
    tp.max_cbs   = tb.nof_cbs > tp.max_cbs ? tb.nof_cbs : tp.max_cbs;
    size_t end   = tb.tb_offset + tb.tbs_bits / 8;
    tp.bytes_end = end > tp.bytes_end ? end : tp.bytes_end;
  }
  return PDC_OK;
}

pdc_status launch_batch(pdc_ctx* ctx, const batch_plan& bp, const pdc_dev_cb* d_cbs, const uint32_t* d_order,
                        cudaStream_t stream)
{
  nvtx_scope        range("pdc: code-block kernels");
  pdc_launch_params prm;
  prm.cbs     = d_cbs;
  prm.graphs  = ctx->d_graphs;
  prm.crc_pow = ctx->d_crc;
  prm.crc_bit = ctx->d_crc_bit;
  int nlists = (bp.n_lane != 0) ? 1 : 0;
  for (int k = 0; k != PDC_NOF_PACKED_LISTS; ++k) {
    nlists += (bp.n_h2[k] != 0);
  }
  // More than one list: fork. The side streams are shared by the context, so concurrent mixed batches take turns.
  std::unique_lock<std::mutex> side_lock(ctx->side_mtx, std::defer_lock);
  const bool                   fork = nlists > 1;
  if (fork) {
    side_lock.lock(); // streams and events were created by pdc_create: nothing is allocated on the data path
    CK(cudaEventRecord(ctx->side_fork, stream));
  }
  uint32_t offs[PDC_NOF_LISTS], acc = 0; // position of every list in the order array
  for (int k = 0; k != PDC_NOF_LISTS; ++k) {
    offs[k] = acc;
    acc += (k < PDC_NOF_PACKED_LISTS) ? bp.n_h2[k] : bp.n_lane;
  }
  int  used = 0; // lists launched so far: the first one stays on the caller's stream
  auto stream_of = [&](int idx) -> cudaStream_t { return (idx == 0) ? stream : ctx->side_stream[idx - 1]; };
  // Longest critical path first: the one-lane kernel (a code block is one to twelve warps running every row in sequence), then
  // the many-layer kernel, then the four-row kernels. Whoever is launched first gets its thread blocks resident first; the
  // many-layer kernel fills the register file (2 x 192 x 168) and would otherwise keep the one-lane kernel waiting.
  static const int launch_order[PDC_NOF_LISTS] = {PDC_NOF_PACKED_LISTS, 4, 5, 0, 1, 2, 3};
  for (int q = 0; q != PDC_NOF_LISTS; ++q) {
    const int      k   = launch_order[q];
    const uint32_t cnt = (k < PDC_NOF_PACKED_LISTS) ? bp.n_h2[k] : bp.n_lane;
    const uint32_t off = offs[k];
    if (cnt == 0) {
      continue;
    }
    cudaStream_t s = stream_of(used);
    if (used != 0) {
      CK(cudaStreamWaitEvent(s, ctx->side_fork, 0));
    }
    prm.order = bp.identity ? nullptr : d_order + off;
    prm.n     = (int)cnt;
    if (k == PDC_NOF_PACKED_LISTS) {
      CK(pdc_launch_lane(prm, bp.block_lane, bp.smem_lane, s));
    } else if (k >= 4) {
      CK(pdc_launch_h2g(prm, (k & 1) + 1, bp.block_h2[k], bp.smem_h2[k], s));
    } else {
      CK(pdc_launch_h2(prm, (k & 1) + 1, (k & 2) ? 384 : 0, bp.block_h2[k], bp.smem_h2[k], s));
    }
    ctx->launches++;
    if (used != 0) {
      CK(cudaEventRecord(ctx->side_join[used - 1], s));
      CK(cudaStreamWaitEvent(stream, ctx->side_join[used - 1], 0));
    }
    ++used;
  }
  return PDC_OK;
}

} // namespace
