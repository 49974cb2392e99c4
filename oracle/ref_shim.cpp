// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// C-ABI shim over the UNMODIFIED srsRAN reference classes, compiled from the sources where they lie under
// /root/reference by oracle/build_ref.sh into oracle/_ref/libsrsran_ref.so (git-ignored; travels to the GPU box).
// It exists to (1) pin the C restatement in oracle/pdc_oracle.c, (2) generate golden fixtures and the LDPC table
// dump, and (3) serve as the "reference" CPU baseline of bench.py.  Nothing under srsran_project_b200/ links it.
//
// Reference classes instantiated here (all from /root/reference/lib/phy/upper/channel_coding):
//   ldpc/ldpc_decoder_{generic,avx2,avx512}.h, ldpc/ldpc_rate_dematcher_{impl,avx2_impl,avx512_impl}.h,
//   crc_calculator_{generic,lut,clmul}_impl.h, ldpc/ldpc_encoder_{generic,avx2}.h, ldpc/ldpc_rate_matcher_impl.h,
//   ldpc/ldpc_segmenter_rx_impl.h, ldpc/ldpc_luts_impl.h.
#include "crc_calculator_clmul_impl.h"
#include "crc_calculator_generic_impl.h"
#include "crc_calculator_lut_impl.h"
#include "ldpc/ldpc_decoder_avx2.h"
#include "ldpc/ldpc_decoder_avx512.h"
#include "ldpc/ldpc_decoder_generic.h"
#include "ldpc/ldpc_encoder_avx2.h"
#include "ldpc/ldpc_encoder_generic.h"
#include "ldpc/ldpc_graph_impl.h"
#include "ldpc/ldpc_luts_impl.h"
#include "ldpc/ldpc_rate_dematcher_avx2_impl.h"
#include "ldpc/ldpc_rate_dematcher_avx512_impl.h"
#include "ldpc/ldpc_rate_dematcher_impl.h"
#include "ldpc/ldpc_rate_matcher_impl.h"
#include "ldpc/ldpc_segmenter_rx_impl.h"
#include "pusch_codeblock_decoder.h"
#include "srsran/phy/upper/channel_coding/ldpc/ldpc.h"
#include "srsran/ran/sch/modulation_scheme.h"
#include <cstdint>
#include <cstring>
#include <memory>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

using namespace srsran;

namespace {

enum flavour_t { FLAVOUR_GENERIC = 0, FLAVOUR_AVX2 = 1, FLAVOUR_AVX512 = 2 };

modulation_scheme mod_from_qm(unsigned qm)
{
  switch (qm) {
    case 1:
      return modulation_scheme::BPSK;
    case 2:
      return modulation_scheme::QPSK;
    case 4:
      return modulation_scheme::QAM16;
    case 6:
      return modulation_scheme::QAM64;
    default:
      return modulation_scheme::QAM256;
  }
}

crc_generator_poly poly_from_id(int id)
{
  // id: 1 = CRC16, 2 = CRC24A, 3 = CRC24B.
  switch (id) {
    case 1:
      return crc_generator_poly::CRC16;
    case 2:
      return crc_generator_poly::CRC24A;
    default:
      return crc_generator_poly::CRC24B;
  }
}

std::unique_ptr<ldpc_decoder> make_decoder(int flavour, bool force)
{
  switch (flavour) {
    case FLAVOUR_AVX512:
      return std::make_unique<ldpc_decoder_avx512>(force);
    case FLAVOUR_AVX2:
      return std::make_unique<ldpc_decoder_avx2>(force);
    default:
      return std::make_unique<ldpc_decoder_generic>(force);
  }
}

std::unique_ptr<ldpc_rate_dematcher> make_dematcher(int flavour)
{
  switch (flavour) {
    case FLAVOUR_AVX512:
      return std::make_unique<ldpc_rate_dematcher_avx512_impl>();
    case FLAVOUR_AVX2:
      return std::make_unique<ldpc_rate_dematcher_avx2_impl>();
    default:
      return std::make_unique<ldpc_rate_dematcher_impl>();
  }
}

std::unique_ptr<crc_calculator> make_crc(int impl, crc_generator_poly poly)
{
  switch (impl) {
    case 2:
      return std::make_unique<crc_calculator_clmul_impl>(poly);
    case 1:
      return std::make_unique<crc_calculator_lut_impl>(poly);
    default:
      return std::make_unique<crc_calculator_generic_impl>(poly);
  }
}

// Per-thread handle bundling one instance of each operator (the reference pools them the same way).
struct ref_handle {
  int                                  flavour;
  std::unique_ptr<ldpc_decoder>        dec_force[2];
  std::unique_ptr<ldpc_rate_dematcher> dematcher;
  std::unique_ptr<crc_calculator>      crc[4];
  // The PCLMUL calculator of the reference indexes out of bounds for 13..15-byte messages
  // (crc_calculator_clmul_impl.cpp:134: data.subspan(16, ...) with data.size() < 16); the LUT one is used there.
  std::unique_ptr<crc_calculator>      crc_lut[4];
  std::unique_ptr<ldpc_encoder>        encoder;
  std::unique_ptr<ldpc_rate_matcher>   matcher;
};

} // namespace

extern "C" {

void* ref_create(int flavour)
{
  auto* h         = new ref_handle;
  h->flavour      = flavour;
  h->dec_force[0] = make_decoder(flavour, false);
  h->dec_force[1] = make_decoder(flavour, true);
  h->dematcher    = make_dematcher(flavour);
  int crc_impl    = (flavour == FLAVOUR_GENERIC) ? 1 : 2;
  for (int id = 1; id <= 3; ++id) {
    h->crc[id]     = make_crc(crc_impl, poly_from_id(id));
    h->crc_lut[id] = make_crc(1, poly_from_id(id));
  }
  if (flavour == FLAVOUR_GENERIC) {
    h->encoder = std::make_unique<ldpc_encoder_generic>();
  } else {
    h->encoder = std::make_unique<ldpc_encoder_avx2>();
  }
  h->matcher = std::make_unique<ldpc_rate_matcher_impl>();
  return h;
}

void ref_destroy(void* handle)
{
  delete static_cast<ref_handle*>(handle);
}

// ldpc_decoder::decode. crc_id: 0 = nullptr, 1 = CRC16, 2 = CRC24A, 3 = CRC24B. Returns iterations or -1 (nullopt).
// out_packed holds K = Kb*Z bits, MSB first.
int ref_ldpc_decode(void*         handle,
                    int           bg,
                    int           Z,
                    int           nof_filler,
                    int           nof_crc_bits,
                    int           max_iter,
                    int           crc_id,
                    int           force_decoding,
                    const int8_t* llrs,
                    int           n_llrs,
                    uint8_t*      out_packed)
{
  auto*                       h = static_cast<ref_handle*>(handle);
  ldpc_decoder::configuration cfg;
  cfg.base_graph      = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  cfg.lifting_size    = static_cast<ldpc::lifting_size_t>(Z);
  cfg.nof_filler_bits = nof_filler;
  cfg.nof_crc_bits    = nof_crc_bits;
  cfg.max_iterations  = max_iter;
  unsigned K          = ((bg == 1) ? 22 : 10) * Z;
  bit_buffer out      = bit_buffer::from_bytes(span<uint8_t>(out_packed, (K + 7) / 8)).first(K);
  span<const log_likelihood_ratio> in(reinterpret_cast<const log_likelihood_ratio*>(llrs), n_llrs);
  crc_calculator*                  crc = (crc_id == 0) ? nullptr : h->crc[crc_id].get();
  unsigned                         crc_bytes = (K - nof_filler) / 8;
  if ((crc_id != 0) && (crc_bytes >= 13) && (crc_bytes <= 15)) {
    crc = h->crc_lut[crc_id].get();
  }
  std::optional<unsigned>          r   = h->dec_force[force_decoding ? 1 : 0]->decode(out, in, crc, cfg);
  return r.has_value() ? static_cast<int>(*r) : -1;
}

// ldpc_rate_dematcher::rate_dematch. buf (in/out) has N = 66Z|50Z entries.
void ref_rate_dematch(void*         handle,
                      int8_t*       buf,
                      int           N,
                      const int8_t* in,
                      int           E,
                      int           new_data,
                      int           bg,
                      int           Z,
                      int           rv,
                      int           Qm,
                      int           Nref,
                      int           nof_filler)
{
  auto*              h = static_cast<ref_handle*>(handle);
  codeblock_metadata md;
  md.tb_common.base_graph        = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  md.tb_common.lifting_size      = static_cast<ldpc::lifting_size_t>(Z);
  md.tb_common.rv                = rv;
  md.tb_common.mod               = mod_from_qm(Qm);
  md.tb_common.Nref              = Nref;
  md.cb_specific.nof_filler_bits = nof_filler;
  md.cb_specific.full_length     = N;
  md.cb_specific.rm_length       = E;
  h->dematcher->rate_dematch(span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(buf), N),
                             span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(in), E),
                             new_data != 0,
                             md);
}

// crc_calculator::calculate over the first nbits of a packed MSB-first buffer. impl: 0 generic, 1 lut, 2 clmul.
uint32_t ref_crc(int impl, int crc_id, const uint8_t* packed, int nbits)
{
  auto                 crc = make_crc(impl, poly_from_id(crc_id));
  std::vector<uint8_t> tmp(packed, packed + (nbits + 7) / 8);
  bit_buffer           bb = bit_buffer::from_bytes(span<uint8_t>(tmp)).first(nbits);
  return crc->calculate(bb);
}

// ldpc_encoder::encode + write_codeblock. msg_packed: K bits (fillers as zero). cb_unpacked: N = 66Z|50Z bytes of 0/1.
void ref_ldpc_encode(void* handle, int bg, int Z, const uint8_t* msg_packed, uint8_t* cb_unpacked)
{
  auto*                       h = static_cast<ref_handle*>(handle);
  ldpc_encoder::configuration cfg;
  cfg.base_graph   = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  cfg.lifting_size = static_cast<ldpc::lifting_size_t>(Z);
  cfg.Nref         = 0;
  unsigned             K = ((bg == 1) ? 22 : 10) * Z;
  unsigned             N = ((bg == 1) ? 66 : 50) * Z;
  std::vector<uint8_t> tmp(msg_packed, msg_packed + (K + 7) / 8);
  bit_buffer           msg = bit_buffer::from_bytes(span<uint8_t>(tmp)).first(K);
  const ldpc_encoder_buffer& enc = h->encoder->encode(msg, cfg);
  enc.write_codeblock(span<uint8_t>(cb_unpacked, N), 0);
}

// ldpc_encoder + ldpc_rate_matcher::rate_match. out_unpacked: E bytes of 0/1.
void ref_ldpc_encode_rm(void*          handle,
                        int            bg,
                        int            Z,
                        const uint8_t* msg_packed,
                        int            rv,
                        int            Qm,
                        int            Nref,
                        int            nof_filler,
                        int            E,
                        uint8_t*       out_unpacked)
{
  auto*                       h = static_cast<ref_handle*>(handle);
  ldpc_encoder::configuration cfg;
  cfg.base_graph   = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  cfg.lifting_size = static_cast<ldpc::lifting_size_t>(Z);
  cfg.Nref         = Nref;
  unsigned             K = ((bg == 1) ? 22 : 10) * Z;
  std::vector<uint8_t> tmp(msg_packed, msg_packed + (K + 7) / 8);
  bit_buffer           msg = bit_buffer::from_bytes(span<uint8_t>(tmp)).first(K);
  const ldpc_encoder_buffer& enc = h->encoder->encode(msg, cfg);

  codeblock_metadata md;
  md.tb_common.base_graph        = cfg.base_graph;
  md.tb_common.lifting_size      = cfg.lifting_size;
  md.tb_common.rv                = rv;
  md.tb_common.mod               = mod_from_qm(Qm);
  md.tb_common.Nref              = Nref;
  md.cb_specific.nof_filler_bits = nof_filler;
  md.cb_specific.rm_length       = E;
  std::vector<uint8_t> packed((E + 7) / 8);
  bit_buffer           out = bit_buffer::from_bytes(span<uint8_t>(packed)).first(E);
  h->matcher->rate_match(out, enc, md);
  for (int i = 0; i != E; ++i) {
    out_unpacked[i] = out.extract<uint8_t>(i, 1);
  }
}

// Dumps one lifted base graph: for each check node m, the adjacency row up to the first NO_EDGE and the shifts.
// cols/shifts are [46][20] row-major; deg is [46]. Returns the number of check nodes.
int ref_graph_dump(int bg, int Z, uint16_t* deg, uint16_t* cols, uint16_t* shifts)
{
  ldpc_base_graph_type bgt   = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  ldpc_graph_impl      graph(bgt, static_cast<ldpc::lifting_size_t>(Z));
  unsigned             M = graph.get_nof_BG_check_nodes();
  for (unsigned m = 0; m != M; ++m) {
    const ldpc::BG_adjacency_row_t& row = graph.get_adjacency_row(m);
    unsigned                        d   = 0;
    for (; d != row.size() && row[d] != ldpc::NO_EDGE; ++d) {
      cols[m * 20 + d]   = row[d];
      shifts[m * 20 + d] = graph.get_lifted_node(m, row[d]);
    }
    deg[m] = d;
  }
  return M;
}

// ldpc_segmenter_rx_impl::segment metadata. Arrays sized 152. Returns the number of codeblocks.
int ref_segment_rx(int       bg,
                   int       rv,
                   int       Qm,
                   int       Nref,
                   int       nof_layers,
                   int       nof_llrs,
                   int       tbs_bits,
                   unsigned* Z,
                   unsigned* full_length,
                   unsigned* rm_length,
                   unsigned* nof_filler,
                   unsigned* cw_offset,
                   unsigned* nof_crc_bits)
{
  ldpc_segmenter_rx_impl seg;
  segmenter_config       cfg;
  cfg.base_graph     = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  cfg.rv             = rv;
  cfg.mod            = mod_from_qm(Qm);
  cfg.Nref           = Nref;
  cfg.nof_layers     = nof_layers;
  cfg.nof_ch_symbols = nof_llrs / Qm;
  std::vector<log_likelihood_ratio>                             llrs(nof_llrs);
  static_vector<described_rx_codeblock, MAX_NOF_SEGMENTS>       cbs;
  seg.segment(cbs, llrs, tbs_bits, cfg);
  for (unsigned i = 0; i != cbs.size(); ++i) {
    const codeblock_metadata& md = cbs[i].second;
    Z[i]                         = md.tb_common.lifting_size;
    full_length[i]               = md.cb_specific.full_length;
    rm_length[i]                 = md.cb_specific.rm_length;
    nof_filler[i]                = md.cb_specific.nof_filler_bits;
    cw_offset[i]                 = md.cb_specific.cw_offset;
    nof_crc_bits[i]              = md.cb_specific.nof_crc_bits;
  }
  return cbs.size();
}

// ---- CPU baseline driver (bench.py --impl reference / cpu_baseline) ---------------------------------------------------

// Bit 0: AVX2, bit 1: AVX512F+BW (ldpc_decoder_avx512), bit 2: AVX512VBMI (ldpc_rate_dematcher_avx512_impl),
// bit 3: PCLMUL. The caller picks the widest flavour the host supports (channel_coding_factories.cpp:114-137 "auto").
int ref_cpu_flags()
{
  __builtin_cpu_init();
  int f = 0;
  f |= __builtin_cpu_supports("avx2") ? 1 : 0;
  f |= (__builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw")) ? 2 : 0;
  f |= __builtin_cpu_supports("avx512vbmi") ? 4 : 0;
  f |= __builtin_cpu_supports("pclmul") ? 8 : 0;
  return f;
}

struct ref_cb_desc {
  int32_t  bg, Z, Qm, rv, crc_id, max_iter, new_data, early_stop;
  uint32_t E, Nref, nof_filler, nof_crc_bits;
  uint64_t llr_offset, soft_offset, out_offset;
  int32_t  decode_only, pad;
};

// Runs n code blocks `reps` times through the UNMODIFIED reference chain on `nthreads` host threads, each thread owning
// one pusch_codeblock_decoder (rate dematcher + LDPC decoder + CRC calculators), exactly like the reference's pool of
// nof_pusch_decoder_threads instances (lib/phy/upper/channel_processors/pusch/factories.cpp:82-94). Code blocks are
// handed out dynamically. decode_only != 0 runs ldpc_decoder::decode alone on llrs (ldpc_decoder_benchmark.cpp:193).
// dec_flavour / dem_flavour: 0 generic, 1 AVX2, 2 AVX512. Returns the wall-clock seconds of the parallel region.
double ref_chain_run(int                dec_flavour,
                     int                dem_flavour,
                     int                nthreads,
                     int                n,
                     const ref_cb_desc* cbs,
                     const int8_t*      llrs,
                     int8_t*            soft,
                     uint8_t*           bits,
                     int32_t*           iters,
                     int                reps)
{
  std::atomic<int>         next{0};
  std::atomic<int>         ready{0};
  std::atomic<bool>        go{false};
  std::vector<std::thread> pool;
  int                      total = n * reps;
  auto                     body  = [&]() {
    pusch_codeblock_decoder::sch_crc crcs;
    int crc_impl = (dec_flavour == FLAVOUR_GENERIC) ? 1 : 2;
    crcs.crc16   = make_crc(crc_impl, crc_generator_poly::CRC16);
    crcs.crc24A  = make_crc(crc_impl, crc_generator_poly::CRC24A);
    crcs.crc24B  = make_crc(crc_impl, crc_generator_poly::CRC24B);
    pusch_codeblock_decoder       chain(make_dematcher(dem_flavour), make_decoder(dec_flavour, false), crcs);
    std::unique_ptr<ldpc_decoder> dec = make_decoder(dec_flavour, false);
    std::unique_ptr<crc_calculator> crc_alone[4];
    for (int id = 1; id <= 3; ++id) {
      crc_alone[id] = make_crc(crc_impl, poly_from_id(id));
    }
    ready++;
    while (!go.load()) {
      std::this_thread::yield();
    }
    for (;;) {
      int k = next.fetch_add(1);
      if (k >= total) {
        break;
      }
      const ref_cb_desc& d = cbs[k % n];
      unsigned           K = ((d.bg == 1) ? 22 : 10) * d.Z;
      unsigned           N = ((d.bg == 1) ? 66 : 50) * d.Z;
      bit_buffer out = bit_buffer::from_bytes(span<uint8_t>(bits + d.out_offset, (K + 7) / 8)).first(K);
      std::optional<unsigned> r;
      if (d.decode_only) {
        ldpc_decoder::configuration cfg;
        cfg.base_graph      = (d.bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
        cfg.lifting_size    = static_cast<ldpc::lifting_size_t>(d.Z);
        cfg.nof_filler_bits = d.nof_filler;
        cfg.nof_crc_bits    = d.nof_crc_bits;
        cfg.max_iterations  = d.max_iter;
        span<const log_likelihood_ratio> in(reinterpret_cast<const log_likelihood_ratio*>(llrs + d.llr_offset), d.E);
        r = dec->decode(out, in, d.crc_id ? crc_alone[d.crc_id].get() : nullptr, cfg);
      } else {
        codeblock_metadata md;
        md.tb_common.base_graph        = (d.bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
        md.tb_common.lifting_size      = static_cast<ldpc::lifting_size_t>(d.Z);
        md.tb_common.rv                = d.rv;
        md.tb_common.mod               = mod_from_qm(d.Qm);
        md.tb_common.Nref              = d.Nref;
        md.cb_specific.nof_filler_bits = d.nof_filler;
        md.cb_specific.nof_crc_bits    = d.nof_crc_bits;
        md.cb_specific.full_length     = N;
        md.cb_specific.rm_length       = d.E;
        span<log_likelihood_ratio> buf(reinterpret_cast<log_likelihood_ratio*>(soft + d.soft_offset), N);
        span<const log_likelihood_ratio> in(reinterpret_cast<const log_likelihood_ratio*>(llrs + d.llr_offset), d.E);
        r = chain.decode(out, buf, in, d.new_data != 0, poly_from_id(d.crc_id), d.early_stop != 0, d.max_iter, false, md);
      }
      iters[k % n] = r.has_value() ? static_cast<int>(*r) : -1;
    }
  };
  for (int t = 0; t != nthreads; ++t) {
    pool.emplace_back(body);
  }
  while (ready.load() != nthreads) {
    std::this_thread::yield();
  }
  auto t0 = std::chrono::steady_clock::now();
  go.store(true);
  for (auto& th : pool) {
    th.join();
  }
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

} // extern "C"
