// TEST INFRASTRUCTURE ONLY. Boundary test harness: the UNMODIFIED reference transport-block decoders
// (pusch_decoder_impl and pusch_decoder_hw_impl, compiled from /root/reference by tests/adapters/build.sh) driving
//   mode 0: the reference's own software operators (ldpc_decoder_avx512 | avx2 | generic + matching dematcher) — the oracle,
//   mode 1: the CUDA ldpc_decoder + ldpc_rate_dematcher (srsran_project_b200/adapters/channel_coding_factories_cuda.h)
//           inside the reference's pusch_codeblock_decoder / pusch_decoder_impl (drop-in boundary 1),
//   mode 2: the CUDA hal::hw_accelerator_pusch_dec inside the reference's pusch_decoder_hw_impl (drop-in boundary 2),
//   mode 4: mode 2 with the HAL metric decorator that integration/srsran_cuda.patch adds to the reference (lib/phy/metrics),
//   mode 3: the CUDA srsran::pusch_decoder (srsran_project_b200/adapters/pusch_decoder_cuda.h): code blocks AND transport
//           block assembly + CRC24A on the device, in place of the reference's TB decoders.
// Downlink mirror: the CUDA srsran::pdsch_encoder, and the CUDA hal::hw_accelerator_pdsch_enc inside the
// reference's UNMODIFIED pdsch_encoder_hw_impl (its segmenter and TB CRC on the host, everything else on the device).
// Driven from Python (tests/test_adapters.py) through a small C ABI; the stimulus comes from tools/stimulus.
#include "../../srsran_project_b200/adapters/channel_coding_factories_cuda.h"
#include "../../srsran_project_b200/adapters/channel_modulation_factories_cuda.h"
#include "../../srsran_project_b200/adapters/hw_accelerator_pdsch_enc_cuda.h"
#include "../../srsran_project_b200/adapters/hw_accelerator_pusch_dec_cuda.h"
#include "../../srsran_project_b200/adapters/pdsch_encoder_cuda.h"
#include "../../srsran_project_b200/adapters/pusch_decoder_cuda.h"
#include "../../srsran_project_b200/adapters/ulsch_demultiplex_cuda.h"
#include "phy/upper/sequence_generators/pseudo_random_generator_impl.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_codeword_buffer.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_buffer.h"
#include "du_low_hal_factory.h" // the reference's, with integration/srsran_cuda.patch applied (tests/adapters/build.sh)
#include "phy_metrics_hw_accelerator_pusch_dec_decorator.h" // added by the same patch (lib/phy/metrics)
#include "crc_calculator_clmul_impl.h"
#include "crc_calculator_lut_impl.h"
#include "ldpc/ldpc_decoder_avx2.h"
#include "ldpc/ldpc_decoder_avx512.h"
#include "ldpc/ldpc_decoder_generic.h"
#include "ldpc/ldpc_rate_dematcher_avx2_impl.h"
#include "ldpc/ldpc_rate_dematcher_avx512_impl.h"
#include "ldpc/ldpc_rate_dematcher_impl.h"
#include "ldpc/ldpc_segmenter_rx_impl.h"
#include "ldpc/ldpc_segmenter_tx_impl.h"
#include "phy/upper/channel_processors/pdsch/pdsch_encoder_hw_impl.h"
#include "pusch_codeblock_decoder.h"
#include "pusch_decoder_hw_impl.h"
#include "pusch_decoder_impl.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_notifier.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_result.h"
#include "srsran/phy/upper/rx_buffer_pool.h"
#include "srsran/phy/upper/unique_rx_buffer.h"
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

using namespace srsran;

namespace {

struct notifier_capture : public pusch_decoder_notifier {
  pusch_decoder_result last;
  bool                 called = false;
  void                 on_sch_data(const pusch_decoder_result& result) override
  {
    last   = result;
    called = true;
  }
};

std::unique_ptr<crc_calculator> make_crc(crc_generator_poly poly)
{
  if (__builtin_cpu_supports("pclmul")) {
    return std::make_unique<crc_calculator_clmul_impl>(poly);
  }
  return std::make_unique<crc_calculator_lut_impl>(poly);
}

std::unique_ptr<ldpc_decoder> make_sw_decoder(int flavour)
{
  if (flavour == 2) {
    return std::make_unique<ldpc_decoder_avx512>(false);
  }
  if (flavour == 1) {
    return std::make_unique<ldpc_decoder_avx2>(false);
  }
  return std::make_unique<ldpc_decoder_generic>(false);
}

std::unique_ptr<ldpc_rate_dematcher> make_sw_dematcher(int flavour)
{
  if (flavour == 2) {
    return std::make_unique<ldpc_rate_dematcher_avx512_impl>();
  }
  if (flavour == 1) {
    return std::make_unique<ldpc_rate_dematcher_avx2_impl>();
  }
  return std::make_unique<ldpc_rate_dematcher_impl>();
}

struct segmenter_factory_local : public ldpc_segmenter_rx_factory {
  std::unique_ptr<ldpc_segmenter_rx> create() override { return std::make_unique<ldpc_segmenter_rx_impl>(); }
};

// Collects what the HAL metric decorator of the patch reports (mode 4).
struct metric_capture : public ldpc_decoder_metric_notifier {
  std::atomic<uint64_t> count{0}, iterations{0}, crc_ok{0}, bits{0}, nanoseconds{0};
  void                  on_new_metric(const ldpc_decoder_metrics& m) override
  {
    count++;
    iterations += m.nof_iterations;
    crc_ok += m.crc_ok ? 1 : 0;
    bits += m.cb_sz.value();
    nanoseconds += static_cast<uint64_t>(m.measurements.duration.count());
  }
};

struct harness {
  metric_capture                             metrics;
  std::shared_ptr<cuda::pusch_dec_context>   context;
  std::unique_ptr<rx_buffer_pool_controller> pool;
  std::unique_ptr<pusch_decoder>             decoder;
  unsigned                                   pool_codeblocks = 0;
};

} // namespace

extern "C" {

// mode: see the header comment. dec_flavour / dem_flavour (mode 0): 0 generic, 1 AVX2, 2 AVX512. cuda_generic (modes 1, 2):
// reproduce the generic decoder instead of the AVX one. external_harq (mode 2): soft buffers live in the device arena.
void* adt_create(int mode, int dec_flavour, int dem_flavour, int cuda_generic, int external_harq, unsigned nof_buffers,
                 unsigned nof_codeblocks)
{
  __builtin_cpu_init();
  auto h             = std::make_unique<harness>();
  h->pool_codeblocks = nof_codeblocks;

  rx_buffer_pool_config pool_cfg = {};
  pool_cfg.max_codeblock_size   = ldpc::MAX_CODEBLOCK_SIZE;
  pool_cfg.nof_buffers          = nof_buffers;
  pool_cfg.nof_codeblocks       = nof_codeblocks;
  pool_cfg.expire_timeout_slots = 100000;
  pool_cfg.external_soft_bits   = (mode == 3) || ((mode == 2 || mode == 4) && (external_harq != 0));
  h->pool                       = create_rx_buffer_pool(pool_cfg);

  if (mode != 0) {
    cuda::pusch_dec_context_config ccfg;
    ccfg.max_batch_cbs   = 152;
    ccfg.max_batch_llrs  = 4u << 20;
    ccfg.harq_codeblocks = nof_codeblocks;
    ccfg.nof_instances   = 4;
    h->context           = std::make_shared<cuda::pusch_dec_context>(ccfg);
  }
  ldpc_decoder_cuda_flavour fl = cuda_generic ? ldpc_decoder_cuda_flavour::generic : ldpc_decoder_cuda_flavour::avx512;

  if (mode == 3) {
    pusch_decoder_factory_cuda_configuration dcfg;
    dcfg.segmenter_factory  = std::make_shared<segmenter_factory_local>();
    dcfg.context            = h->context;
    dcfg.generic_arithmetic = cuda_generic != 0;
    h->decoder              = create_pusch_decoder_factory_cuda(dcfg)->create();
  } else if (mode == 2 || mode == 4) {
    hal::cuda_hwacc_pusch_dec_factory_configuration acfg;
    acfg.context            = h->context;
    acfg.external_harq      = external_harq != 0;
    acfg.generic_arithmetic = cuda_generic != 0;
    acfg.batching.max_parts = 2; // the context above holds one transport block per submission (152 code blocks, 4 MB)
    auto acc_factory        = hal::create_cuda_pusch_dec_acc_factory(acfg);
    if (mode == 4) {
      // as the patched upper_phy_factories.cpp wraps the HAL factory when metrics are enabled
      acc_factory = create_hw_accelerator_pusch_dec_metric_decorator_factory(acc_factory, h->metrics);
    }
    std::vector<std::unique_ptr<hal::hw_accelerator_pusch_dec>> accs;
    for (unsigned i = 0; i != 2; ++i) {
      accs.push_back(acc_factory->create());
    }
    auto hw_pool = std::make_shared<pusch_decoder_hw_impl::hw_decoder_pool>(accs);
    pusch_decoder_hw_impl::sch_crc crcs = {make_crc(crc_generator_poly::CRC16),
                                           make_crc(crc_generator_poly::CRC24A),
                                           make_crc(crc_generator_poly::CRC24B)};
    h->decoder = std::make_unique<pusch_decoder_hw_impl>(std::make_unique<ldpc_segmenter_rx_impl>(), crcs, hw_pool, nullptr);
  } else {
    std::shared_ptr<ldpc_decoder_factory>        dec_factory;
    std::shared_ptr<ldpc_rate_dematcher_factory> dem_factory;
    if (mode == 1) {
      dec_factory = create_ldpc_decoder_factory_cuda({false}, h->context, fl);
      dem_factory = create_ldpc_rate_dematcher_factory_cuda(h->context, fl);
    }
    std::vector<std::unique_ptr<pusch_codeblock_decoder>> cb_decoders;
    for (unsigned i = 0; i != 2; ++i) {
      pusch_codeblock_decoder::sch_crc crcs = {make_crc(crc_generator_poly::CRC16),
                                               make_crc(crc_generator_poly::CRC24A),
                                               make_crc(crc_generator_poly::CRC24B)};
      cb_decoders.push_back(std::make_unique<pusch_codeblock_decoder>(
          (mode == 1) ? dem_factory->create() : make_sw_dematcher(dem_flavour),
          (mode == 1) ? dec_factory->create() : make_sw_decoder(dec_flavour),
          crcs));
    }
    auto cb_pool = std::make_shared<pusch_decoder_impl::codeblock_decoder_pool>(cb_decoders);
    pusch_decoder_impl::sch_crc crcs = {make_crc(crc_generator_poly::CRC16),
                                        make_crc(crc_generator_poly::CRC24A),
                                        make_crc(crc_generator_poly::CRC24B)};
    h->decoder = std::make_unique<pusch_decoder_impl>(
        std::make_unique<ldpc_segmenter_rx_impl>(), cb_pool, std::move(crcs), nullptr, MAX_RB, 4);
  }
  return h.release();
}

// Mode 4: what the metric decorator has reported so far: [code blocks, sum of iterations, CRC ok, sum of code-block bits, ns].
void adt_metrics(void* handle, uint64_t* out)
{
  auto* h = static_cast<harness*>(handle);
  out[0] = h->metrics.count, out[1] = h->metrics.iterations, out[2] = h->metrics.crc_ok, out[3] = h->metrics.bits;
  out[4] = h->metrics.nanoseconds;
}

void adt_destroy(void* handle)
{
  auto* h = static_cast<harness*>(handle);
  if (h) {
    h->decoder.reset();
    h->pool->stop();
    delete h;
  }
}

// Decodes one transmission of one transport block (HARQ process harq_id). Returns 0, or a negative value if the soft
// buffer could not be reserved / the notifier was not called. stats: [tb_crc_ok, nof_codeblocks, nof_observations,
// min_iterations, max_iterations, mean_iterations * 1000].
int adt_decode_tb(void*         handle,
                  unsigned      harq_id,
                  int           new_data,
                  unsigned      rv,
                  int           bg,
                  unsigned      Qm,
                  unsigned      nof_layers,
                  unsigned      Nref,
                  unsigned      max_iter,
                  int           early_stop,
                  unsigned      tbs_bits,
                  const int8_t* llrs,
                  unsigned      nof_llrs,
                  uint8_t*      tb_out,
                  int64_t*      stats)
{
  auto* h = static_cast<harness*>(handle);

  pusch_decoder::configuration cfg = {};
  cfg.base_graph          = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  cfg.rv                  = rv;
  cfg.mod                 = static_cast<modulation_scheme>(Qm);
  cfg.Nref                = Nref;
  cfg.nof_layers          = nof_layers;
  cfg.nof_ldpc_iterations = max_iter;
  cfg.force_decoding      = false;
  cfg.use_early_stop      = early_stop != 0;
  cfg.new_data            = new_data != 0;

  unsigned         nof_cbs = ldpc::compute_nof_codeblocks(units::bits(tbs_bits), cfg.base_graph);
  unique_rx_buffer buffer  = h->pool->get_pool().reserve({}, trx_buffer_identifier(1, harq_id), nof_cbs, new_data != 0);
  if (!buffer.is_valid()) {
    return -1;
  }
  if (new_data) {
    buffer.get().reset_codeblocks_crc();
  }

  std::vector<uint8_t> tb(tbs_bits / 8, 0);
  notifier_capture     notifier;
  pusch_decoder_buffer& in = h->decoder->new_data(tb, std::move(buffer), notifier, cfg);
  span<const log_likelihood_ratio> view(reinterpret_cast<const log_likelihood_ratio*>(llrs), nof_llrs);
  // Feed the codeword in two chunks with the length announced in between, as pusch_decoder_vectortest.cpp:357-375 does.
  unsigned first = nof_llrs / 2;
  in.on_new_softbits(view.first(first));
  h->decoder->set_nof_softbits(units::bits(nof_llrs));
  in.on_new_softbits(view.last(nof_llrs - first));
  in.on_end_softbits();
  if (!notifier.called) {
    return -2;
  }
  std::memcpy(tb_out, tb.data(), tb.size());
  const pusch_decoder_result& r = notifier.last;
  stats[0]                      = r.tb_crc_ok ? 1 : 0;
  stats[1]                      = r.nof_codeblocks_total;
  stats[2]                      = r.ldpc_decoder_stats.get_nof_observations();
  bool any                      = r.ldpc_decoder_stats.get_nof_observations() != 0;
  stats[3]                      = any ? r.ldpc_decoder_stats.get_min() : 0;
  stats[4]                      = any ? r.ldpc_decoder_stats.get_max() : 0;
  stats[5]                      = any ? static_cast<int64_t>(r.ldpc_decoder_stats.get_mean() * 1000.0 + 0.5) : 0;
  return 0;
}

// Latency of ONE ldpc_decoder::decode call (BASELINE config 1 at batch 1: what ldpc_decoder_benchmark.cpp:193-196 times with
// the software decoders), through the CUDA ldpc_decoder adapter or the reference's own decoder. impl: 0 generic, 1 AVX2, 2 AVX512,
// 3 CUDA. Fills us[0..reps) with the wall time of every call in microseconds and bits with the decoded message of the last one.
// Returns the iteration count of the last call (0 = nullopt), -1 = no device.
int adt_ldpc_decode_latency(int impl, int bg, unsigned Z, const int8_t* llrs, unsigned nof_llrs, unsigned max_iter, unsigned reps,
                            double* us, uint8_t* bits)
{
  __builtin_cpu_init();
  std::unique_ptr<ldpc_decoder> dec;
  try {
    if (impl == 3) {
      static std::shared_ptr<ldpc_decoder_factory> factory = create_ldpc_decoder_factory_cuda({false});
      dec = factory->create();
    } else {
      dec = make_sw_decoder(impl);
    }
  } catch (const std::exception&) {
    return -1;
  }
  const unsigned               K = pdc_cb_message_bits(bg, Z);
  dynamic_bit_buffer           message(K);
  ldpc_decoder::configuration  cfg = {};
  cfg.base_graph      = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  cfg.lifting_size    = static_cast<ldpc::lifting_size_t>(Z);
  cfg.nof_filler_bits = 0;
  cfg.nof_crc_bits    = 16;
  cfg.max_iterations  = max_iter;
  span<const log_likelihood_ratio> in(reinterpret_cast<const log_likelihood_ratio*>(llrs), nof_llrs);
  std::optional<unsigned>          it;
  for (unsigned r = 0; r != reps; ++r) {
    const auto t0 = std::chrono::steady_clock::now();
    it            = dec->decode(message, in, nullptr, cfg);
    us[r]         = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
  }
  std::memcpy(bits, message.get_buffer().data(), (K + 7) / 8);
  return it.has_value() ? static_cast<int>(*it) : 0;
}

// TEST INFRASTRUCTURE: an accelerator that answers at once with the right bits (replayed from a table), used ONLY to measure
// what the reference's pusch_decoder_hw_impl costs on the host by itself — the ceiling of the drop-in boundary whatever sits
// behind it. It takes the LLRs (one copy, as any accelerator must) and returns recorded messages.
thread_local unsigned tl_replay_codeword = 0;
class replay_accelerator : public hal::hw_accelerator_pusch_dec
{
public:
  replay_accelerator(const uint8_t* table_, unsigned nof_cbs_, unsigned cb_bytes_) :
    table(table_), nof_cbs(nof_cbs_), cb_bytes(cb_bytes_), scratch(35 * 8448)
  {
  }
  void reserve_queue() override {}
  void free_queue() override {}
  void configure_operation(const hal::hw_pusch_decoder_configuration&, unsigned) override {}
  bool enqueue_operation(span<const int8_t> data, span<const int8_t>, unsigned) override
  {
    std::memcpy(scratch.data(), data.data(), std::min(data.size(), scratch.size()));
    return true;
  }
  bool dequeue_operation(span<uint8_t> data, span<int8_t>, unsigned segment_index) override
  {
    std::memcpy(data.data(), table + ((size_t)tl_replay_codeword * nof_cbs + segment_index) * cb_bytes,
                std::min<size_t>(data.size(), cb_bytes));
    return true;
  }
  void read_operation_outputs(hal::hw_pusch_decoder_outputs& out, unsigned, unsigned) override { out = {true, 1}; }
  void free_harq_context_entry(unsigned) override {}
  bool is_harq_external() const override { return true; }

private:
  const uint8_t*      table;
  unsigned            nof_cbs, cb_bytes;
  std::vector<int8_t> scratch;
};

// Throughput of the drop-in boundary (bench.py leg e2e_hal): nthreads host threads, each with its own UNMODIFIED
// pusch_decoder_hw_impl, all drawing accelerators from one pool created by ONE CUDA HAL factory (factories.cpp:130-138), decode
// transport blocks back to back the way pusch_decoder_hwacc_benchmark.cpp:430-495 drives a decoder (new_data, one
// on_new_softbits with the whole codeword, on_end_softbits; synchronous: no executor). Inputs: n_distinct codewords of
// nof_llrs soft bits each (thread t decodes codeword (t + k) mod n_distinct as its k-th block) and the transport blocks they
// must decode to. devices: CUDA ordinals (one context each). Returns 0 and seconds = wall time of the timed part;
// stats: [transport blocks decoded correctly, transport blocks attempted, GPU submissions, code blocks submitted];
// -1 = no device, -2 = a soft buffer could not be reserved.
// Optional second transmission for adt_bench_hal (HARQ stress): when set, a transport block whose first transmission fails its
// CRC is sent again as redundancy version rv2 from llrs2 ([n_distinct][nof_llrs], same layout as the first transmissions),
// combined into the same soft buffer (new_data = false). Cleared by passing nullptr.
static uint64_t      g_last_retransmitted = 0;
uint64_t             adt_bench_hal_retransmitted() { return g_last_retransmitted; }
static const int8_t* g_retx_llrs = nullptr;
static unsigned      g_retx_rv   = 2;
void adt_bench_hal_set_retx(const int8_t* llrs2, unsigned rv2)
{
  g_retx_llrs = llrs2;
  g_retx_rv   = rv2;
}

int adt_bench_hal(unsigned       nthreads,
                  unsigned       tbs_per_thread,
                  unsigned       warmup_per_thread,
                  const int8_t*  llrs,
                  unsigned       nof_llrs,
                  unsigned       n_distinct,
                  const uint8_t* expected_tbs,
                  unsigned       tbs_bits,
                  int            bg,
                  unsigned       Qm,
                  unsigned       nof_layers,
                  unsigned       max_iter,
                  const int*     devices,
                  unsigned       nof_devices,
                  unsigned       max_parts,
                  unsigned       close_after_us,
                  const uint8_t* replay_table, // not null: [n_distinct][nof_cbs][ceil(K / 8)] messages for the replay accelerator
                  double*        seconds,
                  uint64_t*      stats)
{
  __builtin_cpu_init();
  const ldpc_base_graph_type base_graph = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  const unsigned             nof_cbs    = ldpc::compute_nof_codeblocks(units::bits(tbs_bits), base_graph);
  const unsigned             nof_harq   = 2;

  rx_buffer_pool_config pool_cfg = {};
  pool_cfg.max_codeblock_size   = ldpc::MAX_CODEBLOCK_SIZE;
  pool_cfg.nof_buffers          = nthreads * nof_harq;
  pool_cfg.nof_codeblocks       = nthreads * nof_harq * nof_cbs;
  pool_cfg.expire_timeout_slots = 100000;
  pool_cfg.external_soft_bits   = true;
  std::unique_ptr<rx_buffer_pool_controller> pool = create_rx_buffer_pool(pool_cfg);

  std::shared_ptr<hal::hw_accelerator_pusch_dec_factory> acc_factory;
  if (!replay_table) try {
    hal::cuda_hwacc_pusch_dec_factory_configuration acfg;
    acfg.devices.assign(devices, devices + nof_devices);
    acfg.external_harq           = true;
    acfg.harq_context_size       = pool_cfg.nof_codeblocks;
    acfg.batching.max_parts      = max_parts;
    acfg.batching.close_after_us = close_after_us;
    if (const char* e = std::getenv("ADT_EAGER_INFLIGHT")) { // tuning experiments
      acfg.batching.eager_inflight = (unsigned)std::atoi(e);
    }
    if (const char* e = std::getenv("ADT_NOF_BATCHES")) {
      acfg.batching.nof_batches = (unsigned)std::atoi(e);
    }
    acc_factory                  = hal::create_cuda_pusch_dec_acc_factory(acfg);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "adt_bench_hal: %s\n", e.what());
    return -1;
  }
  std::vector<std::unique_ptr<hal::hw_accelerator_pusch_dec>> accs;
  const unsigned cb_bytes = (pdc_cb_message_bits(bg, ldpc::compute_lifting_size(units::bits(tbs_bits), base_graph, nof_cbs)) + 7) / 8;
  for (unsigned i = 0; i != nthreads; ++i) {
    if (replay_table) {
      accs.push_back(std::make_unique<replay_accelerator>(replay_table, nof_cbs, cb_bytes));
    } else {
      accs.push_back(acc_factory->create());
    }
  }
  auto hw_pool = std::make_shared<pusch_decoder_hw_impl::hw_decoder_pool>(accs);
  std::vector<std::unique_ptr<pusch_decoder>> decoders;
  for (unsigned i = 0; i != nthreads; ++i) {
    pusch_decoder_hw_impl::sch_crc crcs = {make_crc(crc_generator_poly::CRC16),
                                           make_crc(crc_generator_poly::CRC24A),
                                           make_crc(crc_generator_poly::CRC24B)};
    decoders.push_back(std::make_unique<pusch_decoder_hw_impl>(std::make_unique<ldpc_segmenter_rx_impl>(), crcs, hw_pool, nullptr));
  }

  pusch_decoder::configuration cfg = {};
  cfg.base_graph          = base_graph;
  cfg.rv                  = 0;
  cfg.mod                 = static_cast<modulation_scheme>(Qm);
  cfg.Nref                = 0;
  if (const char* e = std::getenv("ADT_NREF")) { // limited-buffer rate matching as the live stack sets it (pusch_processor_impl.cpp:327)
    cfg.Nref = (unsigned)std::atoi(e);
  }
  cfg.nof_layers          = nof_layers;
  cfg.nof_ldpc_iterations = max_iter;
  cfg.force_decoding      = false;
  cfg.use_early_stop      = true;
  cfg.new_data            = true;

  std::mutex            reserve_mtx;
  std::atomic<unsigned> arrived{0}, failures{0};
  std::atomic<uint64_t> ok{0}, attempted{0}, retransmitted{0};
  std::atomic<int>      phase{0};
  const unsigned        tb_bytes = tbs_bits / 8;

  auto decode_one = [&](unsigned t, unsigned k, std::vector<uint8_t>& tb, bool count) {
    const unsigned   w = (t + k) % n_distinct;
    tl_replay_codeword = w;
    unique_rx_buffer buffer;
    for (unsigned attempt = 0; attempt != 2000 && !buffer.is_valid(); ++attempt) {
      // rx_buffer_pool::reserve is called from one thread in the gNB (the upper-PHY receive path); only unlock / release come
      // from the decoder threads. The pool's free list is a lock-free queue whose try_pop answers "empty" while a releasing
      // thread is in the middle of a push (rx_buffer_codeblock_pool.h:82-90): with more threads than cores a pusher can be
      // descheduled there, so a refused reservation is tried again (the gNB would drop the PUSCH and log a warning).
      if (attempt != 0) {
        std::this_thread::yield();
      }
      std::lock_guard<std::mutex> lock(reserve_mtx);
      buffer = pool->get_pool().reserve({}, trx_buffer_identifier(1 + t, k % nof_harq), nof_cbs, true);
    }
    if (!buffer.is_valid()) {
      failures++;
      return;
    }
    buffer.get().reset_codeblocks_crc();
    notifier_capture      notifier;
    pusch_decoder_buffer& in = decoders[t]->new_data(tb, std::move(buffer), notifier, cfg);
    in.on_new_softbits(span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(llrs) + (size_t)w * nof_llrs, nof_llrs));
    in.on_end_softbits();
    if (g_retx_llrs != nullptr && notifier.called && !notifier.last.tb_crc_ok) {
      // HARQ: the same transport block again, another redundancy version, combined with what the accelerator kept
      if (count) {
        retransmitted++;
      }
      unique_rx_buffer again;
      {
        std::lock_guard<std::mutex> lock(reserve_mtx);
        again = pool->get_pool().reserve({}, trx_buffer_identifier(1 + t, k % nof_harq), nof_cbs, false);
      }
      if (!again.is_valid()) {
        failures++;
        return;
      }
      pusch_decoder::configuration cfg2 = cfg;
      cfg2.new_data                     = false;
      cfg2.rv                           = g_retx_rv;
      notifier.called                   = false;
      pusch_decoder_buffer& in2         = decoders[t]->new_data(tb, std::move(again), notifier, cfg2);
      in2.on_new_softbits(span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(g_retx_llrs) + (size_t)w * nof_llrs, nof_llrs));
      in2.on_end_softbits();
    }
    if (count) {
      attempted++;
      if (notifier.called && notifier.last.tb_crc_ok && std::memcmp(tb.data(), expected_tbs + (size_t)w * tb_bytes, tb_bytes) == 0) {
        ok++;
      }
    }
  };
  auto sync = [&](int target) { // all threads + the main thread
    arrived++;
    while (phase.load() < target) {
      std::this_thread::yield();
    }
  };
  std::vector<std::thread> threads;
  for (unsigned t = 0; t != nthreads; ++t) {
    threads.emplace_back([&, t]() {
      std::vector<uint8_t> tb(tb_bytes, 0);
      for (unsigned k = 0; k != warmup_per_thread; ++k) {
        decode_one(t, k, tb, false);
      }
      sync(1);
      for (unsigned k = 0; k != tbs_per_thread; ++k) {
        decode_one(t, warmup_per_thread + k, tb, true);
      }
      sync(2);
    });
  }
  while (arrived.load() < nthreads) {
    std::this_thread::yield();
  }
  uint64_t b0 = 0, c0 = 0;
  if (acc_factory) {
    hal::cuda_pusch_dec_acc_factory_stats(*acc_factory, b0, c0);
  }
  arrived                                     = 0;
  const std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  phase                                       = 1;
  while (arrived.load() < nthreads) {
    std::this_thread::yield();
  }
  *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  phase    = 2;
  for (auto& th : threads) {
    th.join();
  }
  uint64_t b1 = 0, c1 = 0;
  if (acc_factory) {
    hal::cuda_pusch_dec_acc_factory_stats(*acc_factory, b1, c1);
  }
  stats[0] = ok.load(), stats[1] = attempted.load(), stats[2] = b1 - b0, stats[3] = c1 - c0;
  g_last_retransmitted = retransmitted.load();
  if (acc_factory && std::getenv("ADT_TRACE")) {
    uint64_t tm[8];
    hal::cuda_pusch_dec_acc_factory_timing(*acc_factory, tm); // includes the warm-up transport blocks
    const double np = (double)std::max<uint64_t>(tm[0], 1), nb = (double)std::max<uint64_t>(tm[1], 1);
    std::fprintf(stderr,
                 "adt_bench_hal: %u threads, %.3f ms; parts %llu batches %llu; per part: queue %.1f us, wake %.1f us, add %.1f us; "
                 "per batch: submit %.1f us, fetch %.1f us; join wait total %.1f us\n",
                 nthreads, *seconds * 1e3, (unsigned long long)tm[0], (unsigned long long)tm[1], tm[2] / np / 1e3, tm[5] / np / 1e3,
                 tm[7] / np / 1e3, tm[3] / nb / 1e3, tm[4] / nb / 1e3, tm[6] / 1e3);
  }
  decoders.clear();
  hw_pool.reset();
  pool->stop();
  return failures.load() ? -2 : 0;
}

// The CUDA demodulation_mapper adapter behind the reference's interface: one demodulate_soft call.
// mod: 0 pi/2-BPSK, 1 BPSK, 2, 4, 6, 8. Returns 0, or -1 if the factory could not be created (no device).
int adt_demodulate_soft(int8_t* llrs, const float* symbols_ri, const float* noise_vars, unsigned nof_symbols, int mod)
{
  static std::shared_ptr<demodulation_mapper_factory> factory;
  static std::unique_ptr<demodulation_mapper>         mapper;
  try {
    if (!factory) {
      factory = create_demodulation_mapper_factory_cuda();
      mapper  = factory->create();
    }
  } catch (const std::exception&) {
    return -1;
  }
  modulation_scheme scheme = modulation_scheme::QAM256;
  switch (mod) {
    case 0:
      scheme = modulation_scheme::PI_2_BPSK;
      break;
    case 1:
      scheme = modulation_scheme::BPSK;
      break;
    case 2:
      scheme = modulation_scheme::QPSK;
      break;
    case 4:
      scheme = modulation_scheme::QAM16;
      break;
    case 6:
      scheme = modulation_scheme::QAM64;
      break;
    default:
      break;
  }
  const unsigned qm = (mod == 0) ? 1 : static_cast<unsigned>(mod);
  mapper->demodulate_soft(span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(llrs), nof_symbols * qm),
                          span<const cf_t>(reinterpret_cast<const cf_t*>(symbols_ri), nof_symbols),
                          span<const float>(noise_vars, nof_symbols),
                          scheme);
  return 0;
}

// The CUDA pdsch_encoder adapter behind the reference's interface: one encode() call. Returns 0, -1 = no device.
int adt_pdsch_encode(uint8_t*       codeword,
                     const uint8_t* tb_bytes,
                     unsigned       tbs_bits,
                     int            bg,
                     unsigned       rv,
                     unsigned       qm,
                     unsigned       Nref,
                     unsigned       nof_layers,
                     unsigned       nof_ch_symbols)
{
  static std::shared_ptr<pdsch_encoder_factory> factory;
  static std::unique_ptr<pdsch_encoder>              encoder;
  try {
    if (!factory) {
      factory = create_pdsch_encoder_factory_cuda();
      encoder = factory->create();
    }
  } catch (const std::exception&) {
    return -1;
  }
  pdsch_encoder::configuration cfg;
  cfg.base_graph     = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  cfg.rv             = rv;
  cfg.Nref           = Nref;
  cfg.nof_layers     = nof_layers;
  cfg.nof_ch_symbols = nof_ch_symbols;
  cfg.mod            = (qm == 1)   ? modulation_scheme::BPSK
                       : (qm == 2) ? modulation_scheme::QPSK
                       : (qm == 4) ? modulation_scheme::QAM16
                       : (qm == 6) ? modulation_scheme::QAM64
                                   : modulation_scheme::QAM256;
  encoder->encode(span<uint8_t>(codeword, nof_ch_symbols * qm), span<const uint8_t>(tb_bytes, tbs_bits / 8), cfg);
  return 0;
}

// The CUDA hal::hw_accelerator_pdsch_enc inside the reference's pdsch_encoder_hw_impl (pdsch_encoder_hw_impl.cpp:31-186):
// the reference computes the segmentation and hands the accelerator hal::hw_pdsch_encoder_configuration. Returns 0, -1 = no device.
int adt_pdsch_encode_hw(uint8_t*       codeword,
                        const uint8_t* tb_bytes,
                        unsigned       tbs_bits,
                        int            bg,
                        unsigned       rv,
                        unsigned       qm,
                        unsigned       Nref,
                        unsigned       nof_layers,
                        unsigned       nof_ch_symbols)
{
  static std::unique_ptr<pdsch_encoder> encoder;
  try {
    if (!encoder) {
      std::shared_ptr<hal::hw_accelerator_pdsch_enc_factory> hw_factory = hal::create_cuda_pdsch_enc_acc_factory({});
      pdsch_encoder_hw_impl::sch_crc                         crcs       = {
          make_crc(crc_generator_poly::CRC16), make_crc(crc_generator_poly::CRC24A), make_crc(crc_generator_poly::CRC24B)};
      ldpc_segmenter_tx_impl::sch_crc seg_crcs = {
          make_crc(crc_generator_poly::CRC16), make_crc(crc_generator_poly::CRC24A), make_crc(crc_generator_poly::CRC24B)};
      encoder = std::make_unique<pdsch_encoder_hw_impl>(
          crcs, std::make_unique<ldpc_segmenter_tx_impl>(seg_crcs), hw_factory->create());
    }
  } catch (const std::exception&) {
    return -1;
  }
  pdsch_encoder::configuration cfg;
  cfg.base_graph     = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  cfg.rv             = rv;
  cfg.Nref           = Nref;
  cfg.nof_layers     = nof_layers;
  cfg.nof_ch_symbols = nof_ch_symbols;
  cfg.mod            = (qm == 1)   ? modulation_scheme::BPSK
                       : (qm == 2) ? modulation_scheme::QPSK
                       : (qm == 4) ? modulation_scheme::QAM16
                       : (qm == 6) ? modulation_scheme::QAM64
                                   : modulation_scheme::QAM256;
  encoder->encode(span<uint8_t>(codeword, nof_ch_symbols * qm), span<const uint8_t>(tb_bytes, tbs_bits / 8), cfg);
  return 0;
}

// The CUDA srsran::ulsch_demultiplex behind the reference's interface, driven the way pusch_demodulator_impl
// drives it: blocks of descrambled soft bits with their scrambling bits, through get_next_block_view when `use_view`.
// cfg as in oracle/ref_shim_demux.cpp (cfg[15]: 0 = set_csi_part2 before the data, 1 = from the CSI Part 1 end-of-stream
// callback; cfg[17]: block size, 0 = one block), cfg[18]: use_view. Returns 0, -1 = no device.
int adt_ulsch_demultiplex(const uint32_t* cfg,
                          const int8_t*   in,
                          unsigned        n_in,
                          int8_t*         out_sch,
                          int8_t*         out_harq,
                          int8_t*         out_csi1,
                          int8_t*         out_csi2,
                          unsigned        cap,
                          unsigned*       lens)
{
  struct collector : public pusch_decoder_buffer {
    std::vector<log_likelihood_ratio> data;
    unsigned                          ended = 0;
    ulsch_demultiplex*                demux = nullptr;
    pusch_decoder_buffer*             csi2  = nullptr;
    unsigned                          csi2_bits = 0, csi2_enc_bits = 0;
    span<log_likelihood_ratio>        get_next_block_view(unsigned) override { return {}; }
    void on_new_softbits(span<const log_likelihood_ratio> softbits) override { data.insert(data.end(), softbits.begin(), softbits.end()); }
    void on_end_softbits() override
    {
      ++ended;
      if (demux != nullptr) {
        demux->set_csi_part2(*csi2, csi2_bits, csi2_enc_bits);
      }
    }
  };
  static std::shared_ptr<ulsch_demultiplex_factory> factory;
  static std::unique_ptr<ulsch_demultiplex>         demux;
  try {
    if (!factory) {
      factory = create_ulsch_demultiplex_factory_cuda();
      demux   = factory->create();
    }
  } catch (const std::exception&) {
    return -1;
  }
  ulsch_demultiplex::configuration c;
  c.modulation         = (cfg[0] == 1)   ? modulation_scheme::PI_2_BPSK
                         : (cfg[0] == 2) ? modulation_scheme::QPSK
                         : (cfg[0] == 4) ? modulation_scheme::QAM16
                         : (cfg[0] == 6) ? modulation_scheme::QAM64
                                         : modulation_scheme::QAM256;
  c.nof_layers         = cfg[1];
  c.nof_prb            = cfg[2];
  c.start_symbol_index = cfg[3];
  c.nof_symbols        = cfg[4];
  c.nof_harq_ack_rvd   = cfg[5];
  c.dmrs               = (cfg[6] == 1) ? dmrs_type::TYPE1 : dmrs_type::TYPE2;
  c.dmrs_symbol_mask   = symbol_slot_mask(14);
  for (unsigned l = 0; l != 14; ++l) {
    c.dmrs_symbol_mask.set(l, ((cfg[7] >> l) & 1u) != 0);
  }
  c.nof_cdm_groups_without_data = cfg[8];
  c.nof_harq_ack_bits           = cfg[9];
  c.nof_enc_harq_ack_bits       = cfg[10];
  c.nof_csi_part1_bits          = cfg[11];
  c.nof_enc_csi_part1_bits      = cfg[12];
  collector              sch, harq, csi1, csi2;
  pusch_codeword_buffer& cw = demux->demultiplex(sch, harq, csi1, c);
  if (cfg[14] != 0) {
    if (cfg[15] == 0) {
      demux->set_csi_part2(csi2, cfg[13], cfg[14]);
    } else {
      csi1.demux = demux.get(), csi1.csi2 = &csi2, csi1.csi2_bits = cfg[13], csi1.csi2_enc_bits = cfg[14];
    }
  }
  pseudo_random_generator_impl prg;
  prg.init(cfg[16]);
  dynamic_bit_buffer seq(n_in);
  prg.generate(seq);
  const unsigned block = (cfg[17] == 0) ? n_in : cfg[17];
  for (unsigned o = 0; o < n_in;) {
    unsigned                   n    = (n_in - o < block) ? n_in - o : block;
    span<log_likelihood_ratio> view = (cfg[18] != 0) ? cw.get_next_block_view(n) : span<log_likelihood_ratio>();
    std::vector<log_likelihood_ratio> local;
    if (!view.empty()) {
      n = static_cast<unsigned>(view.size()); // the buffer may hand out less than asked for
      std::memcpy(view.data(), in + o, n);
    } else {
      local.assign(in + o, in + o + n);
      view = local;
    }
    dynamic_bit_buffer part(n);
    for (unsigned k = 0; k != n; ++k) {
      part.insert(seq.extract(o + k, 1), k, 1);
    }
    cw.on_new_block(view, part);
    o += n;
  }
  cw.on_end_codeword();
  const collector* all[4]  = {&sch, &harq, &csi1, &csi2};
  int8_t*          outs[4] = {out_sch, out_harq, out_csi1, out_csi2};
  for (int q = 0; q != 4; ++q) {
    lens[q] = static_cast<unsigned>(all[q]->data.size());
    if (lens[q] <= cap && lens[q] != 0) {
      std::memcpy(outs[q], all[q]->data.data(), lens[q]);
    }
    // Every configured stream ends exactly once.
    const bool configured = (q == 0) || (q == 1 && cfg[9] != 0) || (q == 2 && cfg[11] != 0) || (q == 3 && cfg[14] != 0);
    if (all[q]->ended != (configured ? 1u : 0u)) {
      return -2 - q;
    }
  }
  return 0;
}

// Product wiring: the reference's make_du_low_hal_dependencies, patched by integration/srsran_cuda.patch, given a
// `hal: cuda_hwacc:` configuration. Returns a bit mask: 1 = PUSCH decoder factory present, 2 = PDSCH encoder factory present,
// 4 = an accelerator could be created from each factory present; -1 = no device.
int adt_n3_hal_dependencies(int pusch_dec, int pdsch_enc, int external_harq)
{
  du_low_unit_hal_config hal_cfg;
  hal_cfg.cuda_hwacc.emplace();
  hal_cfg.cuda_hwacc->device            = 0;
  hal_cfg.cuda_hwacc->pusch_dec         = pusch_dec != 0;
  hal_cfg.cuda_hwacc->pdsch_enc         = pdsch_enc != 0;
  hal_cfg.cuda_hwacc->external_harq     = external_harq != 0;
  hal_cfg.cuda_hwacc->harq_context_size = 304;
  o_du_low_hal_dependencies deps;
  try {
    deps = make_du_low_hal_dependencies(hal_cfg);
  } catch (const std::exception&) {
    return -1;
  }
  int  mask = (deps.hw_decoder_factory ? 1 : 0) | (deps.hw_encoder_factory ? 2 : 0);
  bool ok   = true;
  if (deps.hw_decoder_factory) {
    ok = ok && (deps.hw_decoder_factory->create() != nullptr);
  }
  if (deps.hw_encoder_factory) {
    ok = ok && (deps.hw_encoder_factory->create() != nullptr);
  }
  return mask | (ok ? 4 : 0);
}

} // extern "C"
