// See pusch_decoder_cuda.h. The host keeps what pusch_decoder_hw_impl keeps (pusch_decoder_hw_impl.cpp:149-449): the
// codeword soft bits until on_end_softbits, the segmentation metadata, the code-block CRC flags of the rx_buffer and the
// iteration statistics. Everything else is one pdc_submit_tbs + pdc_fetch_tbs.
#include "pusch_decoder_cuda.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_buffer.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_notifier.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_result.h"
#include "srsran/phy/upper/rx_buffer.h"
#include "srsran/phy/upper/unique_rx_buffer.h"
#include "srsran/phy/upper/channel_coding/ldpc/ldpc.h"
#include "srsran/ran/pdsch/pdsch_constants.h"
#include "srsran/ran/sch/sch_constants.h"
#include "srsran/support/executors/task_executor.h"
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

using namespace srsran;

namespace {

class pusch_decoder_cuda_impl : public pusch_decoder, private pusch_decoder_buffer
{
public:
  pusch_decoder_cuda_impl(std::unique_ptr<ldpc_segmenter_rx>       segmenter_,
                          std::shared_ptr<cuda::pusch_dec_context> context_,
                          task_executor*                           executor_,
                          unsigned                                 max_softbits,
                          bool                                     generic_) :
    segmenter(std::move(segmenter_)),
    context(std::move(context_)),
    executor(executor_),
    generic(generic_),
    softbits_buffer(max_softbits)
  {
    descs.reserve(MAX_NOF_SEGMENTS);
    results.resize(MAX_NOF_SEGMENTS);
  }

  // See interface for the documentation.
  pusch_decoder_buffer& new_data(span<uint8_t>           transport_block_,
                                 unique_rx_buffer        rm_buffer,
                                 pusch_decoder_notifier& notifier,
                                 const configuration&    cfg) override
  {
    transport_block = transport_block_;
    softbuffer      = std::move(rm_buffer);
    result_notifier = &notifier;
    current_config  = cfg;
    softbits_count  = 0;
    return *this;
  }

  // The whole transport block is one submission: nothing can start before the last soft bit (pusch_decoder.h:96-102 allows
  // an implementation to ignore the hint).
  void set_nof_softbits(units::bits /*nof_softbits*/) override {}

private:
  // See pusch_decoder_buffer for the documentation.
  span<log_likelihood_ratio> get_next_block_view(unsigned block_size) override
  {
    srsran_assert(softbits_count + block_size <= softbits_buffer.size(),
                  "Codeword of {} soft bits exceeds the decoder buffer ({}).",
                  softbits_count + block_size,
                  softbits_buffer.size());
    return span<log_likelihood_ratio>(softbits_buffer).subspan(softbits_count, block_size);
  }

  void on_new_softbits(span<const log_likelihood_ratio> softbits) override
  {
    span<log_likelihood_ratio> block = get_next_block_view(softbits.size());
    if (block.data() != softbits.data()) {
      std::memcpy(block.data(), softbits.data(), softbits.size());
    }
    softbits_count += softbits.size();
  }

  void on_end_softbits() override
  {
    bool deferred = false;
    if (executor != nullptr) {
      deferred = executor->defer([this]() { run(); });
    }
    if (!deferred) {
      run();
    }
  }

  void run()
  {
    const unsigned modulation_order = get_bits_per_symbol(current_config.mod);
    srsran_assert(softbits_count % modulation_order == 0, "The codeword is not a whole number of symbols.");

    segmenter_config seg_cfg;
    seg_cfg.base_graph     = current_config.base_graph;
    seg_cfg.rv             = current_config.rv;
    seg_cfg.mod            = current_config.mod;
    seg_cfg.Nref           = current_config.Nref;
    seg_cfg.nof_layers     = current_config.nof_layers;
    seg_cfg.nof_ch_symbols = softbits_count / modulation_order;

    span<const log_likelihood_ratio> llrs    = span<const log_likelihood_ratio>(softbits_buffer).first(softbits_count);
    const unsigned                   tb_size = transport_block.size() * 8;
    codeblocks.clear();
    segmenter->segment(codeblocks, llrs, tb_size, seg_cfg);
    const unsigned nof_cbs = codeblocks.size();
    srsran_assert(nof_cbs == softbuffer->get_nof_codeblocks(), "Wrong number of codeblocks.");

    span<bool> cb_crcs = softbuffer->get_codeblocks_crc();
    if (current_config.new_data) {
      std::fill(cb_crcs.begin(), cb_crcs.end(), false); // pusch_decoder_hw_impl.cpp:205-208
    }

    // CB CRC: CRC24B in a segmented TB, else CRC24A above 3824 bits, else CRC16 (pusch_decoder_hw_impl.cpp:78-94).
    const uint8_t crc_type = (nof_cbs > 1) ? PDC_CRC24B : ((ldpc::compute_tb_crc_size(units::bits(tb_size)).value() == 24) ? PDC_CRC24A : PDC_CRC16);
    const int8_t* base     = reinterpret_cast<const int8_t*>(llrs.data());
    descs.clear();
    for (unsigned cb = 0; cb != nof_cbs; ++cb) {
      const codeblock_metadata& m = codeblocks[cb].second;
      pdc_cb_desc               d = {};
      d.bg         = (m.tb_common.base_graph == ldpc_base_graph_type::BG1) ? 1 : 2;
      d.Qm         = modulation_order;
      d.rv         = m.tb_common.rv;
      d.crc_type   = crc_type;
      d.max_iter   = (current_config.nof_ldpc_iterations > 255) ? 255 : current_config.nof_ldpc_iterations;
      d.Z          = static_cast<uint16_t>(m.tb_common.lifting_size);
      d.Nref       = m.tb_common.Nref;
      d.nof_filler = m.cb_specific.nof_filler_bits;
      d.harq_slot  = static_cast<int32_t>(softbuffer->get_absolute_codeblock_id(cb));
      if (cb_crcs[cb]) {
        // Decoded in an earlier transmission: not enqueued again (pusch_decoder_hw_impl.cpp:236-243); its bits wait in
        // the accelerator next to its soft bits.
        d.flags = PDC_F_DECODED;
      } else {
        d.flags = (current_config.new_data ? PDC_F_NEW_DATA : 0) | (current_config.use_early_stop ? PDC_F_EARLY_STOP : 0) |
                  (current_config.force_decoding ? PDC_F_FORCE_DECODING : 0) | (generic ? PDC_F_GENERIC : 0);
        d.E          = codeblocks[cb].first.size();
        d.llr_offset = reinterpret_cast<const int8_t*>(codeblocks[cb].first.data()) - base;
      }
      descs.push_back(d);
    }
    pdc_tb_desc tb = {};
    tb.first_cb = 0, tb.nof_cbs = nof_cbs, tb.tbs_bits = tb_size, tb.tb_offset = 0;

    pusch_decoder_result stats;
    stats.tb_crc_ok            = false;
    stats.nof_codeblocks_total = nof_cbs;
    stats.ldpc_decoder_stats.reset();

    pdc_ticket    ticket = 0;
    pdc_tb_result tb_res = {};
    tb_bytes.resize((transport_block.size() + 3) / 4 * 4);
    // PDC_ERR_CAPACITY = every submission slot of the context is in flight (more decoder instances at work than
    // nof_instances, or pdc_decode_device borrowing one): back-pressure, not a failure. The reference never drops a
    // transport block for pool contention, so wait for a slot like the other adapters do.
    pdc_status st;
    while ((st = pdc_submit_tbs(context->get(), descs.data(), nof_cbs, &tb, 1, base, softbits_count, nullptr, 0, &ticket)) ==
           PDC_ERR_CAPACITY) {
      std::this_thread::yield();
    }
    if (st == PDC_OK) {
      st = pdc_fetch_tbs(context->get(), ticket, tb_bytes.data(), tb_bytes.size(), &tb_res, results.data(), nullptr, 0);
    }
    if (st != PDC_OK) {
      // Never throws on the data path: the transport block is reported as not decoded and stays in its HARQ process.
      std::fprintf(stderr, "pusch_decoder_cuda: transport block of %u code blocks failed (status %d: %s)\n", nof_cbs, (int)st,
                   pdc_last_error(context->get()));
      softbuffer.unlock();
      result_notifier->on_sch_data(stats);
      return;
    }
    for (unsigned cb = 0; cb != nof_cbs; ++cb) {
      if (descs[cb].flags & PDC_F_DECODED) {
        continue;
      }
      // pusch_decoder_hw_impl.cpp:384-391: iterations of a passing code block, the maximum for a failing one.
      cb_crcs[cb] = results[cb].crc_pass != 0;
      stats.ldpc_decoder_stats.update(cb_crcs[cb] ? static_cast<unsigned>(results[cb].iterations)
                                                  : current_config.nof_ldpc_iterations);
    }
    if (tb_res.written) {
      std::memcpy(transport_block.data(), tb_bytes.data(), transport_block.size()); // :406, :417
    }
    stats.tb_crc_ok = tb_res.tb_crc_ok != 0;
    if (tb_res.written && !stats.tb_crc_ok) {
      softbuffer->reset_codeblocks_crc(); // a false positive among the code blocks: decode all of them again (:424-427)
    }
    if (stats.tb_crc_ok) {
      // The transport block is done: its soft bits are no longer needed (pusch_decoder_hw_impl.cpp:413-428 frees the
      // accelerator's HARQ context entries here). The arena slots read as zeros from now on (host-side note, no CUDA call).
      for (unsigned cb = 0; cb != nof_cbs; ++cb) {
        pdc_harq_release(context->get(), descs[cb].harq_slot);
      }
      softbuffer.release();
    } else {
      softbuffer.unlock();
    }
    result_notifier->on_sch_data(stats);
  }

  std::unique_ptr<ldpc_segmenter_rx>                       segmenter;
  std::shared_ptr<cuda::pusch_dec_context>                 context;
  task_executor*                                           executor;
  bool                                                     generic;
  std::vector<log_likelihood_ratio>                        softbits_buffer;
  unsigned                                                 softbits_count = 0;
  span<uint8_t>                                            transport_block;
  unique_rx_buffer                                         softbuffer;
  pusch_decoder_notifier*                                  result_notifier = nullptr;
  pusch_decoder::configuration                             current_config;
  static_vector<described_rx_codeblock, MAX_NOF_SEGMENTS>  codeblocks;
  std::vector<pdc_cb_desc>                                 descs;
  std::vector<pdc_cb_result>                               results;
  std::vector<uint8_t>                                     tb_bytes;
};

class pusch_decoder_factory_cuda : public pusch_decoder_factory
{
public:
  explicit pusch_decoder_factory_cuda(const pusch_decoder_factory_cuda_configuration& cfg_) : cfg(cfg_)
  {
    if (!cfg.segmenter_factory || !cfg.context) {
      throw std::runtime_error("pusch_decoder_factory_cuda: a segmenter factory and an accelerator context are required");
    }
    if (cfg.max_codeword_softbits == 0) {
      cfg.max_codeword_softbits = pdsch_constants::CODEWORD_MAX_SIZE.value(); // as pusch_decoder_hw_impl.h:89
    }
  }
  std::unique_ptr<pusch_decoder> create() override
  {
    return std::make_unique<pusch_decoder_cuda_impl>(
        cfg.segmenter_factory->create(), cfg.context, cfg.executor, cfg.max_codeword_softbits, cfg.generic_arithmetic);
  }

private:
  pusch_decoder_factory_cuda_configuration cfg;
};

} // namespace

std::shared_ptr<pusch_decoder_factory>
srsran::create_pusch_decoder_factory_cuda(const pusch_decoder_factory_cuda_configuration& config)
{
  return std::make_shared<pusch_decoder_factory_cuda>(config);
}
