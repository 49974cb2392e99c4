// hal::hw_accelerator_pdsch_enc over pdc_pdsch_encode (include/pusch_dec_cuda.h), transport-block mode.
//
// Call sequence of pdsch_encoder_hw_impl::encode (pdsch_encoder_hw_impl.cpp:31-186) in TB mode:
//   reserve_queue(); configure_operation(hw_cfg, 0); enqueue_operation(transport_block, {}, 0);
//   dequeue_operation(codeword /* one bit per byte */, codeblock_packed, 0); free_queue();
// hw_cfg carries the segmentation the reference's segmenter computed (nof_segments, lifting_size, cw_length_a / _b,
// nof_short_segments: pdsch_encoder_hw_impl.cpp:188-276). The device derives C and Z itself from (base graph, TBS); the
// rate-matched lengths are taken from hw_cfg since it has no layer count, and nof_segments / lifting_size are
// cross-checked against what the device will use.
#include "hw_accelerator_pdsch_enc_cuda.h"
#include "srsran/support/srsran_assert.h"
#include <vector>

using namespace srsran;
using namespace hal;

namespace {

class hw_accelerator_pdsch_enc_cuda : public hw_accelerator_pdsch_enc
{
public:
  explicit hw_accelerator_pdsch_enc_cuda(std::shared_ptr<cuda::pusch_dec_context> context_) : context(std::move(context_)) {}

  // See hw_accelerator interface for the documentation.
  bool enqueue_operation(span<const uint8_t> data, span<const uint8_t> /*aux_data*/, unsigned cb_index) override
  {
    srsran_assert(configured, "enqueue_operation called before configure_operation.");
    srsran_assert(cb_index == 0, "The CUDA PDSCH encoder works in TB mode: one operation per transport block.");
    if (pending) {
      return false; // One operation in flight per instance: the caller dequeues and retries (pdsch_encoder_hw_impl.cpp:105-119).
    }
    srsran_assert(data.size() * 8 == config.nof_tb_bits, "Transport block of {} bytes, configured for {} bits.", data.size(), config.nof_tb_bits);

    pdc_pdsch_tb_desc d  = {};
    d.bg                 = (config.base_graph_index == ldpc_base_graph_type::BG1) ? 1 : 2;
    d.Qm                 = static_cast<uint8_t>(get_bits_per_symbol(config.modulation));
    d.rv                 = static_cast<uint8_t>(config.rv);
    d.tbs_bits           = config.nof_tb_bits;
    d.nof_ch_symbols     = 0; // explicit rate-matched lengths
    d.Nref               = config.Nref;
    d.cw_length_a        = config.cw_length_a;
    d.cw_length_b        = (config.nof_short_segments < config.nof_segments) ? config.cw_length_b : 0;
    d.nof_short_segments = config.nof_short_segments;

    const size_t cw_bits = static_cast<size_t>(config.nof_short_segments) * config.cw_length_a +
                           static_cast<size_t>(config.nof_segments - config.nof_short_segments) * config.cw_length_b;
    codeword.resize(cw_bits);
    const pdc_status st = pdc_pdsch_encode(context->get(), &d, 1, data.data(), data.size(), codeword.data(), codeword.size());
    srsran_assert(st == PDC_OK, "pdc_pdsch_encode failed ({}): {}", static_cast<int>(st), pdc_last_error(context->get()));
    pending = true;
    return true;
  }

  // See hw_accelerator interface for the documentation.
  bool dequeue_operation(span<uint8_t> data, span<uint8_t> packed_data, unsigned segment_index) override
  {
    if (!pending || segment_index != 0) {
      return false;
    }
    if (config.do_unpack) {
      srsran_assert(data.size() == codeword.size(), "Codeword span of {} bits, {} encoded.", data.size(), codeword.size());
      std::copy(codeword.begin(), codeword.end(), data.begin());
    }
    if (!packed_data.empty()) {
      // Packed codeword, MSB first (what the bbdev accelerator writes and pdsch_encoder_hw_impl sizes codeblock_packed for).
      srsran_assert(packed_data.size() * 8 >= codeword.size(), "Packed codeword span too small.");
      const size_t nfull = codeword.size() / 8;
      for (size_t i = 0; i != nfull; ++i) {
        const uint8_t* b = &codeword[8 * i];
        packed_data[i]   = static_cast<uint8_t>((b[0] << 7) | (b[1] << 6) | (b[2] << 5) | (b[3] << 4) | (b[4] << 3) | (b[5] << 2) |
                                              (b[6] << 1) | b[7]);
      }
      if (codeword.size() % 8 != 0) {
        unsigned v = 0;
        for (size_t k = 8 * nfull; k != codeword.size(); ++k) {
          v |= static_cast<unsigned>(codeword[k]) << (7 - (k - 8 * nfull));
        }
        packed_data[nfull] = static_cast<uint8_t>(v);
      }
    }
    pending = false;
    return true;
  }

  // See hw_accelerator_pdsch_enc interface for the documentation.
  void reserve_queue() override {} // Submissions are ordered by the context, there is no per-operation queue to reserve.
  void free_queue() override {}

  void configure_operation(const hw_pdsch_encoder_configuration& config_, unsigned cb_index) override
  {
    srsran_assert(!config_.cb_mode && cb_index == 0, "The CUDA PDSCH encoder works in TB mode.");
    srsran_assert(config_.attach_tb_crc, "TB mode: the accelerator attaches the transport block CRC.");
    srsran_assert(config_.attach_cb_crc == (config_.nof_segments > 1), "TB mode: code-block CRCs go with segmentation.");
    // The device repeats the segmentation (TS 38.212 5.2.2); what the reference's segmenter found must agree.
    const int bg = (config_.base_graph_index == ldpc_base_graph_type::BG1) ? 1 : 2;
    uint32_t  nof_cbs = 0, lifting_size = 0;
    const pdc_status st = pdc_pdsch_segmentation(bg, config_.nof_tb_bits, &nof_cbs, &lifting_size, nullptr, nullptr);
    srsran_assert(st == PDC_OK, "Invalid transport block size {}.", config_.nof_tb_bits);
    srsran_assert(nof_cbs == config_.nof_segments && lifting_size == config_.lifting_size,
                  "Segmentation mismatch: host C={} Z={}, device C={} Z={}.",
                  config_.nof_segments,
                  config_.lifting_size,
                  nof_cbs,
                  lifting_size);
    srsran_assert(config_.nof_short_segments >= 1 && config_.nof_short_segments <= config_.nof_segments, "Invalid number of short segments.");
    config     = config_;
    configured = true;
  }

  bool is_cb_mode_supported() const override { return false; }

  // Large enough for any transport block and packed codeword (pdsch_constants::CODEWORD_MAX_SIZE), so the caller never
  // falls back to CB mode (pdsch_encoder_hw_impl.cpp:36-43).
  unsigned get_max_supported_buff_size() const override { return 1u << 24; }

private:
  std::shared_ptr<cuda::pusch_dec_context> context;
  hw_pdsch_encoder_configuration           config     = {};
  bool                                     configured = false;
  bool                                     pending    = false;
  std::vector<uint8_t>                     codeword;
};

class hw_accelerator_pdsch_enc_factory_cuda : public hw_accelerator_pdsch_enc_factory
{
public:
  explicit hw_accelerator_pdsch_enc_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context_) : context(std::move(context_)) {}
  std::unique_ptr<hw_accelerator_pdsch_enc> create() override { return std::make_unique<hw_accelerator_pdsch_enc_cuda>(context); }

private:
  std::shared_ptr<cuda::pusch_dec_context> context;
};

} // namespace

std::shared_ptr<hw_accelerator_pdsch_enc_factory>
srsran::hal::create_cuda_pdsch_enc_acc_factory(const cuda_hwacc_pdsch_enc_factory_configuration& accelerator_config)
{
  std::shared_ptr<cuda::pusch_dec_context> context = accelerator_config.context;
  if (!context) {
    context = std::make_shared<cuda::pusch_dec_context>(cuda::pusch_dec_context_config{});
  }
  return std::make_shared<hw_accelerator_pdsch_enc_factory_cuda>(std::move(context));
}
