// pdsch_encoder over pdc_pdsch_encode (include/pusch_dec_cuda.h): one transport block per call, synchronous like the
// reference's encode(). Same contracts as ldpc_segmenter_tx_impl.cpp:40-55 (asserted here, error codes at the C ABI).
#include "pdsch_encoder_cuda.h"
#include "srsran/support/srsran_assert.h"

using namespace srsran;

namespace {

class pdsch_encoder_cuda : public pdsch_encoder
{
public:
  explicit pdsch_encoder_cuda(std::shared_ptr<cuda::pusch_dec_context> context_) : context(std::move(context_)) {}

  // See interface for documentation.
  void encode(span<uint8_t> codeword, span<const uint8_t> transport_block, const configuration& cfg) override
  {
    srsran_assert(!transport_block.empty(), "Argument transport_block should not be empty.");
    srsran_assert((cfg.nof_layers >= 1) && (cfg.nof_layers <= 4), "Invalid number of layers.");
    srsran_assert(cfg.nof_ch_symbols % cfg.nof_layers == 0,
                  "The number of channel symbols should be a multiple of the product between the number of layers.");
    const unsigned qm = get_bits_per_symbol(cfg.mod);
    srsran_assert(codeword.size() == static_cast<size_t>(cfg.nof_ch_symbols) * qm, "Wrong codeword length.");
    pdc_pdsch_tb_desc d = {};
    d.bg                = (cfg.base_graph == ldpc_base_graph_type::BG1) ? 1 : 2;
    d.Qm                = static_cast<uint8_t>(qm);
    d.rv                = static_cast<uint8_t>(cfg.rv);
    d.nof_layers        = static_cast<uint8_t>(cfg.nof_layers);
    d.tbs_bits          = static_cast<uint32_t>(transport_block.size() * 8);
    d.nof_ch_symbols    = cfg.nof_ch_symbols;
    d.Nref              = cfg.Nref;
    const pdc_status st = pdc_pdsch_encode(
        context->get(), &d, 1, transport_block.data(), transport_block.size(), codeword.data(), codeword.size());
    srsran_assert(st == PDC_OK, "pdc_pdsch_encode failed ({}): {}", static_cast<int>(st), pdc_last_error(context->get()));
  }

private:
  std::shared_ptr<cuda::pusch_dec_context> context;
};

} // namespace

std::unique_ptr<pdsch_encoder> pdsch_encoder_factory_cuda::create()
{
  return std::make_unique<pdsch_encoder_cuda>(context);
}

std::shared_ptr<pdsch_encoder_factory>
srsran::create_pdsch_encoder_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context)
{
  return std::make_shared<pdsch_encoder_factory_cuda>(std::move(context));
}

std::shared_ptr<pdsch_encoder_factory> srsran::create_pdsch_encoder_factory_cuda()
{
  return create_pdsch_encoder_factory_cuda(std::make_shared<cuda::pusch_dec_context>(cuda::pusch_dec_context_config{}));
}
