// demodulation_mapper over pdc_demodulate (include/pusch_dec_cuda.h). Synchronous like the reference's demodulate_soft: copy
// in, one kernel, copy out. A deployment that keeps the equaliser outputs on the device calls pdc_demodulate_device instead
// and hands the LLRs to pdc_decode_device without the PCIe hop.
#include "channel_modulation_factories_cuda.h"
#include "srsran/support/srsran_assert.h"

using namespace srsran;

namespace {

uint8_t to_pdc_mod(modulation_scheme mod)
{
  switch (mod) {
    case modulation_scheme::PI_2_BPSK:
      return PDC_MOD_PI2_BPSK;
    case modulation_scheme::BPSK:
      return PDC_MOD_BPSK;
    case modulation_scheme::QPSK:
      return PDC_MOD_QPSK;
    case modulation_scheme::QAM16:
      return PDC_MOD_QAM16;
    case modulation_scheme::QAM64:
      return PDC_MOD_QAM64;
    case modulation_scheme::QAM256:
      return PDC_MOD_QAM256;
    default:
      srsran_assertion_failure("Invalid modulation scheme."); // demodulation_mapper_impl.cpp:103-104
      return PDC_MOD_QPSK;
  }
}

class demodulation_mapper_cuda : public demodulation_mapper
{
public:
  explicit demodulation_mapper_cuda(std::shared_ptr<cuda::pusch_dec_context> context_) : context(std::move(context_)) {}

  // See interface for documentation.
  void demodulate_soft(span<log_likelihood_ratio> llrs,
                       span<const cf_t>           symbols,
                       span<const float>          noise_vars,
                       modulation_scheme          mod) override
  {
    // Same contracts as demodulation_mapper_impl.cpp:84-85.
    srsran_assert(symbols.size() == noise_vars.size(), "Inputs symbols and noise_vars must have the same length.");
    srsran_assert(symbols.size() * get_bits_per_symbol(mod) == llrs.size(), "Input and output lengths are incompatible.");
    if (symbols.empty()) {
      return;
    }
    pdc_demod_seg seg    = {};
    seg.nof_symbols      = static_cast<uint32_t>(symbols.size());
    seg.mod              = to_pdc_mod(mod);
    const pdc_status st  = pdc_demodulate(context->get(),
                                         &seg,
                                         1,
                                         reinterpret_cast<const float*>(symbols.data()),
                                         noise_vars.data(),
                                         symbols.size(),
                                         reinterpret_cast<int8_t*>(llrs.data()),
                                         llrs.size());
    srsran_assert(st == PDC_OK, "pdc_demodulate failed ({}): {}", static_cast<int>(st), pdc_last_error(context->get()));
  }

private:
  std::shared_ptr<cuda::pusch_dec_context> context;
};

class demodulation_mapper_factory_cuda : public demodulation_mapper_factory
{
public:
  explicit demodulation_mapper_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context_) :
    context(std::move(context_))
  {
  }
  std::unique_ptr<demodulation_mapper> create() override { return std::make_unique<demodulation_mapper_cuda>(context); }

private:
  std::shared_ptr<cuda::pusch_dec_context> context;
};

} // namespace

std::shared_ptr<demodulation_mapper_factory>
srsran::create_demodulation_mapper_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context)
{
  return std::make_shared<demodulation_mapper_factory_cuda>(std::move(context));
}

std::shared_ptr<demodulation_mapper_factory> srsran::create_demodulation_mapper_factory_cuda()
{
  return create_demodulation_mapper_factory_cuda(
      std::make_shared<cuda::pusch_dec_context>(cuda::pusch_dec_context_config{}));
}
