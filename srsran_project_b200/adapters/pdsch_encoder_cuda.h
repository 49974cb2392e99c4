// CUDA srsran::pdsch_encoder (the downlink mirror): the whole of pdsch_encoder_impl::encode — TB CRC,
// segmentation, CB CRC, LDPC encoding, rate matching, concatenation — as one submission to the device. It sits beside
// create_pdsch_encoder_factory_sw / _hw (include/srsran/phy/upper/channel_processors/pdsch/factories.h) and returns objects
// with the interface of pdsch_encoder.h:43-74.
#pragma once

#include "cuda_pusch_dec_context.h"
#include "srsran/phy/upper/channel_processors/pdsch/factories.h"
#include "srsran/phy/upper/channel_processors/pdsch/pdsch_encoder.h"
#include <memory>

namespace srsran {

/// A pdsch_encoder_factory (factories.h:42-47): what create_pdsch_processor_factory_sw and friends take.
class pdsch_encoder_factory_cuda : public pdsch_encoder_factory
{
public:
  explicit pdsch_encoder_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context_) : context(std::move(context_)) {}
  std::unique_ptr<pdsch_encoder> create() override;

private:
  std::shared_ptr<cuda::pusch_dec_context> context;
};

/// Creates the factory; throws std::runtime_error if no CUDA device is usable (there is no CPU fallback).
std::shared_ptr<pdsch_encoder_factory> create_pdsch_encoder_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context);
std::shared_ptr<pdsch_encoder_factory> create_pdsch_encoder_factory_cuda();

} // namespace srsran
