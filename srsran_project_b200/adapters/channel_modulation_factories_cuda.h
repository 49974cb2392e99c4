// "cuda" flavour of the reference's demodulation-mapper factory. It sits beside
// create_demodulation_mapper_factory (include/srsran/phy/upper/channel_modulation/channel_modulation_factories.h:41-48) and
// returns objects with the same interface (demodulation_mapper.h:47-69), so that pusch_demodulator_impl uses them unchanged.
#pragma once

#include "cuda_pusch_dec_context.h"
#include "srsran/phy/upper/channel_modulation/channel_modulation_factories.h"

namespace srsran {

/// \brief Creates a demodulation mapper factory whose mappers run on a CUDA device.
///
/// Every call of demodulate_soft is one segment of pdc_demodulate: the LLRs are those of the reference built for an AVX512
/// host, bit for bit (the AVX512 / AVX2 / one-by-one loops of demodulation_mapper_{qpsk,qam16,qam64,qam256}.cpp and their
/// remainders). Throws std::runtime_error if no CUDA device is usable (there is no CPU fallback).
std::shared_ptr<demodulation_mapper_factory>
create_demodulation_mapper_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context);

/// Convenience overload with a private context on device 0.
std::shared_ptr<demodulation_mapper_factory> create_demodulation_mapper_factory_cuda();

} // namespace srsran
