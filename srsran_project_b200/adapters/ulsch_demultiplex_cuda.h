// CUDA srsran::ulsch_demultiplex: the UL-SCH demultiplexer of the PUSCH processor
// (include/srsran/phy/upper/channel_processors/pusch/ulsch_demultiplex.h:36-103, lib/.../ulsch_demultiplex_impl.cpp) over
// pdc_ulsch_demultiplex. It sits beside create_ulsch_demultiplex_factory_sw (pusch/factories.h:155-162) and is what
// pusch_processor_configuration::demux_factory takes.
#pragma once

#include "cuda_pusch_dec_context.h"
#include "srsran/phy/upper/channel_processors/pusch/factories.h"
#include "srsran/phy/upper/channel_processors/pusch/ulsch_demultiplex.h"
#include <memory>

namespace srsran {

/// Creates the factory; throws std::runtime_error if no CUDA device is usable (there is no CPU fallback).
std::shared_ptr<ulsch_demultiplex_factory> create_ulsch_demultiplex_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context);
std::shared_ptr<ulsch_demultiplex_factory> create_ulsch_demultiplex_factory_cuda();

} // namespace srsran
