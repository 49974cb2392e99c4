// CUDA implementation of the reference's hardware-accelerated PDSCH encoder HAL interface
// (include/srsran/hal/phy/upper/channel_processors/hw_accelerator_pdsch_enc.h:81-106), the drop-in that
// pdsch_encoder_hw_impl (lib/phy/upper/channel_processors/pdsch/pdsch_encoder_hw_impl.cpp:31-186) drives in place of the
// bbdev / ACC100 accelerator. It works in transport-block mode: one enqueue carries the whole transport block, the device
// attaches the CRCs, segments, encodes, rate-matches and concatenates (pdc_pdsch_encode, include/pusch_dec_cuda.h), one
// dequeue hands back the codeword, unpacked and packed.
#pragma once

#include "cuda_pusch_dec_context.h"
#include "srsran/hal/phy/upper/channel_processors/hw_accelerator_pdsch_enc_factory.h"

namespace srsran {
namespace hal {

/// Configuration of the CUDA PDSCH encoder accelerator factory (the counterpart of
/// bbdev_hwacc_pdsch_enc_factory_configuration, hw_accelerator_factories.h).
struct cuda_hwacc_pdsch_enc_factory_configuration {
  /// Shared accelerator context (one per GPU); created on device 0 with defaults when empty.
  std::shared_ptr<cuda::pusch_dec_context> context;
};

/// \brief Creates a CUDA PDSCH encoder hardware accelerator factory.
/// \return A valid factory. Throws std::runtime_error if no CUDA device is usable (there is no CPU fallback).
std::shared_ptr<hw_accelerator_pdsch_enc_factory>
create_cuda_pdsch_enc_acc_factory(const cuda_hwacc_pdsch_enc_factory_configuration& accelerator_config);

} // namespace hal
} // namespace srsran
