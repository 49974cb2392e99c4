// CUDA implementation of the reference's hardware-accelerated PUSCH decoder HAL interface
// (include/srsran/hal/phy/upper/channel_processors/pusch/hw_accelerator_pusch_dec.h:83-115), the drop-in that
// pusch_decoder_hw_impl (lib/phy/upper/channel_processors/pusch/pusch_decoder_hw_impl.cpp:149-364) drives in place of the
// bbdev / ACC100 accelerator (whose implementation is not part of the open-source tree:
// lib/hal/phy/upper/channel_processors/pusch/hw_accelerator_factories.cpp:28-36 returns nullptr).
#pragma once

#include "cuda_pusch_dec_context.h"
#include "pusch_dec_batcher.h"
#include "srsran/hal/phy/upper/channel_processors/pusch/hw_accelerator_pusch_dec_factory.h"
#include <vector>

namespace srsran {
namespace hal {

/// Configuration of the CUDA PUSCH decoder accelerator factory (the counterpart of
/// bbdev_hwacc_pusch_dec_factory_configuration, hw_accelerator_factories.h:33-45).
struct cuda_hwacc_pusch_dec_factory_configuration {
  /// Shared accelerator context of a single-GPU set-up; created on `devices` (default: device 0) when empty.
  std::shared_ptr<cuda::pusch_dec_context> context;
  /// Several GPUs: one context per listed CUDA device is created by the factory (ignored when `context` is given). Code
  /// blocks are placed by their absolute code-block identifier (GPU = id mod G, arena slot = id / G), which is what the
  /// reference's external-HARQ mode keys the soft buffers with (ext_harq_buffer_context_repository.h:36-107): a code
  /// block's retransmissions always reach the GPU that holds its soft buffer, whatever else the transport block carries,
  /// and the code blocks of one transport block are decoded by all GPUs at once. No collective, no peer traffic.
  std::vector<int> devices;
  /// Keeps the HARQ soft buffers in the device arena, keyed by the absolute code-block identifier (the reference's
  /// "external HARQ" mode, ext_harq_buffer_context_repository.h:36-107). When false the soft buffer travels from and to
  /// the host with every operation, as with force_local_harq of the bbdev factory.
  bool external_harq = true;
  /// Reproduce ldpc_decoder_generic instead of ldpc_decoder_avx512 / _avx2.
  bool generic_arithmetic = false;
  /// Capacity of the HARQ arena in code blocks over all GPUs (rx_buffer_pool_config::nof_codeblocks); used when the
  /// factory creates the contexts.
  unsigned harq_context_size = 16 * 152;
  /// Cross-instance batching (pusch_dec_batcher.h): transport blocks per GPU launch, closing time, staging buffers.
  cuda::pusch_dec_batcher_config batching;
};

/// \brief Creates a CUDA PUSCH decoder hardware accelerator factory.
/// \return A valid factory. Throws std::runtime_error if no CUDA device is usable (there is no CPU fallback).
std::shared_ptr<hw_accelerator_pusch_dec_factory>
create_cuda_pusch_dec_acc_factory(const cuda_hwacc_pusch_dec_factory_configuration& accelerator_config);

/// Batches submitted / code blocks carried so far by the accelerators of a factory created above (tests, bench).
void cuda_pusch_dec_acc_factory_stats(const hw_accelerator_pusch_dec_factory& factory, uint64_t& batches, uint64_t& code_blocks);
/// Where the time went (pusch_dec_batcher::timing summed over the GPUs): parts, batches, then nanoseconds of queueing (per part),
/// submit, fetch (per batch), wake-up (per part), join wait, add (copy into page-locked memory).
void cuda_pusch_dec_acc_factory_timing(const hw_accelerator_pusch_dec_factory& factory, uint64_t out[8]);

} // namespace hal
} // namespace srsran
