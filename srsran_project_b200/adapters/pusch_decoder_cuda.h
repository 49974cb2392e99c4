// CUDA implementation of the reference's transport-block decoder interface srsran::pusch_decoder
// (include/srsran/phy/upper/channel_processors/pusch/pusch_decoder.h:52-103): segmentation on the host (metadata only), then
// ONE accelerator submission per transport block that dematches, decodes, checks every code block AND assembles the transport
// block with its CRC24A on the device (pdc_submit_tbs). It replaces pusch_decoder_impl / pusch_decoder_hw_impl
// (lib/phy/upper/channel_processors/pusch/pusch_decoder_impl.cpp:89-505, pusch_decoder_hw_impl.cpp:96-449) as a whole; the
// selection site is the same as theirs, upper_phy_factories.cpp:788-798.
//
// HARQ state lives in the accelerator (soft bits AND decoded bits of every code block, keyed by the absolute code-block
// identifier of the rx_buffer pool, which must be created with external_soft_bits = true, rx_buffer_codeblock_pool.h:62-70);
// the rx_buffer keeps what it keeps for the reference's accelerator path: the per-code-block CRC flags.
#pragma once

#include "cuda_pusch_dec_context.h"
#include "srsran/phy/upper/channel_coding/channel_coding_factories.h"
#include "srsran/phy/upper/channel_processors/pusch/factories.h"

namespace srsran {

/// Configuration of the CUDA PUSCH decoder factory (counterpart of pusch_decoder_factory_hw_configuration, factories.h:75-81).
struct pusch_decoder_factory_cuda_configuration {
  /// Codeword segmenter (metadata only; the reference's ldpc_segmenter_rx is used as it is).
  std::shared_ptr<ldpc_segmenter_rx_factory> segmenter_factory;
  /// Shared accelerator context (one per GPU). Its harq_codeblocks must cover the rx_buffer pool's nof_codeblocks.
  std::shared_ptr<cuda::pusch_dec_context> context;
  /// Optional executor for the submission + wait (as pusch_decoder_hw_impl does, :149-163); nullptr = caller's thread.
  task_executor* executor = nullptr;
  /// Largest codeword in soft bits (pdsch_constants::CODEWORD_MAX_SIZE by default, as pusch_decoder_hw_impl.h:89).
  unsigned max_codeword_softbits = 0;
  /// Reproduce ldpc_decoder_generic instead of ldpc_decoder_avx512 / _avx2.
  bool generic_arithmetic = false;
};

/// \brief Creates a PUSCH decoder factory whose decoders run on the CUDA accelerator, transport block assembly included.
/// Throws std::runtime_error when the configuration is unusable (no context: there is no CPU fallback).
std::shared_ptr<pusch_decoder_factory> create_pusch_decoder_factory_cuda(const pusch_decoder_factory_cuda_configuration& config);

} // namespace srsran
