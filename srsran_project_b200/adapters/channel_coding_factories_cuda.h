// "cuda" flavours of the reference's channel-coding factories. They sit beside create_ldpc_decoder_factory_sw /
// create_ldpc_rate_dematcher_factory_sw (include/srsran/phy/upper/channel_coding/channel_coding_factories.h:53-83) and
// return objects with the same interfaces (ldpc_decoder.h:37-73, ldpc_rate_dematcher.h:35-57), so that
// pusch_codeblock_decoder, the PHY metric decorators and ldpc_decoder_benchmark use them unchanged.
#pragma once

#include "cuda_pusch_dec_context.h"
#include "srsran/phy/upper/channel_coding/channel_coding_factories.h"

namespace srsran {

/// Arithmetic flavour the CUDA decoder reproduces bit for bit. The reference ships numerically different decoders behind
/// one interface: AVX2 == AVX512 (scaling (x * 52428) >> 16) and generic (scaling round(0.8f * x)).
enum class ldpc_decoder_cuda_flavour { avx512, generic };

/// \brief Creates an LDPC decoder factory whose decoders run on a CUDA device.
/// \param[in] cfg     Same configuration structure as the software factories (force_decoding).
/// \param[in] context Shared accelerator context (one per GPU).
/// \param[in] flavour Reference decoder whose results are reproduced.
/// \return A valid factory. Throws std::runtime_error if no CUDA device is usable (there is no CPU fallback).
std::shared_ptr<ldpc_decoder_factory>
create_ldpc_decoder_factory_cuda(const ldpc_decoder_factory::ldpc_decoder_factory_configuration& cfg,
                                 std::shared_ptr<cuda::pusch_dec_context>                        context,
                                 ldpc_decoder_cuda_flavour flavour = ldpc_decoder_cuda_flavour::avx512);

/// \brief Creates an LDPC rate dematcher factory whose dematchers run on a CUDA device.
std::shared_ptr<ldpc_rate_dematcher_factory>
create_ldpc_rate_dematcher_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context,
                                        ldpc_decoder_cuda_flavour flavour = ldpc_decoder_cuda_flavour::avx512);

/// Convenience overloads that create a private context on device 0 (what a string-keyed
/// create_ldpc_decoder_factory_sw("cuda", cfg) branch would call).
std::shared_ptr<ldpc_decoder_factory>
create_ldpc_decoder_factory_cuda(const ldpc_decoder_factory::ldpc_decoder_factory_configuration& cfg);
std::shared_ptr<ldpc_rate_dematcher_factory> create_ldpc_rate_dematcher_factory_cuda();

} // namespace srsran
