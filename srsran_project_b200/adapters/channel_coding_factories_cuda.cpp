// ldpc_decoder / ldpc_rate_dematcher implementations that run on a CUDA device through the C ABI of
// include/pusch_dec_cuda.h. Synchronous, one code block per call, exactly as the reference's software operators:
//   srsran::ldpc_decoder::decode              include/srsran/phy/upper/channel_coding/ldpc/ldpc_decoder.h:37-73
//   srsran::ldpc_rate_dematcher::rate_dematch include/srsran/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher.h:35-57
// Contract violations that the reference guards with srsran_assert are guarded with srsran_assert here too (the C ABI
// reports them as PDC_ERR_INVALID_ARG); a decode failure is std::nullopt, never an exception.
#include "channel_coding_factories_cuda.h"
#include "pusch_dec_batcher.h"
#include "srsran/phy/upper/channel_coding/crc_calculator.h"
#include "srsran/phy/upper/channel_coding/ldpc/ldpc_decoder.h"
#include "srsran/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher.h"
#include "srsran/support/srsran_assert.h"
#include <algorithm>
#include <cstring>
#include <thread>
#include <vector>

using namespace srsran;

namespace {

uint8_t crc_type_of(crc_generator_poly poly)
{
  switch (poly) {
    case crc_generator_poly::CRC16:
      return PDC_CRC16;
    case crc_generator_poly::CRC24A:
      return PDC_CRC24A;
    case crc_generator_poly::CRC24B:
      return PDC_CRC24B;
    default:
      srsran_assert(false, "CUDA LDPC decoder: only CRC16, CRC24A and CRC24B can drive the early stop.");
      return PDC_CRC_NONE;
  }
}

// Runs one code block and waits for it. Instances are pooled and driven concurrently by the reference
// (bounded_unique_object_pool, factories.cpp:82-94; pusch_decoder_impl forks one task per code block,
// pusch_decoder_impl.cpp:309-394): the calls of all threads that arrive while the device is busy travel as ONE submission
// (pusch_dec_batcher.h), each caller picks its own code block out of the batch.
pdc_status run_one(cuda::pusch_dec_batcher& batcher,
                   const pdc_cb_desc&       desc,
                   const int8_t*            llrs,
                   size_t                   nof_llrs,
                   int8_t*                  soft,
                   size_t                   nof_soft,
                   uint8_t*                 bits,
                   size_t                   nof_bytes,
                   pdc_cb_result&           res)
{
  cuda::pusch_dec_batcher::part* p;
  while ((p = batcher.join(1, nof_llrs + 16, 200)) == nullptr) {
    // every staging buffer is taken by other threads: the call slept until one was recycled (or 200 us)
  }
  if (!batcher.add(p, desc, llrs, nof_llrs, soft, nof_soft)) {
    batcher.release(p);
    return PDC_ERR_CAPACITY; // a code block larger than the context was created for
  }
  batcher.seal(p);
  while (batcher.poll(p, 100) != cuda::pusch_dec_batcher::status::done) {
  }
  res = batcher.result(p, 0);
  if (bits && res.status == 0 && !(desc.flags & PDC_F_DEMATCH_ONLY)) {
    std::memcpy(bits, batcher.bits(p, 0), std::min<size_t>(nof_bytes, (pdc_cb_message_bits(desc.bg, desc.Z) + 7) / 8));
  }
  if (soft && desc.harq_slot < 0 && !(desc.flags & PDC_F_DECODE_ONLY) && res.status == 0) {
    std::memcpy(soft, batcher.soft(p, 0), std::min<size_t>(nof_soft, pdc_cb_buffer_len(desc.bg, desc.Z)));
  }
  const pdc_status st = (res.status == 1) ? PDC_ERR_INVALID_ARG : PDC_OK; // 1: rejected or failed submission
  batcher.release(p);
  return st;
}

class ldpc_decoder_cuda : public ldpc_decoder
{
public:
  ldpc_decoder_cuda(std::shared_ptr<cuda::pusch_dec_context> context_, bool force_decoding_, bool generic_) :
    context(std::move(context_)),
    batcher(cuda::shared_batcher_of(context)),
    force_decoding(force_decoding_),
    generic(generic_),
    bits((PDC_MAX_K_BYTES + 3) / 4 * 4)
  {
  }

  std::optional<unsigned> decode(bit_buffer&                      output,
                                 span<const log_likelihood_ratio> input,
                                 crc_calculator*                  crc,
                                 const configuration&             cfg) override
  {
    const int      bg = (cfg.base_graph == ldpc_base_graph_type::BG1) ? 1 : 2;
    const unsigned Z  = cfg.lifting_size;
    const unsigned K  = pdc_cb_message_bits(bg, Z);
    const unsigned N  = pdc_cb_buffer_len(bg, Z);
    srsran_assert(K != 0, "Invalid base graph / lifting size.");
    // Same size contract as ldpc_decoder_impl::decode (ldpc_decoder_impl.cpp:64-80).
    srsran_assert(output.size() == K, "The output size {} is not equal to the message length {}.", output.size(), K);
    srsran_assert((input.size() >= K + 2 * Z) && (input.size() <= N),
                  "The input size {} is not valid for message length {} and lifting size {}.",
                  input.size(),
                  K,
                  Z);
    srsran_assert(cfg.max_iterations > 0, "Max iterations must be different to 0");

    pdc_cb_desc d = {};
    d.bg          = bg;
    d.Qm          = 1;
    d.rv          = 0;
    d.crc_type    = (crc != nullptr) ? crc_type_of(crc->get_generator_poly()) : PDC_CRC_NONE;
    d.max_iter    = (cfg.max_iterations > 255) ? 255 : cfg.max_iterations;
    d.flags       = PDC_F_DECODE_ONLY | ((crc != nullptr) ? PDC_F_EARLY_STOP : 0) |
              (force_decoding ? PDC_F_FORCE_DECODING : 0) | (generic ? PDC_F_GENERIC : 0);
    d.Z          = Z;
    d.E          = input.size();
    d.nof_filler = cfg.nof_filler_bits;
    d.harq_slot  = -1;

    pdc_cb_result res = {};
    pdc_status    st  = run_one(*batcher,
                            d,
                            reinterpret_cast<const int8_t*>(input.data()),
                            input.size(),
                            nullptr,
                            0,
                            bits.data(),
                            bits.size(),
                            res);
    srsran_assert(st == PDC_OK, "CUDA LDPC decoder failed: status {} ({}).", (int)st, pdc_last_error(context->get()));

    // status 2: not enough soft bits and a CRC calculator was given: the reference returns without touching the output
    // (ldpc_decoder_impl.cpp:91-97).
    if (res.status == 0) {
      span<uint8_t> out_bytes  = output.get_buffer();
      unsigned      full_bytes = K / 8;
      std::memcpy(out_bytes.data(), bits.data(), full_bytes);
      if (K % 8 != 0) {
        // Keep the bits of the last byte that do not belong to the message.
        uint8_t mask          = static_cast<uint8_t>(0xff00u >> (K % 8));
        out_bytes[full_bytes] = (out_bytes[full_bytes] & ~mask) | (bits[full_bytes] & mask);
      }
    }
    if (res.iterations > 0) {
      return static_cast<unsigned>(res.iterations);
    }
    return std::nullopt;
  }

private:
  static constexpr size_t PDC_MAX_K_BYTES = 8448 / 8;

  std::shared_ptr<cuda::pusch_dec_context> context;
  std::shared_ptr<cuda::pusch_dec_batcher> batcher;
  bool                                     force_decoding;
  bool                                     generic;
  std::vector<uint8_t>                     bits;
};

class ldpc_rate_dematcher_cuda : public ldpc_rate_dematcher
{
public:
  ldpc_rate_dematcher_cuda(std::shared_ptr<cuda::pusch_dec_context> context_, bool generic_) :
    context(std::move(context_)), batcher(cuda::shared_batcher_of(context)), generic(generic_)
  {
  }

  void rate_dematch(span<log_likelihood_ratio>       output,
                    span<const log_likelihood_ratio> input,
                    bool                             new_data,
                    const codeblock_metadata&        cfg) override
  {
    // The reference infers the base graph from the output length, BG1 first (ldpc_rate_dematcher_impl.cpp:81-96).
    const unsigned N  = output.size();
    int            bg = 0;
    unsigned       Z  = 0;
    if (N % 66 == 0) {
      bg = 1, Z = N / 66;
    } else if (N % 50 == 0) {
      bg = 2, Z = N / 50;
    }
    srsran_assert(bg != 0 && pdc_cb_buffer_len(bg, Z) == N, "LDPC rate dematching: invalid input length.");
    const unsigned Qm = get_bits_per_symbol(cfg.tb_common.mod);
    srsran_assert(input.size() % Qm == 0, "The input length should be a multiple of the modulation order.");

    pdc_cb_desc d = {};
    d.bg          = bg;
    d.Qm          = Qm;
    d.rv          = cfg.tb_common.rv;
    d.crc_type    = PDC_CRC_NONE;
    d.max_iter    = 1;
    d.flags       = PDC_F_DEMATCH_ONLY | (new_data ? PDC_F_NEW_DATA : 0) | (generic ? PDC_F_GENERIC : 0);
    d.Z           = Z;
    d.E           = input.size();
    d.Nref        = cfg.tb_common.Nref;
    d.nof_filler  = cfg.cb_specific.nof_filler_bits;
    d.harq_slot   = -1;

    // The write set of the dematcher does not cover the whole buffer (ldpc_rate_dematcher_impl.cpp:128-201), so the
    // current content always travels to the device and back.
    pdc_cb_result res = {};
    pdc_status    st  = run_one(*batcher,
                            d,
                            reinterpret_cast<const int8_t*>(input.data()),
                            input.size(),
                            reinterpret_cast<int8_t*>(output.data()),
                            output.size(),
                            nullptr,
                            0,
                            res);
    srsran_assert(st == PDC_OK, "CUDA LDPC rate dematcher failed: status {} ({}).", (int)st, pdc_last_error(context->get()));
  }

private:
  std::shared_ptr<cuda::pusch_dec_context> context;
  std::shared_ptr<cuda::pusch_dec_batcher> batcher;
  bool                                     generic;
};

class ldpc_decoder_factory_cuda : public ldpc_decoder_factory
{
public:
  ldpc_decoder_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context_, bool force_decoding_, bool generic_) :
    context(std::move(context_)), force_decoding(force_decoding_), generic(generic_)
  {
  }
  std::unique_ptr<ldpc_decoder> create() override
  {
    return std::make_unique<ldpc_decoder_cuda>(context, force_decoding, generic);
  }

private:
  std::shared_ptr<cuda::pusch_dec_context> context;
  bool                                     force_decoding;
  bool                                     generic;
};

class ldpc_rate_dematcher_factory_cuda : public ldpc_rate_dematcher_factory
{
public:
  ldpc_rate_dematcher_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context_, bool generic_) :
    context(std::move(context_)), generic(generic_)
  {
  }
  std::unique_ptr<ldpc_rate_dematcher> create() override
  {
    return std::make_unique<ldpc_rate_dematcher_cuda>(context, generic);
  }

private:
  std::shared_ptr<cuda::pusch_dec_context> context;
  bool                                     generic;
};

std::shared_ptr<cuda::pusch_dec_context> make_default_context()
{
  cuda::pusch_dec_context_config cfg;
  cfg.max_batch_cbs  = 64;               // concurrent single-code-block calls that may share one submission
  cfg.max_batch_llrs = 16u * 35u * 8448u; // 16 x MAX_CODEBLOCK_RM_SIZE (ldpc.h:122)
  cfg.nof_instances  = 8;
  return std::make_shared<cuda::pusch_dec_context>(cfg);
}

} // namespace

std::shared_ptr<ldpc_decoder_factory>
srsran::create_ldpc_decoder_factory_cuda(const ldpc_decoder_factory::ldpc_decoder_factory_configuration& cfg,
                                         std::shared_ptr<cuda::pusch_dec_context>                        context,
                                         ldpc_decoder_cuda_flavour                                       flavour)
{
  return std::make_shared<ldpc_decoder_factory_cuda>(
      std::move(context), cfg.force_decoding, flavour == ldpc_decoder_cuda_flavour::generic);
}

std::shared_ptr<ldpc_rate_dematcher_factory>
srsran::create_ldpc_rate_dematcher_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context,
                                                ldpc_decoder_cuda_flavour                flavour)
{
  return std::make_shared<ldpc_rate_dematcher_factory_cuda>(std::move(context),
                                                            flavour == ldpc_decoder_cuda_flavour::generic);
}

std::shared_ptr<ldpc_decoder_factory>
srsran::create_ldpc_decoder_factory_cuda(const ldpc_decoder_factory::ldpc_decoder_factory_configuration& cfg)
{
  return create_ldpc_decoder_factory_cuda(cfg, make_default_context(), ldpc_decoder_cuda_flavour::avx512);
}

std::shared_ptr<ldpc_rate_dematcher_factory> srsran::create_ldpc_rate_dematcher_factory_cuda()
{
  return create_ldpc_rate_dematcher_factory_cuda(make_default_context(), ldpc_decoder_cuda_flavour::avx512);
}
