// ulsch_demultiplex over pdc_ulsch_demultiplex (include/pusch_dec_cuda.h).
//
// The reference demultiplexes OFDM symbol by OFDM symbol while the demodulator streams blocks in (on_new_block). Here the
// blocks of a codeword are gathered (the demodulator writes straight into the gathering buffer through get_next_block_view) and
// the whole codeword is demultiplexed on the device at on_end_codeword; the four consumers then get their streams in the
// reference's order of completion (HARQ-ACK, CSI Part 1, CSI Part 2, UL-SCH). What a consumer receives is bit-identical; only
// the moment differs (its soft bits arrive at the end of the codeword instead of after each OFDM symbol).
//
// CSI Part 2: set_csi_part2 called before the first block means "known from the start" (the reference's unit test). Called
// later, it can only come from the CSI Part 1 feedback (pusch_processor_impl.cpp:73-98), which fires when the CSI Part 1 stream
// ends: the codeword is then demultiplexed a second time with PDC_DEMUX_F_CSI2_AFTER_CSI1, which reproduces what the reference
// does when the call arrives inside the OFDM symbol that completes CSI Part 1.
//
// The interface hands over the scrambling sequence, not its seed, and the device regenerates the sequence from c_init for the
// placeholder bits: the seed is recovered from the sequence itself (c = x1 ^ x2, x1 is fixed, and the x2 register at position
// 1600 runs backwards to position 0, TS 38.211 5.2.1).
#include "ulsch_demultiplex_cuda.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_codeword_buffer.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_buffer.h"
#include "srsran/support/srsran_assert.h"
#include <cstring>
#include <vector>

using namespace srsran;

namespace {

// First 31 bits of x1(1600 + n).
uint32_t x1_at_1600()
{
  uint64_t x = 1; // bit i = x1(p + i)
  for (unsigned p = 0; p != 1600; ++p) {
    const uint64_t nb = ((x >> 3) ^ x) & 1u;
    x                 = (x >> 1) | (nb << 30);
  }
  return static_cast<uint32_t>(x & 0x7fffffffu);
}

// c_init whose sequence starts with the given bits (bit i of `first` = c(i), nof_bits <= 31 of them known; unknown ones taken as 0,
// which yields a seed that still reproduces every known bit).
uint32_t recover_c_init(uint32_t first, unsigned nof_bits)
{
  static const uint32_t x1 = x1_at_1600();
  const uint32_t        known = (nof_bits >= 31) ? 0x7fffffffu : ((1u << nof_bits) - 1u);
  uint64_t              x     = (first ^ x1) & known; // bit i = x2(1600 + i)
  for (unsigned p = 0; p != 1600; ++p) {
    // x2(n) = x2(n + 31) ^ x2(n + 3) ^ x2(n + 2) ^ x2(n + 1): with bit i = x2(p + i), the bit before the window is
    // b30 ^ b2 ^ b1 ^ b0.
    const uint64_t pb = ((x >> 30) ^ (x >> 2) ^ (x >> 1) ^ x) & 1u;
    x                 = ((x << 1) | pb) & 0x7fffffffu;
  }
  return static_cast<uint32_t>(x);
}

class ulsch_demultiplex_cuda : public ulsch_demultiplex, private pusch_codeword_buffer
{
public:
  explicit ulsch_demultiplex_cuda(std::shared_ptr<cuda::pusch_dec_context> context_) : context(std::move(context_)) {}

  // See ulsch_demultiplex interface for documentation.
  void set_csi_part2(pusch_decoder_buffer& csi_part2_, unsigned nof_csi_part2_bits_, unsigned nof_csi_part2_enc_bits_) override
  {
    csi_part2                   = &csi_part2_;
    desc.nof_csi_part2_bits     = nof_csi_part2_bits_;
    desc.nof_enc_csi_part2_bits = nof_csi_part2_enc_bits_;
    desc.flags                  = (count != 0) ? PDC_DEMUX_F_CSI2_AFTER_CSI1 : 0;
  }

  pusch_codeword_buffer& demultiplex(pusch_decoder_buffer& sch_data_,
                                     pusch_decoder_buffer& harq_ack_,
                                     pusch_decoder_buffer& csi_part1_,
                                     const configuration&  config) override
  {
    sch_data  = &sch_data_;
    harq_ack  = (config.nof_harq_ack_bits != 0) ? &harq_ack_ : nullptr;
    csi_part1 = (config.nof_csi_part1_bits != 0) ? &csi_part1_ : nullptr;
    csi_part2 = nullptr;
    std::memset(&desc, 0, sizeof(desc));
    desc.Qm                          = static_cast<uint8_t>(get_bits_per_symbol(config.modulation));
    desc.nof_layers                  = static_cast<uint8_t>(config.nof_layers);
    desc.start_symbol_index          = static_cast<uint8_t>(config.start_symbol_index);
    desc.nof_symbols                 = static_cast<uint8_t>(config.nof_symbols);
    desc.dmrs_type                   = (config.dmrs == dmrs_type::TYPE1) ? 1 : 2;
    desc.nof_cdm_groups_without_data = static_cast<uint8_t>(config.nof_cdm_groups_without_data);
    desc.nof_prb                     = static_cast<uint16_t>(config.nof_prb);
    for (unsigned l = 0; l != config.dmrs_symbol_mask.size() && l != 14; ++l) {
      desc.dmrs_symbol_mask |= static_cast<uint16_t>(config.dmrs_symbol_mask.test(l) ? (1u << l) : 0u);
    }
    desc.nof_harq_ack_rvd       = config.nof_harq_ack_rvd;
    desc.nof_harq_ack_bits      = config.nof_harq_ack_bits;
    desc.nof_enc_harq_ack_bits  = config.nof_enc_harq_ack_bits;
    desc.nof_csi_part1_bits     = config.nof_csi_part1_bits;
    desc.nof_enc_csi_part1_bits = config.nof_enc_csi_part1_bits;
    uint32_t         lens[5];
    const pdc_status st = pdc_ulsch_demux_lengths(&desc, lens);
    srsran_assert(st == PDC_OK, "Invalid UL-SCH demultiplexer configuration.");
    codeword.resize(lens[0]);
    count      = 0;
    first_bits = 0;
    return *this;
  }

private:
  // See pusch_codeword_buffer interface for documentation.
  span<log_likelihood_ratio> get_next_block_view(unsigned block_size) override
  {
    // Without UCI the reference hands out the decoder's own buffer (ulsch_demultiplex_impl.cpp:253-260): same here.
    if (count == 0 && bypass()) {
      return sch_data->get_next_block_view(block_size);
    }
    block_size = std::min<size_t>(block_size, codeword.size() - count);
    return span<log_likelihood_ratio>(codeword).subspan(count, block_size);
  }

  void on_new_block(span<const log_likelihood_ratio> new_data, const bit_buffer& new_scrambling_seq) override
  {
    srsran_assert(new_data.size() == new_scrambling_seq.size(), "Demodulated and descrambled data sizes must be equal.");
    if (count == 0 && bypass()) {
      sch_data->on_new_softbits(new_data);
      return;
    }
    srsran_assert(count + new_data.size() <= codeword.size(), "More soft bits than the allocation holds.");
    if (new_data.data() != codeword.data() + count) {
      std::memcpy(codeword.data() + count, new_data.data(), new_data.size());
    }
    // The first 31 bits of the scrambling sequence identify its seed.
    for (unsigned k = 0; k != new_data.size() && count + k < 31; ++k) {
      first_bits |= static_cast<uint32_t>(new_scrambling_seq.extract(k, 1)) << (count + k);
    }
    count += new_data.size();
  }

  void on_end_codeword() override
  {
    if (bypass() && count == 0) {
      sch_data->on_end_softbits();
      sch_data = nullptr;
      return;
    }
    srsran_assert(count == codeword.size(), "The codeword is incomplete: {} of {} soft bits.", count, codeword.size());
    desc.c_init = recover_c_init(first_bits, static_cast<unsigned>(std::min<size_t>(count, 31)));
    run();
    deliver(harq_ack, desc.harq_ack_offset, lens[2]);
    // The end of CSI Part 1 may configure CSI Part 2 (feedback): demultiplex again with it.
    const bool csi2_before = (csi_part2 != nullptr);
    deliver(csi_part1, desc.csi_part1_offset, lens[3]);
    if (!csi2_before && csi_part2 != nullptr) {
      run();
    }
    deliver(csi_part2, desc.csi_part2_offset, lens[4]);
    deliver(sch_data, desc.sch_offset, lens[1]);
  }

  bool bypass() const { return harq_ack == nullptr && csi_part1 == nullptr && csi_part2 == nullptr; }

  void run()
  {
    pdc_status st = pdc_ulsch_demux_lengths(&desc, lens);
    srsran_assert(st == PDC_OK, "The UCI fields do not fit the allocation.");
    uint64_t o = 0;
    desc.sch_offset = o, o += (lens[1] + 3u) & ~3u;
    desc.harq_ack_offset = o, o += (lens[2] + 3u) & ~3u;
    desc.csi_part1_offset = o, o += (lens[3] + 3u) & ~3u;
    desc.csi_part2_offset = o, o += (lens[4] + 3u) & ~3u;
    streams.resize(o + 4);
    st = pdc_ulsch_demultiplex(context->get(),
                               &desc,
                               1,
                               reinterpret_cast<const int8_t*>(codeword.data()),
                               codeword.size(),
                               reinterpret_cast<int8_t*>(streams.data()),
                               streams.size());
    srsran_assert(st == PDC_OK, "pdc_ulsch_demultiplex failed ({}): {}", static_cast<int>(st), pdc_last_error(context->get()));
  }

  void deliver(pusch_decoder_buffer*& buffer, uint64_t offset, uint32_t nof_softbits)
  {
    if (buffer == nullptr) {
      return;
    }
    if (nof_softbits != 0) {
      buffer->on_new_softbits(span<const log_likelihood_ratio>(streams).subspan(offset, nof_softbits));
    }
    pusch_decoder_buffer* b = buffer;
    buffer                  = nullptr;
    b->on_end_softbits();
  }

  std::shared_ptr<cuda::pusch_dec_context> context;
  pusch_decoder_buffer*                    sch_data  = nullptr;
  pusch_decoder_buffer*                    harq_ack  = nullptr;
  pusch_decoder_buffer*                    csi_part1 = nullptr;
  pusch_decoder_buffer*                    csi_part2 = nullptr;
  pdc_ulsch_demux_desc                     desc      = {};
  uint32_t                                 lens[5]   = {};
  std::vector<log_likelihood_ratio>        codeword, streams;
  size_t                                   count      = 0;
  uint32_t                                 first_bits = 0;
};

class ulsch_demultiplex_factory_cuda : public ulsch_demultiplex_factory
{
public:
  explicit ulsch_demultiplex_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context_) : context(std::move(context_)) {}
  std::unique_ptr<ulsch_demultiplex> create() override { return std::make_unique<ulsch_demultiplex_cuda>(context); }

private:
  std::shared_ptr<cuda::pusch_dec_context> context;
};

} // namespace

std::shared_ptr<ulsch_demultiplex_factory> srsran::create_ulsch_demultiplex_factory_cuda(std::shared_ptr<cuda::pusch_dec_context> context)
{
  return std::make_shared<ulsch_demultiplex_factory_cuda>(std::move(context));
}

std::shared_ptr<ulsch_demultiplex_factory> srsran::create_ulsch_demultiplex_factory_cuda()
{
  return create_ulsch_demultiplex_factory_cuda(std::make_shared<cuda::pusch_dec_context>(cuda::pusch_dec_context_config{}));
}
