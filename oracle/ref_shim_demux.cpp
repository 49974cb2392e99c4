// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// C-ABI shim over the UNMODIFIED srsRAN UL-SCH demultiplexer: ulsch_demultiplex_impl, compiled by
// oracle/build_ref.sh into oracle/_ref/libsrsran_ref.so. Reference files:
// lib/phy/upper/channel_processors/pusch/ulsch_demultiplex_impl.{h,cpp}; the scrambling sequence handed to on_new_block comes
// from the reference's own pseudo_random_generator_impl, as pusch_demodulator_impl.cpp:370-405 produces it.
#include "lib/phy/upper/channel_processors/pusch/ulsch_demultiplex_impl.h"
#include "lib/phy/upper/sequence_generators/pseudo_random_generator_impl.h"
#include "srsran/adt/bit_buffer.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_buffer.h"
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

using namespace srsran;

namespace {

// Collects the soft bits of one stream; set_csi_part2 can be triggered when the stream ends (the product's CSI Part 1
// feedback, pusch_processor_impl.cpp:73-98).
struct collector : public pusch_decoder_buffer {
  std::vector<log_likelihood_ratio> data;
  bool                              ended = false;
  ulsch_demultiplex*                demux = nullptr; // when set: on_end_softbits configures CSI Part 2
  pusch_decoder_buffer*             csi2  = nullptr;
  unsigned                          csi2_bits = 0, csi2_enc_bits = 0;

  span<log_likelihood_ratio> get_next_block_view(unsigned) override { return {}; }
  void on_new_softbits(span<const log_likelihood_ratio> softbits) override { data.insert(data.end(), softbits.begin(), softbits.end()); }
  void on_end_softbits() override
  {
    ended = true;
    if (demux != nullptr) {
      demux->set_csi_part2(*csi2, csi2_bits, csi2_enc_bits);
    }
  }
};

unsigned copy_out(int8_t* dst, unsigned cap, const collector& c)
{
  const unsigned n = static_cast<unsigned>(c.data.size());
  if (dst != nullptr && n <= cap) {
    std::memcpy(dst, c.data.data(), n);
  }
  return n;
}

} // namespace

extern "C" {

// cfg: [0] qm (1, 2, 4, 6, 8) [1] nof_layers [2] nof_prb [3] start_symbol [4] nof_symbols [5] nof_harq_ack_rvd
//      [6] dmrs_type (1 | 2) [7] dmrs_symbol_mask (bit l = symbol l) [8] nof_cdm_groups_without_data [9] nof_harq_ack_bits
//      [10] nof_enc_harq_ack_bits [11] nof_csi_part1_bits [12] nof_enc_csi_part1_bits [13] nof_csi_part2_bits
//      [14] nof_enc_csi_part2_bits [15] csi2_mode: 0 = set_csi_part2 right after demultiplex() (the reference's unit test),
//      1 = from the end-of-stream callback of CSI Part 1 (the product's feedback path) [16] c_init [17] block size of the
//      on_new_block calls (0 = everything in one call).
// in: n_in descrambled soft bits. out_*: capacity cap each; lens[4] = soft bits delivered to SCH, HARQ-ACK, CSI1, CSI2.
// Returns 0.
int ref_ulsch_demultiplex(const uint32_t* cfg,
                          const int8_t*   in,
                          unsigned        n_in,
                          int8_t*         out_sch,
                          int8_t*         out_harq,
                          int8_t*         out_csi1,
                          int8_t*         out_csi2,
                          unsigned        cap,
                          unsigned*       lens)
{
  ulsch_demultiplex::configuration c;
  c.modulation         = (cfg[0] == 1)   ? modulation_scheme::PI_2_BPSK
                         : (cfg[0] == 2) ? modulation_scheme::QPSK
                         : (cfg[0] == 4) ? modulation_scheme::QAM16
                         : (cfg[0] == 6) ? modulation_scheme::QAM64
                                         : modulation_scheme::QAM256;
  c.nof_layers         = cfg[1];
  c.nof_prb            = cfg[2];
  c.start_symbol_index = cfg[3];
  c.nof_symbols        = cfg[4];
  c.nof_harq_ack_rvd   = cfg[5];
  c.dmrs               = (cfg[6] == 1) ? dmrs_type::TYPE1 : dmrs_type::TYPE2;
  c.dmrs_symbol_mask   = symbol_slot_mask(14);
  for (unsigned l = 0; l != 14; ++l) {
    c.dmrs_symbol_mask.set(l, ((cfg[7] >> l) & 1u) != 0);
  }
  c.nof_cdm_groups_without_data = cfg[8];
  c.nof_harq_ack_bits           = cfg[9];
  c.nof_enc_harq_ack_bits       = cfg[10];
  c.nof_csi_part1_bits          = cfg[11];
  c.nof_enc_csi_part1_bits      = cfg[12];

  // Value-initialised as the reference's factory creates it (std::make_unique<ulsch_demultiplex_impl>()): the counters start at 0.
  std::unique_ptr<ulsch_demultiplex_impl> demux_ptr = std::make_unique<ulsch_demultiplex_impl>();
  ulsch_demultiplex_impl&                 demux     = *demux_ptr;
  collector                               sch, harq, csi1, csi2;
  pusch_codeword_buffer&                  cw = demux.demultiplex(sch, harq, csi1, c);
  if (cfg[14] != 0) {
    if (cfg[15] == 0) {
      demux.set_csi_part2(csi2, cfg[13], cfg[14]);
    } else {
      csi1.demux = &demux, csi1.csi2 = &csi2, csi1.csi2_bits = cfg[13], csi1.csi2_enc_bits = cfg[14];
    }
  }

  // Scrambling sequence of the whole codeword, as the demodulator hands it over block by block.
  pseudo_random_generator_impl prg;
  prg.init(cfg[16]);
  dynamic_bit_buffer seq(n_in);
  prg.generate(seq);

  const unsigned                    block = (cfg[17] == 0) ? n_in : cfg[17];
  std::vector<log_likelihood_ratio> data(in, in + n_in);
  for (unsigned o = 0; o < n_in; o += block) {
    const unsigned     n = (n_in - o < block) ? n_in - o : block;
    dynamic_bit_buffer part(n);
    for (unsigned k = 0; k != n; ++k) {
      part.insert(seq.extract(o + k, 1), k, 1);
    }
    cw.on_new_block(span<const log_likelihood_ratio>(data).subspan(o, n), part);
  }
  cw.on_end_codeword();

  lens[0] = copy_out(out_sch, cap, sch);
  lens[1] = copy_out(out_harq, cap, harq);
  lens[2] = copy_out(out_csi1, cap, csi1);
  lens[3] = copy_out(out_csi2, cap, csi2);
  return 0;
}

} // extern "C"
