// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// C-ABI shim over the UNMODIFIED srsRAN PDSCH encoder: pdsch_encoder_impl driving ldpc_segmenter_tx_impl,
// ldpc_encoder_avx2 (or generic) and ldpc_rate_matcher_impl, compiled by oracle/build_ref.sh into
// oracle/_ref/libsrsran_ref.so. Reference files: lib/phy/upper/channel_processors/pdsch/pdsch_encoder_impl.{h,cpp},
// lib/phy/upper/channel_coding/ldpc/ldpc_segmenter_tx_impl.{h,cpp}.
#include "crc_calculator_lut_impl.h"
#include "ldpc/ldpc_encoder_avx2.h"
#include "ldpc/ldpc_encoder_generic.h"
#include "ldpc/ldpc_rate_matcher_impl.h"
#include "ldpc/ldpc_segmenter_tx_impl.h"
#include "lib/phy/upper/channel_processors/pdsch/pdsch_encoder_impl.h"
#include <cstdint>
#include <memory>

using namespace srsran;

extern "C" {

// pdsch_encoder::encode. codeword: nof_ch_symbols * qm bytes, one bit each (nof_ch_symbols counts every layer). generic != 0: ldpc_encoder_generic.
// Returns 0.
int ref_pdsch_encode(uint8_t*       codeword,
                     const uint8_t* tb_bytes,
                     unsigned       tbs_bits,
                     int            bg,
                     unsigned       rv,
                     unsigned       qm,
                     unsigned       Nref,
                     unsigned       nof_layers,
                     unsigned       nof_ch_symbols,
                     int            generic)
{
  ldpc_segmenter_tx_impl::sch_crc crcs;
  crcs.crc16  = std::make_unique<crc_calculator_lut_impl>(crc_generator_poly::CRC16);
  crcs.crc24A = std::make_unique<crc_calculator_lut_impl>(crc_generator_poly::CRC24A);
  crcs.crc24B = std::make_unique<crc_calculator_lut_impl>(crc_generator_poly::CRC24B);
  std::unique_ptr<ldpc_encoder> enc;
  if (generic) {
    enc = std::make_unique<ldpc_encoder_generic>();
  } else {
    enc = std::make_unique<ldpc_encoder_avx2>();
  }
  pdsch_encoder_impl encoder(
      std::make_unique<ldpc_segmenter_tx_impl>(crcs), std::move(enc), std::make_unique<ldpc_rate_matcher_impl>());
  pdsch_encoder::configuration cfg;
  cfg.base_graph     = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  cfg.rv             = rv;
  cfg.Nref           = Nref;
  cfg.nof_layers     = nof_layers;
  cfg.nof_ch_symbols = nof_ch_symbols;
  switch (qm) {
    case 1:
      cfg.mod = modulation_scheme::BPSK;
      break;
    case 2:
      cfg.mod = modulation_scheme::QPSK;
      break;
    case 4:
      cfg.mod = modulation_scheme::QAM16;
      break;
    case 6:
      cfg.mod = modulation_scheme::QAM64;
      break;
    default:
      cfg.mod = modulation_scheme::QAM256;
      break;
  }
  encoder.encode(span<uint8_t>(codeword, nof_ch_symbols * qm), span<const uint8_t>(tb_bytes, tbs_bits / 8), cfg);
  return 0;
}
}
