// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// C-ABI shim over the UNMODIFIED srsRAN soft demodulator and scrambling-sequence generator, compiled by
// oracle/build_ref.sh into oracle/_ref/libsrsran_ref.so next to ref_shim.cpp. The reference classes used:
//   lib/phy/upper/channel_modulation/demodulation_mapper_impl.h (+ the per-modulation files, built with the AVX512 flags
//   -march=native gives on an AVX512 host), lib/phy/upper/sequence_generators/pseudo_random_generator_impl.h.
#include "lib/phy/upper/channel_modulation/demodulation_mapper_impl.h"
#include "lib/phy/upper/sequence_generators/pseudo_random_generator_impl.h"
#include "srsran/adt/complex.h"
#include "srsran/phy/upper/log_likelihood_ratio.h"
#include <cstdint>

using namespace srsran;

static modulation_scheme demod_mod(int m)
{
  switch (m) {
    case 0:
      return modulation_scheme::PI_2_BPSK;
    case 1:
      return modulation_scheme::BPSK;
    case 2:
      return modulation_scheme::QPSK;
    case 4:
      return modulation_scheme::QAM16;
    case 6:
      return modulation_scheme::QAM64;
    default:
      return modulation_scheme::QAM256;
  }
}

extern "C" {

// demodulation_mapper::demodulate_soft (demodulation_mapper.h:66-69); mod: 0 pi/2-BPSK, 1 BPSK, 2, 4, 6, 8.
void ref_demodulate_soft(int8_t* llrs, const float* symbols_ri, const float* noise_vars, unsigned nof_symbols, int mod)
{
  demodulation_mapper_impl demapper;
  const unsigned           qm = (mod == 0) ? 1 : (unsigned)mod;
  demapper.demodulate_soft(span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(llrs), nof_symbols * qm),
                           span<const cf_t>(reinterpret_cast<const cf_t*>(symbols_ri), nof_symbols),
                           span<const float>(noise_vars, nof_symbols),
                           demod_mod(mod));
}

// pseudo_random_generator::init + advance + apply_xor on soft bits: what pusch_demodulator_impl.cpp:60-222
// (revert_scrambling) does with the sequence generated at :401-405.
void ref_descramble(int8_t* out, const int8_t* in, unsigned n, unsigned c_init, unsigned offset)
{
  pseudo_random_generator_impl prg;
  prg.init(c_init);
  prg.advance(offset);
  prg.apply_xor(span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(out), n),
                span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(in), n));
}

// The sequence itself, one bit per byte.
void ref_prg_bits(uint8_t* out, unsigned n, unsigned c_init, unsigned offset)
{
  pseudo_random_generator_impl prg;
  prg.init(c_init);
  prg.advance(offset);
  std::vector<uint8_t> zero(n, 0);
  prg.apply_xor(span<uint8_t>(out, n), span<const uint8_t>(zero.data(), n));
}
}
