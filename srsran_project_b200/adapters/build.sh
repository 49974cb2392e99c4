#!/usr/bin/env bash
# Builds libsrsran_cuda_adapters.so: the C++ adapters that put libpuschdec_cuda.so behind srsRAN's operator interfaces
# (ldpc_decoder, ldpc_rate_dematcher, hal::hw_accelerator_pusch_dec, pusch_decoder, demodulation_mapper, pdsch_encoder, hal::hw_accelerator_pdsch_enc, ulsch_demultiplex). Only the reference's PUBLIC headers are needed
# (-I$REF/include and its vendored fmt); without a reference tree the prebuilt library (if any) is kept, since the
# adapters cannot be compiled without the interfaces they implement.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
PKG="$(dirname "$HERE")"
OUT="$PKG/libsrsran_cuda_adapters.so"
if [ ! -d "$REF/include/srsran" ]; then
  echo "reference headers not found at $REF: keeping prebuilt $OUT (if any)" >&2
  exit 0
fi
# The environment's CXX may be a wrapper that links libstdc++ statically; several private copies of libstdc++ in one
# process (this library, the adapters, numpy) corrupt each other's locale facets. Use the system compiler, shared runtime.
CXX=${PDC_CXX:-/usr/bin/g++}
FLAGS="-std=c++17 -O2 -fPIC -DFMT_HEADER_ONLY -DASSERTS_ENABLED -w -I$REF/include -I$REF/external/fmt/include -I$REF/external"
$CXX $FLAGS -shared -o "$OUT" "$HERE/channel_coding_factories_cuda.cpp" "$HERE/hw_accelerator_pusch_dec_cuda.cpp" "$HERE/pusch_dec_batcher.cpp" \
  "$HERE/pusch_decoder_cuda.cpp" "$HERE/channel_modulation_factories_cuda.cpp" "$HERE/pdsch_encoder_cuda.cpp" \
  "$HERE/hw_accelerator_pdsch_enc_cuda.cpp" "$HERE/ulsch_demultiplex_cuda.cpp" \
  -L"$PKG" -lpuschdec_cuda -lpthread -Wl,-rpath,'$ORIGIN'
echo "built $OUT"
# The boundary test harness (reference TB decoders over the adapters) is test infrastructure; built here because it
# needs the same reference tree.
bash "$PKG/../tests/adapters/build.sh"
