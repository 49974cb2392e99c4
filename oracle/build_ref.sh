#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY. Builds oracle/_ref/libsrsran_ref.so from the UNMODIFIED reference sources where they lie
# under /root/reference (nothing is copied into this repo). The reference's own CMake tree cannot configure in this
# image (MbedTLS / yaml-cpp / gtest / SCTP absent), so the translation units of the PUSCH code-block
# decode path are compiled one by one with the reference's Release flags (CMakeLists.txt:206,564) and the per-file SIMD
# flags of lib/phy/upper/channel_coding/{,ldpc/}CMakeLists.txt. Baseline ISA is x86-64-v3 (instead of -march=native)
# so that the library also loads on the GPU box's host, whose CPU model may differ from this container's.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/_ref"
OBJ="$OUT/obj"
if [ ! -d "$REF/lib/phy/upper/channel_coding/ldpc" ]; then
  echo "reference tree not found at $REF: keeping prebuilt $OUT (if any)" >&2
  exit 0
fi
mkdir -p "$OBJ"
# The environment's CXX may be a wrapper that links libstdc++ statically; several private copies of libstdc++ in one
# process (this library, the adapters, numpy) corrupt each other's locale facets. Use the system compiler, shared runtime.
CXX=${PDC_CXX:-/usr/bin/g++}
INC="-I$REF/include -I$REF/external/fmt/include -I$REF/external -I$REF/lib -I$REF/lib/phy/upper/channel_coding -I$REF/lib/phy/upper/channel_processors/pusch"
BASE="-std=c++17 -O3 -fPIC -march=x86-64-v3 -mtune=generic -fno-trapping-math -fno-math-errno -DFMT_HEADER_ONLY -DASSERTS_ENABLED -w"

CC_DIR=lib/phy/upper/channel_coding
CM_DIR=lib/phy/upper/channel_modulation
AVX512_NATIVE="-mavx512f -mavx512bw -mavx512vl -mavx512dq -mavx512cd"
declare -A EXTRA=(
  ["$CC_DIR/ldpc/ldpc_decoder_avx2.cpp"]="-mavx2"
  ["$CC_DIR/ldpc/ldpc_decoder_avx512.cpp"]="-mavx512f -mavx512bw"
  ["$CC_DIR/ldpc/ldpc_encoder_avx2.cpp"]="-mavx2"
  ["$CC_DIR/ldpc/ldpc_rate_dematcher_avx2_impl.cpp"]="-mavx2"
  ["$CC_DIR/ldpc/ldpc_rate_dematcher_avx512_impl.cpp"]="-mavx512f -mavx512bw -mavx512vbmi"
  ["$CC_DIR/crc_calculator_clmul_impl.cpp"]="-mpclmul -msse4.1"
  # Soft demodulator: these files take the global -march in the reference's build (native); on the AVX512
  # hosts the metric names that is the HAVE_AVX512 flavour of every modulation.
  ["$CM_DIR/demodulation_mapper_impl.cpp"]="$AVX512_NATIVE"
  ["$CM_DIR/demodulation_mapper_qpsk.cpp"]="$AVX512_NATIVE"
  ["$CM_DIR/demodulation_mapper_qam16.cpp"]="$AVX512_NATIVE"
  ["$CM_DIR/demodulation_mapper_qam64.cpp"]="$AVX512_NATIVE"
  ["$CM_DIR/demodulation_mapper_qam256.cpp"]="$AVX512_NATIVE"
)
SRCS=(
  $CC_DIR/ldpc/ldpc_graph_impl.cpp
  $CC_DIR/ldpc/ldpc_luts_impl.cpp
  $CC_DIR/ldpc/ldpc_decoder_impl.cpp
  $CC_DIR/ldpc/ldpc_decoder_generic.cpp
  $CC_DIR/ldpc/ldpc_decoder_avx2.cpp
  $CC_DIR/ldpc/ldpc_decoder_avx512.cpp
  $CC_DIR/ldpc/ldpc_rate_dematcher_impl.cpp
  $CC_DIR/ldpc/ldpc_rate_dematcher_avx2_impl.cpp
  $CC_DIR/ldpc/ldpc_rate_dematcher_avx512_impl.cpp
  $CC_DIR/ldpc/ldpc_encoder_impl.cpp
  $CC_DIR/ldpc/ldpc_encoder_generic.cpp
  $CC_DIR/ldpc/ldpc_encoder_avx2.cpp
  $CC_DIR/ldpc/ldpc_rate_matcher_impl.cpp
  $CC_DIR/ldpc/ldpc_segmenter_rx_impl.cpp
  $CC_DIR/crc_calculator_generic_impl.cpp
  $CC_DIR/crc_calculator_lut_impl.cpp
  $CC_DIR/crc_calculator_clmul_impl.cpp
  lib/phy/upper/log_likelihood_ratio.cpp
  lib/srsvec/compare.cpp
  lib/srsvec/bit.cpp
  lib/phy/upper/channel_processors/pusch/pusch_codeblock_decoder.cpp
  $CM_DIR/demodulation_mapper_impl.cpp
  $CM_DIR/demodulation_mapper_qpsk.cpp
  $CM_DIR/demodulation_mapper_qam16.cpp
  $CM_DIR/demodulation_mapper_qam64.cpp
  $CM_DIR/demodulation_mapper_qam256.cpp
  lib/phy/upper/sequence_generators/pseudo_random_generator_impl.cpp
  $CC_DIR/ldpc/ldpc_segmenter_tx_impl.cpp
  lib/phy/upper/channel_processors/pdsch/pdsch_encoder_impl.cpp
  lib/phy/upper/channel_processors/pusch/ulsch_demultiplex_impl.cpp
)
# Extra reference units may be appended by the caller (e.g. the whole pusch_decoder for the TB-level harness).
SRCS+=(${REF_EXTRA_SRCS:-})

pids=()
objs=()
for s in "${SRCS[@]}"; do
  o="$OBJ/$(echo "$s" | tr '/' '_' | sed 's/\.cpp$/.o/')"
  objs+=("$o")
  if [ ! -f "$o" ] || [ "$REF/$s" -nt "$o" ]; then
    $CXX $BASE ${EXTRA[$s]:-} $INC -c "$REF/$s" -o "$o" &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done

$CXX $BASE $INC -c "$HERE/ref_shim.cpp" -o "$OBJ/ref_shim.o"
$CXX $BASE $INC -I$REF -c "$HERE/ref_shim_demod.cpp" -o "$OBJ/ref_shim_demod.o"
$CXX $BASE $INC -I$REF -c "$HERE/ref_shim_pdsch.cpp" -o "$OBJ/ref_shim_pdsch.o"
$CXX $BASE $INC -I$REF -c "$HERE/ref_shim_demux.cpp" -o "$OBJ/ref_shim_demux.o"
$CXX -shared -o "$OUT/libsrsran_ref.so" "$OBJ/ref_shim.o" "$OBJ/ref_shim_demod.o" "$OBJ/ref_shim_pdsch.o" "$OBJ/ref_shim_demux.o" "${objs[@]}" -lpthread
echo "built $OUT/libsrsran_ref.so"
