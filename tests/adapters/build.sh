#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY. Builds tests/adapters/_build/libadapter_harness.so: the UNMODIFIED reference transport-block
# decoders (pusch_decoder_impl, pusch_decoder_hw_impl, rx_buffer_pool) and the reference's
# pdsch_encoder_hw_impl, compiled from the sources where they lie under
# /root/reference, linked with the CUDA adapters, behind the small C ABI of harness.cpp. Reuses the reference objects that
# oracle/build_ref.sh already compiled (oracle/_ref/obj). Nothing is copied into the repo; outputs are git-ignored.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ROOT="$(cd "$HERE/../.." && pwd)"
OUT="$HERE/_build"
OBJ="$OUT/obj"
if [ ! -d "$REF/lib/phy/upper/channel_processors/pusch" ]; then
  echo "reference tree not found at $REF: keeping prebuilt $OUT (if any)" >&2
  exit 0
fi
[ -d "$ROOT/oracle/_ref/obj" ] || bash "$ROOT/oracle/build_ref.sh"
mkdir -p "$OBJ"
# The environment's CXX may be a wrapper that links libstdc++ statically; several private copies of libstdc++ in one
# process (this library, the adapters, numpy) corrupt each other's locale facets. Use the system compiler, shared runtime.
CXX=${PDC_CXX:-/usr/bin/g++}
INC="-I$REF/include -I$REF/external/fmt/include -I$REF/external -I$REF/lib -I$REF/lib/phy/upper/channel_coding -I$REF/lib/phy/upper/channel_processors/pusch"
BASE="-std=c++17 -O2 -fPIC -march=x86-64-v3 -mtune=generic -DFMT_HEADER_ONLY -DASSERTS_ENABLED -w"
SRCS=(
  lib/phy/upper/channel_processors/pusch/pusch_decoder_impl.cpp
  lib/phy/upper/channel_processors/pusch/pusch_decoder_hw_impl.cpp
  lib/phy/upper/channel_processors/pdsch/pdsch_encoder_hw_impl.cpp
  lib/phy/upper/rx_buffer_pool_impl.cpp
  lib/support/cpu_architecture_info.cpp
  lib/srsvec/copy.cpp
  lib/srsvec/zero.cpp
)
while IFS= read -r f; do SRCS+=("${f#$REF/}"); done < <(find "$REF/lib/srslog" -name '*.cpp' | sort)
pids=()
objs=()
for s in "${SRCS[@]}"; do
  [ -f "$REF/$s" ] || continue
  o="$OBJ/$(echo "$s" | tr '/' '_' | sed 's/\.cpp$/.o/')"
  objs+=("$o")
  if [ ! -f "$o" ] || [ "$REF/$s" -nt "$o" ]; then
    $CXX $BASE $INC -c "$REF/$s" -o "$o" &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
ref_objs=$(ls "$ROOT"/oracle/_ref/obj/*.o | grep -v ref_shim)
$CXX $BASE $INC -c "$HERE/harness.cpp" -o "$OBJ/harness.o"
$CXX -shared -o "$OUT/libadapter_harness.so" "$OBJ/harness.o" "${objs[@]}" $ref_objs \
  -L"$ROOT/srsran_project_b200" -lsrsran_cuda_adapters -lpuschdec_cuda -lpthread \
  -Wl,-rpath,'$ORIGIN/../../../srsran_project_b200'
echo "built $OUT/libadapter_harness.so"
