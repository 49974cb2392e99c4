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
# Product wiring: the reference's DU-low HAL factory with integration/srsran_cuda.patch applied, in a scratch copy
# under _build/ (the reference tree itself is never touched), so that the test can call the patched make_du_low_hal_dependencies.
N3="$OUT/n3"
rm -rf "$N3"
mkdir -p "$N3/apps/units/flexible_o_du" "$N3/lib/phy/upper" "$N3/lib/phy/metrics"
cp -r "$REF/apps/units/flexible_o_du/o_du_low" "$N3/apps/units/flexible_o_du/"
cp "$REF/lib/phy/upper/upper_phy_factories.cpp" "$REF/lib/phy/upper/CMakeLists.txt" "$N3/lib/phy/upper/"
cp "$REF/CMakeLists.txt" "$N3/"
chmod -R u+w "$N3"
(cd "$N3" && patch -s -p1 --no-backup-if-mismatch < "$ROOT/integration/srsran_cuda.patch")
N3INC="-I$N3/apps/units/flexible_o_du/o_du_low -I$N3/lib/phy/metrics -I$REF -I$ROOT/srsran_project_b200/adapters"
$CXX $BASE $INC $N3INC -DSRSRAN_CUDA_FOUND -c "$N3/apps/units/flexible_o_du/o_du_low/du_low_hal_factory.cpp" -o "$OBJ/n3_du_low_hal_factory.o"
$CXX $BASE $INC $N3INC -c "$HERE/harness.cpp" -o "$OBJ/harness.o"
$CXX -shared -o "$OUT/libadapter_harness.so" "$OBJ/harness.o" "$OBJ/n3_du_low_hal_factory.o" "${objs[@]}" $ref_objs \
  -L"$ROOT/srsran_project_b200" -lsrsran_cuda_adapters -lpuschdec_cuda -lpthread \
  -Wl,-rpath,'$ORIGIN/../../../srsran_project_b200'
rm -rf "$N3" # the patched scratch copy of the reference files has served: only objects stay under _build/
echo "built $OUT/libadapter_harness.so"
