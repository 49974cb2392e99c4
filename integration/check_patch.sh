#!/usr/bin/env bash
# Product wiring: applies integration/srsran_cuda.patch to a scratch copy of the srsRAN files it touches and
# syntax-checks the patched translation units against the reference's headers and this repository's adapters, with and without
# SRSRAN_CUDA_FOUND. Nothing is written to the reference tree or to this repository. Usage: integration/check_patch.sh [REF]
set -euo pipefail
REF=${1:-${REF:-/root/reference}}
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ADP="$HERE/../srsran_project_b200/adapters"
[ -d "$REF/apps/units/flexible_o_du/o_du_low" ] || { echo "reference tree not found at $REF" >&2; exit 2; }
TMP="$(mktemp -d)"
trap 'rm -rf "$TMP"' EXIT
mkdir -p "$TMP/apps/units/flexible_o_du" "$TMP/lib/phy/upper" "$TMP/lib/phy/metrics"
cp -r "$REF/apps/units/flexible_o_du/o_du_low" "$TMP/apps/units/flexible_o_du/"
cp "$REF/lib/phy/upper/upper_phy_factories.cpp" "$REF/lib/phy/upper/CMakeLists.txt" "$TMP/lib/phy/upper/"
cp "$REF/CMakeLists.txt" "$TMP/"
chmod -R u+w "$TMP"
(cd "$TMP" && patch -p1 --no-backup-if-mismatch < "$HERE/srsran_cuda.patch")
CXX=${PDC_CXX:-/usr/bin/g++}
INC="-I$REF/include -I$REF/external/fmt/include -I$REF/external -I$REF/external/CLI -I$REF -I$REF/lib -I$ADP"
for defs in "-DSRSRAN_CUDA_FOUND" ""; do
  for f in apps/units/flexible_o_du/o_du_low/du_low_hal_factory.cpp apps/units/flexible_o_du/o_du_low/du_low_config_cli11_schema.cpp \
           apps/units/flexible_o_du/o_du_low/du_low_config_validator.cpp lib/phy/upper/upper_phy_factories.cpp; do
    $CXX -std=c++17 -fsyntax-only -DFMT_HEADER_ONLY $defs -w -I"$TMP/$(dirname "$f")" -I"$TMP" $INC -I"$REF/$(dirname "$f")" "$TMP/$f"
    echo "ok: $f ${defs:-(without SRSRAN_CUDA_FOUND)}"
  done
done
echo "patch applies and the patched files compile (du_low_config_yaml_writer.cpp is patched too but needs yaml-cpp, which this machine lacks: not compiled)"
