#!/usr/bin/env bash
# Compiles tools/sass_premise.cu for sm_100a and prints, per kernel, the arithmetic SASS between the loads and the store.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
TMP=$(mktemp -d)
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -cubin -o "$TMP/p.cubin" "$HERE/sass_premise.cu"
cuobjdump -sass "$TMP/p.cubin" | python3 -c '
import re, sys
txt = sys.stdin.read()
skip = ("LDC", "LDG", "STG", "S2R", "S2UR", "IMAD.WIDE", "EXIT", "BRA", "NOP", "LDCU", "ULDC", "MOV", "IMAD.MOV")
print("kernel: arithmetic instructions between loads and store (address arithmetic, loads, stores, moves excluded)")
for part in re.split(r"\n\s*Function : ", txt)[1:]:
    name = part.split("\n")[0].strip()
    ops = []
    for l in part.split("\n"):
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(@!?U?P\d\s+)?([A-Za-z0-9_.]+)", l)
        if m and not m.group(2).startswith(skip):
            ops.append(m.group(2))
    print("%-14s %2d  %s" % (name, len(ops), " ".join(ops)))
'
rm -rf "$TMP"
