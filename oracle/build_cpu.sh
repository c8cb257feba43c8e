#!/usr/bin/env bash
# TEST INFRASTRUCTURE. Builds the C restatement oracle/oracle_cpu.c into oracle/_build/liboracle_cpu.so.
# -ffp-contract=off: only the explicit fmaf() calls fuse (see the header of oracle_cpu.c).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
mkdir -p "$HERE/_build"
SRC="$HERE/oracle_cpu.c"
OUT="$HERE/_build/liboracle_cpu.so"
if [ ! -f "$OUT" ] || [ "$SRC" -nt "$OUT" ]; then
    gcc -O2 -std=c99 -fPIC -shared -fopenmp -ffp-contract=off -fno-fast-math -Wall -Wno-unknown-pragmas \
        -Wno-misleading-indentation -o "$OUT" "$SRC" -lm
    echo "[build_cpu] built $OUT"
fi
