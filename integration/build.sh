#!/usr/bin/env bash
# Builds the reference-side binding (integration/gpqhe_b200_shim.c) and the reference's own test driver twice:
#   _build/libgpqhe_b200.so   the shim: GPQHE evaluation symbols -> libhectr_b200.so (compiled against the reference's
#                             headers where they lie; needs oracle/_ref/libgpqhe.so and hectr_b200/libhectr_b200.so)
#   _build/test-hectr         $REF/tests/hectr.c, unmodified, linked as the reference's tests/Makefile links it
#   _build/test-hectr-b200    the SAME source with -lgpqhe_b200 first in the link order: he_gemv, he_sub, ... resolve to
#                             the shim, everything else to the unmodified libhectr.so / libgpqhe.so
# Outputs are git-ignored (they travel to the GPU box with the snapshot).  Nothing is copied from the reference.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ROOT="$HERE/.."
REF="${REF:-/root/reference}"
OUT="$HERE/_build"
SHIM="$ROOT/oracle/shim"
[ -d "$REF/GPQHE/src" ] || { echo "integration/build: $REF not present; keeping prebuilt $OUT"; exit 0; }
[ -f "$ROOT/oracle/_ref/libgpqhe.so" ] || { echo "integration/build: run oracle/build_ref.sh first"; exit 1; }
[ -f "$ROOT/hectr_b200/libhectr_b200.so" ] || { echo "integration/build: run hectr_b200/csrc/build.sh first"; exit 1; }
mkdir -p "$OUT"
RP="-Wl,-rpath,\$ORIGIN -Wl,-rpath,\$ORIGIN/../../oracle/_ref -Wl,-rpath,\$ORIGIN/../../hectr_b200"
gcc -w -O2 -fPIC -shared -I"$SHIM" -I"$REF/GPQHE/src" -I"$ROOT/include" "$HERE/gpqhe_b200_shim.c" \
    -L"$ROOT/hectr_b200" -lhectr_b200 -L"$ROOT/oracle/_ref" -lgpqhe -l:libgcrypt.so.20 -ldl -lm $RP -o "$OUT/libgpqhe_b200.so"
if [ -f "$ROOT/oracle/_ref/libhectr.so" ]; then
  CF="-w -I$SHIM -Og -U_FORTIFY_SOURCE -D_FORTIFY_SOURCE=0"
  (cd "$REF/tests" && gcc $CF hectr.c -L"$ROOT/oracle/_ref" -lhectr -lgpqhe -lpmu -l:libgcrypt.so.20 -lm $RP -o "$OUT/test-hectr")
  (cd "$REF/tests" && gcc $CF hectr.c -L"$OUT" -Wl,--no-as-needed -lgpqhe_b200 -Wl,--as-needed -L"$ROOT/oracle/_ref" -lhectr -lgpqhe -lpmu -l:libgcrypt.so.20 -lm $RP \
      -o "$OUT/test-hectr-b200")
else
  echo "integration/build: oracle/_ref/libhectr.so missing (no LAPACKE); test-hectr skipped"
fi
echo "integration/build: ok -> $OUT"; ls -la "$OUT"
