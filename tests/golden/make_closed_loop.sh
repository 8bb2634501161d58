#!/usr/bin/env bash
# Regenerates tests/golden/closed_loop/ (run where /root/reference exists, after integration/build.sh):
#   fixture_*.bin  the reference's own committed results, tests/results/cstr-{hempc,mpc}.bin (41 x {uint k; double x[3], u[2]})
#   refrun_*.bin   the UNMODIFIED reference run here: integration/_build/test-hectr cstr-mpc / cstr-hempc (2.5 min of CPU)
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${REF:-/root/reference}"
OUT="$HERE/closed_loop"
mkdir -p "$OUT"
cp "$REF/tests/results/cstr-hempc.bin" "$OUT/fixture_cstr-hempc.bin"
cp "$REF/tests/results/cstr-mpc.bin" "$OUT/fixture_cstr-mpc.bin"
T="$(mktemp -d)"; mkdir -p "$T/results"
(cd "$T" && "$HERE/../../integration/_build/test-hectr" cstr-mpc >/dev/null && "$HERE/../../integration/_build/test-hectr" cstr-hempc >/dev/null 2>&1)
cp "$T/results/cstr-hempc.bin" "$OUT/refrun_cstr-hempc.bin"
cp "$T/results/cstr-mpc.bin" "$OUT/refrun_cstr-mpc.bin"
rm -rf "$T"
