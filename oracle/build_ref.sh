#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Compiles the UNMODIFIED reference (GPQHE + HECTR + libpmu)
# from the sources where they lie under $REF (default /root/reference) into
# oracle/_ref/ (git-ignored).  Nothing is copied into the repo.  Recipe = SURVEY.md
# Appendix B: the reference's own flags (-Og -fomit-frame-pointer -DSUPERCOP) plus an
# -O2 twin for the CPU baseline, declaration-only shim headers for the two absent
# third-party headers (libgcrypt, LAPACKE).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${REF:-/root/reference}"
OUT="$HERE/_ref"
SHIM="$HERE/shim"
[ -d "$REF/GPQHE/src" ] || { echo "build_ref: $REF not present; keeping prebuilt $OUT"; exit 0; }
mkdir -p "$OUT"
OB_DIR="$(python - <<'PY'
import os, glob
try:
    import scipy
    d = os.path.join(os.path.dirname(scipy.__file__), os.pardir, "scipy.libs")
    c = glob.glob(os.path.join(d, "libscipy_openblas-*.so"))
    print(os.path.realpath(os.path.dirname(c[0])) if c else "")
except Exception:
    print("")
PY
)"
OB_LIB="$(basename "$(ls "$OB_DIR"/libscipy_openblas-*.so 2>/dev/null | head -1)" 2>/dev/null || true)"
G="types.c precomp.c reduce.c ntt.c rns.c poly.c fips202.c canemb.c rng.c symmetric-shake.c sample.c he-mem.c \
   he-info.c he-encode.c he-encrypt.c he-rescale.c he-kem.c he-add.c he-mult.c he-automorphism.c he-algo.c he-bootstrap.c"
for variant in "Og:-Og" "O2:-O2"; do
  tag="${variant%%:*}"; opt="${variant##*:}"
  suffix=""; [ "$tag" = "O2" ] && suffix="_O2"
  (cd "$REF/GPQHE/src" && gcc -w -I"$SHIM" $opt -fomit-frame-pointer -DSUPERCOP $G -fPIC -shared \
      -l:libgcrypt.so.20 -lm -o "$OUT/libgpqhe${suffix}.so")
done
gcc -w -Og -fPIC -shared "$REF/libpmu/pmu.c" -o "$OUT/libpmu.so"
if [ -n "$OB_LIB" ]; then
  H="ode.c cstr.c matrices.c ctr.c dare.c dlqr.c dlqe.c mpc.c hempc.c quadprog.c"
  (cd "$REF/src" && gcc -w -I"$SHIM" -Og -fomit-frame-pointer $H -fPIC -shared -L"$OUT" -lgpqhe -lpmu \
      -L"$OB_DIR" -l:"$OB_LIB" -Wl,-rpath,'$ORIGIN' -Wl,-rpath,"$OB_DIR" -lm -o "$OUT/libhectr.so")
  (cd "$REF/src" && gcc -w -I"$SHIM" -O2 -fomit-frame-pointer $H -fPIC -shared -L"$OUT" -l:libgpqhe_O2.so -lpmu \
      -L"$OB_DIR" -l:"$OB_LIB" -Wl,-rpath,'$ORIGIN' -Wl,-rpath,"$OB_DIR" -lm -o "$OUT/libhectr_O2.so")
else
  echo "build_ref: scipy OpenBLAS not found; libhectr.so skipped"
fi
# helper: MPI <-> little-endian two's-complement words (our own code, links the reference .so)
gcc -w -I"$SHIM" -O2 -fPIC -shared "$HERE/ref_helpers.c" -l:libgcrypt.so.20 -o "$OUT/libref_helpers.so"
echo "build_ref: ok -> $OUT"; ls -la "$OUT"
