#!/usr/bin/env bash
# Development variants of the library (logn 12 and 16 only) for A/B timing on the GPU box: build_variants/lib_<tag>.so, run with
# profiles/r02/run_variants.sh.  Knobs that exist in the shipped sources are listed; the CTA-order and fused-pass variants of
# variants_order*.log / fused_passes.txt were source patches (git history: 77f43ad, and the HB32_TILE_ORDER / HB32_MA_DMAJOR
# experiment described in profiles/r02/README.md).
set -u
cd "$(dirname "${BASH_SOURCE[0]}")/../.."
mkdir -p build_variants
build() { tag="$1"; shift; OUT="$PWD/build_variants/lib_$tag.so" bash hectr_b200/csrc/build.sh -DHB_DEV_FAST "$@" > "build_variants/$tag.log" 2>&1 && echo "ok $tag" || echo "FAILED $tag"; }
build base &
build minbbig2 -DHB_MINB_BIG=2 &          # N >= 2^14 transforms: 2 CTAs of up to 128 registers per SM
build ksminb2 -DHB_KS_MINB=2 &            # 59-bit key-switch kernels: 2 CTAs per SM
build nostream -DHB_NOSTREAM &            # no evict-first cache hints
build ksb4_kss3 -DHB32_KSB_MINB=4 -DHB32_KSS_MINB=3 &
wait
