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
# knobs of the later A/B runs (variants_tails.log, variants_tw_tr.log, variants_occ.log, variants_mont.log):
#   -DHB32_KSS_UNROLL=2|8|16  -DHB32_KSS_PUNROLL=4  -DHB32_KSB_PUNROLL=4     loop unrolling of the 30-bit key-switch kernels
#   -DHB_FSH_MINB=1|5|7|8  -DHB_FH_MINB=1|10|12                                resident CTAs per SM of the hybrid tails
#   -DHB32_TW_TR=0                                                             natural twiddle table in the stride-1 pass
#   -DHB32_KSB_MINB=4|5  -DHB32_KSS_MINB=3  -DHB32_PN_MINB=3  -DHB32_MA_MINB=2  resident CTAs per SM of the 30-bit kernels
#   -DHB32_MA_MONT=0                                                           Shoup pairs instead of Montgomery-form diagonals
