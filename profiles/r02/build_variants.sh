#!/usr/bin/env bash
# Development variants of the library (logn 12 and 16 only) for A/B timing on the GPU box: build_variants/lib_<tag>.so
set -u
cd "$(dirname "${BASH_SOURCE[0]}")/../.."
mkdir -p build_variants
build() { tag="$1"; shift; OUT="$PWD/build_variants/lib_$tag.so" bash hectr_b200/csrc/build.sh -DHB_DEV_FAST "$@" > "build_variants/$tag.log" 2>&1 && echo "ok $tag" || echo "FAILED $tag"; }
build w_base &
build w_tile1024 -DHB32_TILE_ORDER=1024 &
build w_tile512_madmajor -DHB32_TILE_ORDER=512 -DHB32_MA_DMAJOR &
build w_madmajor -DHB32_MA_DMAJOR &
wait
