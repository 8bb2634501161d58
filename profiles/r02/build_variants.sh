#!/usr/bin/env bash
# Development variants of the library (logn 12 and 16 only) for A/B timing on the GPU box: build_variants/lib_<tag>.so
set -u
cd "$(dirname "${BASH_SOURCE[0]}")/../.."
mkdir -p build_variants
build() { tag="$1"; shift; OUT="$PWD/build_variants/lib_$tag.so" bash hectr_b200/csrc/build.sh -DHB_DEV_FAST "$@" > "build_variants/$tag.log" 2>&1 && echo "ok $tag" || echo "FAILED $tag"; }
build h_base &
build h_kss4 -DHB32_KSS_MINB=4 &
build h_ksb5 -DHB32_KSB_MINB=5 -DHB32_KSS_MINB=4 &
build h_ksb6 -DHB32_KSB_MINB=6 -DHB32_KSS_MINB=4 &
wait
