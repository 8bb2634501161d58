#!/usr/bin/env bash
# Development variants of the library (logn 12 and 16 only) for A/B timing on the GPU box: build_variants/lib_<tag>.so
set -u
cd "$(dirname "${BASH_SOURCE[0]}")/../.."
mkdir -p build_variants
build() { tag="$1"; shift; OUT="$PWD/build_variants/lib_$tag.so" bash hectr_b200/csrc/build.sh -DHB_DEV_FAST "$@" > "build_variants/$tag.log" 2>&1 && echo "ok $tag" || echo "FAILED $tag"; }
build base &
build minbbig2 -DHB_MINB_BIG=2 &
build ksminb2 -DHB_KS_MINB=2 &
build nostream -DHB_NOSTREAM &
wait
