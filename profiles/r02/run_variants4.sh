#!/usr/bin/env bash
mkdir -p gpurun_out
( echo "== shipped"; python profiles/prof_driver.py --case hempc --batch 4096 --reps 3 2>&1 | head -13
for v in build_variants/lib_*.so; do
  [ -f "$v" ] || continue
  echo "== $v"; HB_LIB="$PWD/$v" python profiles/prof_driver.py --case hempc --batch 4096 --reps 3 2>&1 | head -13
done ) | tee gpurun_out/variants_reread2.log
timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
