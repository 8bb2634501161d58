#!/usr/bin/env bash
# variants of the tails (resident CTAs per SM) against the shipped library, then the whole GPU suite on the shipped library
mkdir -p gpurun_out
( echo "== shipped"; python profiles/prof_driver.py --case hempc --batch 4096 --reps 3 2>&1 | head -12
for v in build_variants/lib_*.so; do
  [ -f "$v" ] || continue
  echo "== $v"; HB_LIB="$PWD/$v" python profiles/prof_driver.py --case hempc --batch 4096 --reps 3 2>&1 | head -12
done ) | tee gpurun_out/variants_tails.log | grep -E "^==|ctr_hempc|k_ks_finish"
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/pytest11.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest11.log
