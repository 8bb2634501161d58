#!/usr/bin/env bash
# A/B of build_variants/lib_*.so against the shipped library: per-kernel times of one 4096-instance evaluation, result checked
# against the shipped library's by prof_driver (--check)
mkdir -p gpurun_out
echo "== shipped"; python profiles/prof_driver.py --case hempc --batch 4096 --reps 3 2>&1 | head -12
for v in build_variants/lib_*.so; do
  [ -f "$v" ] || continue
  echo "== $v"; HB_LIB="$PWD/$v" python profiles/prof_driver.py --case hempc --batch 4096 --reps 3 2>&1 | head -12
done | tee gpurun_out/variants_run.log
python -m pytest tests/test_gpu_shapes.py tests/test_gpu_parity.py -m gpu -q -x -k "config_E or shortcuts or paired or gemv_and_hempc or synthetic" 2>&1 | tail -3
