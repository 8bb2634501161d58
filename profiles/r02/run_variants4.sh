#!/usr/bin/env bash
mkdir -p gpurun_out
( echo "== shipped"; python profiles/prof_driver.py --case hempc --batch 4096 --reps 3 2>&1 | head -13
for v in build_variants/lib_*.so; do
  [ -f "$v" ] || continue
  echo "== $v"; HB_LIB="$PWD/$v" python profiles/prof_driver.py --case hempc --batch 4096 --reps 3 2>&1 | head -13
done ) | tee gpurun_out/variants_keymont.log
python -m pytest tests/test_gpu_shapes.py tests/test_gpu_parity.py tests/test_gpu_golden.py -m gpu -q -x -k "not config_C" 2>&1 | tail -3
