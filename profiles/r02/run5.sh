#!/usr/bin/env bash
# GPU call: variant timings, then ncu launch list + --set full captures of the hot kernels of the hybrid build
set -u
O=gpurun_out/r02b; mkdir -p "$O"
DRV="python profiles/prof_driver.py"
for v in build_variants/lib_v_*.so; do
  [ -f "$v" ] || continue
  echo "== $v"; HB_LIB="$PWD/$v" $DRV --case hempc --batch 4096 --reps 2 2>&1 | head -13
done > "$O/variants.log" 2>&1
cat "$O/variants.log"
$DRV --case hempc --batch 4096 --reps 2 > "$O/hempc_plain.log" 2>&1; echo "hempc rc=$?"; cat "$O/hempc_plain.log"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$O/launches_hempc.csv" \
    $DRV --case hempc --batch 4096 --reps 1 > "$O/ncu_launches.log" 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on \
    -k regex:'k_ks_limb|k_ks_baby|k32_ks|k_ks_finish|k32_mulacc_intt_all|k32_recon_smod4|k32_poly_ntt_all' -s 11 -c 11 \
    -o "$O/prof_hempc" -f $DRV --case hempc --batch 4096 --reps 1 > "$O/ncu_full_hempc.log" 2>&1
echo "ncu full hempc rc=$?"
for r in "$O"/*.ncu-rep; do
  [ -f "$r" ] || continue
  b="${r%.ncu-rep}"
  ncu -i "$r" --page raw --csv > "$b.raw.csv" 2>/dev/null
  ncu -i "$r" --page details --csv > "$b.details.csv" 2>/dev/null
  ncu -i "$r" --page source --csv --print-source sass 2>/dev/null | gzip -9 > "$b.source_sass.csv.gz"
  sz=$(stat -c %s "$r"); [ "$sz" -gt 20000000 ] && rm -f "$r"
done
ls -la "$O"
