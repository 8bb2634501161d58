#!/usr/bin/env bash
# ncu --set full over the nine hot kernels of ONE batched evaluation (the warm-up evaluation's nine are skipped)
set -u
O=gpurun_out/final; mkdir -p "$O"
DRV="python profiles/prof_driver.py"
$DRV --case hempc --batch 4096 --reps 1 > "$O/plain_again.log" 2>&1 || exit 1
ncu --set full --clock-control none --import-source on \
    -k regex:'k_ks_limb|k_ks_baby|k32_ks|k_ks_finish|k32_mulacc_intt_all|k32_recon_smod4|k32_poly_ntt_all' -s 9 -c 9 \
    -o "$O/prof_hempc" -f $DRV --case hempc --batch 4096 --reps 1 > "$O/ncu_full_hempc.log" 2>&1
echo "ncu full hempc rc=$?"
r="$O/prof_hempc.ncu-rep"; b="${r%.ncu-rep}"
ncu -i "$r" --page raw --csv > "$b.raw.csv" 2>/dev/null
ncu -i "$r" --page details --csv > "$b.details.csv" 2>/dev/null
ncu -i "$r" --page source --csv --print-source sass 2>/dev/null | gzip -9 > "$b.source_sass.csv.gz"
sz=$(stat -c %s "$r"); [ "$sz" -gt 20000000 ] && rm -f "$r"
ls -la "$O" | head -30
