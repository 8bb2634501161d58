#!/usr/bin/env bash
# Capture call for the shipped sources (tests ran in run4.sh): bench line, reference arm, ncu launch list + --set full captures
set -u
O=gpurun_out/final; mkdir -p "$O"
DRV="python profiles/prof_driver.py"
timeout 1200 python bench.py --steps 20 --warmup 3 > "$O/bench.json" 2> "$O/bench.err"; echo "bench rc=$?"; tail -c 300 "$O/bench.err"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > "$O/bench_reference.json" 2>> "$O/bench.err"; echo "reference arm rc=$?"
$DRV --case hempc --batch 4096 --reps 2 > "$O/hempc_plain.log" 2>&1; cat "$O/hempc_plain.log"
$DRV --case ntt --logn 16 --limbs 1024 > "$O/ntt16_plain.log" 2>&1; cat "$O/ntt16_plain.log"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$O/launches_hempc.csv" \
    $DRV --case hempc --batch 4096 --reps 1 > "$O/ncu_launches.log" 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on \
    -k regex:'k_ks_limb|k_ks_baby|k32_ks|k_ks_finish|k32_mulacc_intt_all|k32_recon_smod4|k32_poly_ntt_all' -s 9 -c 9 \
    -o "$O/prof_hempc" -f $DRV --case hempc --batch 4096 --reps 1 > "$O/ncu_full_hempc.log" 2>&1
echo "ncu full hempc rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_ntt' -s 2 -c 2 \
    -o "$O/prof_ntt16" -f $DRV --case ntt --logn 16 --reps 1 > "$O/ncu_full_ntt.log" 2>&1
echo "ncu full ntt rc=$?"
for r in "$O"/*.ncu-rep; do
  [ -f "$r" ] || continue
  b="${r%.ncu-rep}"
  ncu -i "$r" --page raw --csv > "$b.raw.csv" 2>/dev/null
  ncu -i "$r" --page details --csv > "$b.details.csv" 2>/dev/null
  ncu -i "$r" --page source --csv --print-source sass 2>/dev/null | gzip -9 > "$b.source_sass.csv.gz"
  sz=$(stat -c %s "$r"); [ "$sz" -gt 20000000 ] && rm -f "$r"
done
ls -la "$O"
