#!/usr/bin/env bash
# GPU call 2 of round 2: ncu launch list + --set full captures of the hot kernels, compute-sanitizer logs, variant A/B timings.
set -u
O=gpurun_out/r02; mkdir -p "$O/sanitizer"
DRV="python profiles/prof_driver.py"
$DRV --case hempc --batch 4096 --reps 2 > "$O/hempc_plain.log" 2>&1; echo "hempc rc=$?"; cat "$O/hempc_plain.log"
$DRV --case ntt --logn 16 --limbs 1024 > "$O/ntt16_plain.log" 2>&1; echo "ntt rc=$?"; cat "$O/ntt16_plain.log"
# --- variants (development builds, logn 12 / 16 only) ---
for v in build_variants/lib_*.so; do
  [ -f "$v" ] || continue
  echo "== $v"
  HB_LIB="$PWD/$v" $DRV --case hempc --batch 4096 --reps 2 2>&1 | head -8
  HB_LIB="$PWD/$v" $DRV --case ntt --logn 16 --limbs 1024 2>&1
done > "$O/variants.log" 2>&1
cat "$O/variants.log"
# --- ncu: launch list of one batched evaluation, then full captures ---
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$O/launches_hempc.csv" \
    $DRV --case hempc --batch 4096 --reps 1 > "$O/ncu_launches.log" 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on \
    -k regex:'k_ks_limb|k_ks_baby|k_ks_finish|k32_mulacc_intt_all|k32_recon_smod4|k32_poly_ntt_all' -s 9 -c 9 \
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
# --- compute-sanitizer: the smoke call (cluster-free N = 4096 path) and one N = 2^16 NTT (clusters + distributed shared memory) ---
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python -c "import __graft_entry__ as g; g.smoke()" > "$O/sanitizer/${tool}_smoke.log" 2>&1
  echo "$tool smoke rc=$?" | tee -a "$O/sanitizer/${tool}_smoke.log"
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 $DRV --case ntt --logn 16 --limbs 16 --reps 1 > "$O/sanitizer/${tool}_ntt16.log" 2>&1
  echo "$tool ntt16 rc=$?" | tee -a "$O/sanitizer/${tool}_ntt16.log"
done
tail -4 "$O"/sanitizer/*.log
ls -la "$O"
