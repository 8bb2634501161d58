#!/usr/bin/env bash
# Run on the GPU box (gpurun): tests, bench, ncu launch list, ncu --set full on the top kernels.
# usage: bash profiles/capture.sh <tag> [quick]
TAG="${1:-r01}"; MODE="${2:-full}"
O=gpurun_out/$TAG; mkdir -p "$O"
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > "$O/clocks.csv" 2>/dev/null &
SMI=$!
python -m pytest tests -m gpu -x -q > "$O/pytest_gpu.log" 2>&1; echo "pytest rc=$?" | tee -a "$O/pytest_gpu.log"
tail -3 "$O/pytest_gpu.log"
python profiles/prof_driver.py --case hempc --batch 4096 > "$O/hempc_plain.log" 2>&1; echo "hempc rc=$?"; cat "$O/hempc_plain.log"
python profiles/prof_driver.py --case ntt --logn 16 > "$O/ntt16_plain.log" 2>&1; echo "ntt rc=$?"; cat "$O/ntt16_plain.log"
for ln in 13 14 15; do python profiles/prof_driver.py --case ntt --logn $ln --limbs $((8192 >> (ln-13))) 2>&1 | sed "s/^/logn=$ln /"; done | tee "$O/ntt_small.log"
python profiles/prof_driver.py --case mul --logn 15 --logq 881 --batch 8 > "$O/mul15_plain.log" 2>&1; echo "mul rc=$?"; cat "$O/mul15_plain.log"
if [ "$MODE" = "full" ]; then
python bench.py --steps 3 --warmup 3 > "$O/bench.json" 2> "$O/bench.err"; echo "bench rc=$?"; cat "$O/bench.json"
fi
if [ "$MODE" != "quick" ]; then
python profiles/prof_driver.py --case hempc --batch 4096 --reps 1 > "$O/plain1.log" 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file "$O/launches_hempc.csv" \
    python profiles/prof_driver.py --case hempc --batch 4096 --reps 1 > "$O/ncu_launches.log" 2>&1
echo "ncu launches rc=$?"
python profiles/prof_driver.py --case hempc --batch 4096 --reps 1 > "$O/plain2.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_ks_limb|k_ks_baby|k32_mulacc_intt_all|k_ks_finish|k32_recon_smod4|k32_poly_ntt_all' -s 14 -c 14 \
    -o "$O/prof_hempc" -f python profiles/prof_driver.py --case hempc --batch 4096 --reps 1 > "$O/ncu_full_hempc.log" 2>&1
echo "ncu full hempc rc=$?"
python profiles/prof_driver.py --case ntt --logn 16 --reps 1 > "$O/plain3.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_ntt' -s 2 -c 2 \
    -o "$O/prof_ntt16" -f python profiles/prof_driver.py --case ntt --logn 16 --reps 1 > "$O/ncu_full_ntt.log" 2>&1
echo "ncu full ntt rc=$?"
fi
# the .ncu-rep files are too large to travel (64 MiB cap): export what we read, then drop reports over 24 MiB
for r in "$O"/*.ncu-rep; do
  [ -f "$r" ] || continue
  b="${r%.ncu-rep}"
  ncu -i "$r" --page raw --csv > "$b.raw.csv" 2>/dev/null
  ncu -i "$r" --page details --csv > "$b.details.csv" 2>/dev/null
  ncu -i "$r" --page source --csv --print-source sass 2>/dev/null | gzip -9 > "$b.source_sass.csv.gz"
  sz=$(stat -c %s "$r"); [ "$sz" -gt 25000000 ] && rm -f "$r"
done
kill $SMI 2>/dev/null
ls -la "$O"
