#!/usr/bin/env bash
# turn the files of gpurun_out/final (run_final*.sh) into the committed digests of profiles/r02 and profiles/traffic.json
set -u
cd "$(dirname "${BASH_SOURCE[0]}")/../.."
F=gpurun_out/final
python profiles/summarize.py $F/prof_hempc.raw.csv $F/prof_ntt16.raw.csv --batch 4096 > profiles/r02/ncu_summary.md
python profiles/sass_summary.py $F/prof_hempc.source_sass.csv.gz $F/prof_ntt16.source_sass.csv.gz > profiles/r02/sass_dynamic_mix.txt
names="k_ks_limbILi12E k_ks_babyILi12E k_ks_limb_sumILi12E k32_ks_babyILi12ELi2E k32_ks_limb_sumILi12ELi2E k_ks_finish_hybILi4ELi3ELi8E k_ks_finish_sum_hybILi4ELi3ELi8ELi3E k_ks_finish_fastILi2ELi4ELi7ELi3E k_ntt_fwdILi16ELb0E k32_mulacc_intt_allILi12ELi4E k32_poly_ntt_allILi12ELi2E"
(for kn in $names; do echo "== $kn (static SASS opcode histogram, cuobjdump -sass hectr_b200/libhectr_b200.so)"; bash profiles/sass_mix.sh hectr_b200/libhectr_b200.so $kn 16; done) > profiles/r02/sass_static_mix.txt
cp $F/prof_hempc.details.csv profiles/r02/hempc_details.csv
cp $F/prof_ntt16.details.csv profiles/r02/ntt16_details.csv
cp $F/launches_hempc.csv $F/hempc_plain.log $F/ntt16_plain.log profiles/r02/
tail -1 $F/bench.json > profiles/r02/bench.json
tail -1 $F/bench_reference.json > profiles/r02/bench_reference_arm.json
