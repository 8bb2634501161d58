#!/usr/bin/env bash
mkdir -p gpurun_out/final
timeout 1500 python bench.py --steps 20 --warmup 3 > gpurun_out/final/bench.json 2> gpurun_out/final/bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/final/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final/bench_reference.json 2>> gpurun_out/final/bench.err; echo "reference arm rc=$?"
timeout 600 python bench.py > gpurun_out/final/bench_default.json 2>> gpurun_out/final/bench.err; echo "default invocation rc=$?"
