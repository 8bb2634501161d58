#!/usr/bin/env bash
# full GPU test suite, all failures reported, with durations
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --durations=30 > gpurun_out/pytest_full.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_full.log
grep -vE "^\s*$" gpurun_out/pytest_full.log | tail -70 | cut -c1-300
cat gpurun_out/closed_loop.txt 2>/dev/null
