#!/usr/bin/env bash
# last call of the round: whole GPU suite, then the capture of run_final_nopytest.sh
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_last.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_last.log
bash profiles/r02/run_final_nopytest.sh
