#!/usr/bin/env bash
# whole GPU suite on the shipped library, then the default bench invocation
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/pytest12.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest12.log
timeout 900 python bench.py > gpurun_out/bench12.json 2> gpurun_out/bench12.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench12.err
python - <<'PY'
import json
d = [json.loads(l) for l in open("gpurun_out/bench12.json") if l.startswith("{")][-1]
print("bench12", round(d["value"]), "it/s  e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"], 2), "traffic", d["roofline"].get("traffic"), "frac", d["roofline"]["frac"], "parity", d["parity_spot"]["all_equal"])
PY
