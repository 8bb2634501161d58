#!/usr/bin/env bash
# GPU call: full test suite, bench line, A/B without the hybrid basis, per-kernel times
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/pytest8.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest8.log
grep -vE "^\s*$" gpurun_out/pytest8.log | tail -25 | cut -c1-300
FAST="--steps 8 --warmup 3 --no-ntt --no-cpu-baseline --no-closed-loop"
timeout 600 python bench.py $FAST > gpurun_out/bench8.json 2> gpurun_out/bench8.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench8.err
HB_HEMPC_NOPAIR=1 timeout 600 python bench.py $FAST --no-parity-spot > gpurun_out/bench8_nopair.json 2>> gpurun_out/bench8.err
python - <<'PY'
import json
for f in ("bench8", "bench8_nopair"):
    try:
        d = [json.loads(l) for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1]
        print(f, round(d["value"]), "it/s  e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"], 2), "parity", d.get("parity_spot"))
        for k, v in list(d["roofline"]["kernels"].items())[:12]:
            print("   ", k, v["launches"], round(v["ms"] / d["steps"], 3), "ms/step")
    except Exception as ex:
        print(f, "unreadable", ex)
PY
