#!/usr/bin/env bash
# GPU call: the allocation-free entry point alone in a fresh process, then the whole suite (it is collected last),
# then the bench line and the NTT table
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_zz_alone.log 2>&1
echo "zz alone rc=$?" | tee -a gpurun_out/pytest_zz_alone.log
tail -5 gpurun_out/pytest_zz_alone.log | cut -c1-300
timeout 2400 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/pytest10.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest10.log
grep -vE "^\s*$" gpurun_out/pytest10.log | tail -22 | cut -c1-300
timeout 900 python bench.py --steps 10 --warmup 3 --no-closed-loop > gpurun_out/bench10.json 2> gpurun_out/bench10.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench10.err
python - <<'PY'
import json
try:
    d = [json.loads(l) for l in open("gpurun_out/bench10.json") if l.startswith("{")][-1]
    print("bench10", round(d["value"]), "it/s  e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"], 2), "parity", d.get("parity_spot"))
    for k, v in list(d["roofline"]["kernels"].items())[:12]:
        print("   ", k, v["launches"], round(v["ms"] / d["steps"], 3), "ms/step")
    print(json.dumps(d.get("ntt_table", d.get("ntt")))[:1500])
except Exception as ex:
    print("unreadable", ex)
PY
