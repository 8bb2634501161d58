#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
FAST="--steps 8 --warmup 3 --no-ntt --no-cpu-baseline --no-closed-loop"
timeout 600 python bench.py $FAST > gpurun_out/bench4.json 2> gpurun_out/bench4.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench4.err
python - <<'PY'
import json
for f in ("bench4",):
    try:
        d = [json.loads(l) for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1]
        print(f, round(d["value"]), "it/s  e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"], 2), "parity", d.get("parity_spot"))
        for k, v in list(d["roofline"]["kernels"].items())[:14]:
            print("   ", k, v["launches"], round(v["ms"] / d["steps"], 3), "ms/step")
    except Exception as ex:
        print(f, "unreadable", ex)
PY
for v in build_variants/lib_h*.so; do
  [ -f "$v" ] || continue
  echo "== $v"; HB_LIB="$PWD/$v" python profiles/prof_driver.py --case hempc --batch 4096 --reps 2 2>&1 | head -14
done
