#!/usr/bin/env bash
# GPU call 1 of round 2: full GPU test suite with durations, the bench line, and A/B lines for the two new he_gemv shortcuts.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q --durations=25 > gpurun_out/pytest1.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest1.log
tail -40 gpurun_out/pytest1.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench1.json 2> gpurun_out/bench1.err
echo "bench rc=$?"; tail -c 600 gpurun_out/bench1.err
FAST="--steps 6 --warmup 3 --no-ntt --no-cpu-baseline --no-parity-spot --no-closed-loop"
HB_GEMV_NOSUM=1 timeout 300 python bench.py $FAST > gpurun_out/bench1_nosum.json 2>> gpurun_out/bench1.err
HB_GEMV_NOSUM=1 HB_GEMV_NOSKIP=1 timeout 300 python bench.py $FAST > gpurun_out/bench1_nosum_noskip.json 2>> gpurun_out/bench1.err
HB_GEMV_NOSKIP=1 timeout 300 python bench.py $FAST > gpurun_out/bench1_noskip.json 2>> gpurun_out/bench1.err
python - <<'PY'
import json
for f in ("bench1", "bench1_nosum", "bench1_nosum_noskip", "bench1_noskip"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["value"]), "it/s  e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"], 2))
        for k, v in list(d["roofline"]["kernels"].items())[:9]:
            print("   ", k, v)
    except Exception as ex:
        print(f, "unreadable", ex)
PY
