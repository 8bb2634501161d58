#!/usr/bin/env bash
# multi-GPU bench lines (strong scaling over 4096 instances): usage run_multi.sh <N>
N="${1:-2}"
mkdir -p gpurun_out
FAST="--steps 20 --warmup 3 --no-ntt --no-cpu-baseline --no-closed-loop"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
   bench.py --gpus $N $FAST > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "rc=$?"; tail -c 400 gpurun_out/bench_n$N.err
python - <<PY
import json
d = [json.loads(l) for l in open("gpurun_out/bench_n$N.json") if l.startswith("{")][-1]
print("N=$N", round(d["value"]), "it/s  e2e", round(d["e2e"]["value"]), "single call", round(d["e2e"]["single_blocking_call"]["value"]), "parity", d.get("parity_spot"))
PY
