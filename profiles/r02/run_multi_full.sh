#!/usr/bin/env bash
# the driver's multi-GPU command line, all secondary blocks on: usage run_multi_full.sh <N>
N="${1:-2}"
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
   bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_full_n$N.json 2> gpurun_out/bench_full_n$N.err
echo "rc=$?"; tail -c 300 gpurun_out/bench_full_n$N.err
python - <<PY
import json
d = [json.loads(l) for l in open("gpurun_out/bench_full_n$N.json") if l.startswith("{")][-1]
print("N=$N", round(d["value"]), "it/s  e2e", round(d["e2e"]["value"]), "keys", sorted(d.keys()))
PY
