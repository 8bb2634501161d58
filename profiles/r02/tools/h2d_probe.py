"""Platform probe behind the 8-GPU end-to-end number: aggregate pinned-host -> device (and back) bandwidth when 1, 2, 4, 8 GPUs
copy at the same time, one process, one stream and one pinned buffer pair per GPU.  Prints GB/s per direction and both at once."""
import subprocess
import sys
import time

import torch

MB = 256
n = torch.cuda.device_count()
print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout[:3000])
try:
    print(open("/sys/devices/system/node/online").read().strip(), "NUMA nodes online;",
          subprocess.run(["nproc"], capture_output=True, text=True).stdout.strip(), "CPUs")
except Exception as ex:
    print("numa:", ex)
host_in = [torch.empty(MB << 20, dtype=torch.uint8).pin_memory() for _ in range(n)]
host_out = [torch.empty(MB << 20, dtype=torch.uint8).pin_memory() for _ in range(n)]
dev = [torch.empty(MB << 20, dtype=torch.uint8, device=f"cuda:{i}") for i in range(n)]
dev2 = [torch.empty(MB << 20, dtype=torch.uint8, device=f"cuda:{i}") for i in range(n)]
s_in = [torch.cuda.Stream(device=i) for i in range(n)]
s_out = [torch.cuda.Stream(device=i) for i in range(n)]


def run(gpus, mode, reps=8):
    for i in gpus:
        torch.cuda.synchronize(i)
    t0 = time.perf_counter()
    for _ in range(reps):
        for i in gpus:
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s_in[i]):
                    dev[i].copy_(host_in[i], non_blocking=True)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s_out[i]):
                    host_out[i].copy_(dev2[i], non_blocking=True)
    for i in gpus:
        torch.cuda.synchronize(i)
    dt = time.perf_counter() - t0
    nbytes = len(gpus) * reps * (MB << 20) * (2 if mode == "both" else 1)
    return nbytes / dt / 1e9


for k in (1, 2, 4, 8):
    if k > n:
        break
    gpus = list(range(k))
    run(gpus, "both", 2)
    print(f"{k} GPUs: h2d {run(gpus, 'h2d'):7.1f} GB/s   d2h {run(gpus, 'd2h'):7.1f} GB/s   both {run(gpus, 'both'):7.1f} GB/s (sum of directions)")
if n >= 8:
    for gpus in ([0, 1], [0, 4], [0, 2, 4, 6], [4, 5, 6, 7]):
        print(f"GPUs {gpus}: h2d {run(gpus, 'h2d'):7.1f} GB/s")
