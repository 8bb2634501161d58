#!/usr/bin/env python
"""bench.py -- encrypted CSTR MPC iterations/s (BASELINE.json metric) on N B200s of one node.

One *step* = one pass of the ciphertext half of ctr_hempc (src/hempc.c:253-266: 2 he_sub, 2 he_gemv, he_add,
he_neg, he_copy_ct, he_moddown, he_add) over a batch of independent controller instances (config E: 4096
instances of the `cstr-hempc` parameter set N=4096, q=2^109, Delta=2^50, slots=16; shared keys and gain
matrices).  Instances are sharded contiguously across ranks (strong scaling, no data-path collective); with
N>1 the result ciphertexts are gathered with one NCCL all_gather per step.

    python bench.py [--gpus N --steps K --warmup W]          # our arm
    python bench.py --impl reference [...]                     # the reference's own CPU path (oracle/_ref), all host cores

Prints ONE JSON line (rank 0).  Inputs are synthetic: uniform random ciphertext words in the centred range of
q_l (what an RLWE ciphertext looks like), rotation keys made by the engine's own he_genswk from a sampled secret
(real switching keys: the bound-checked shortcuts of he_gemv apply exactly as they do to the reference's keys),
random gain matrices with the hempc block structure (8x3 and 8x2 blocks in a 16x16 frame, src/hempc.c:232-238).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PARAMS = dict(logn=12, logq=109, slots=16, logDelta=50)     # src/ctr.c:510-518
METRIC = "encrypted CSTR MPC iterations/s"
UNIT = "iterations/s"


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            d = json.load(open(path))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------------
# clocks sampler
# ------------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons DURING the timed region.  NVML (the library nvidia-smi prints from) polled every few
    milliseconds from a thread, so that even a 50 ms region (8 GPUs) gets samples; `nvidia-smi -lms` when pynvml is missing."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    MASKS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}   # nvml.h

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self.proc = None
        self.thread = None
        self.nvml = None
        self.stop_flag = False

    def _start_nvml(self) -> bool:
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            smax = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            reasons_fn = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            reasons_fn(h)
        except Exception:
            return False
        self.nvml = pynvml

        def run():
            while not self.stop_flag:
                try:
                    sm = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                    mask = int(reasons_fn(h))
                except Exception:
                    break
                f = [str(sm), str(smax)] + ["Active" if mask & self.MASKS[n] else "Not Active" for n in self.NAMES]
                self.samples.append(", ".join(f))
                time.sleep(0.010)
        self.thread = threading.Thread(target=run, daemon=True)
        self.thread.start()
        return True

    def start(self):
        if self._start_nvml():
            return
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index), "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        def run():
            for line in self.proc.stdout:
                self.samples.append(line.strip())
        self.thread = threading.Thread(target=run, daemon=True)
        self.thread.start()

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
        elif not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, smax, reasons = [], None, set()
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                smax = float(f[1])
            except ValueError:
                continue
            for name, v in zip(self.NAMES, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm), "source": "nvml (the library behind nvidia-smi), 10 ms period" if self.nvml is not None else "nvidia-smi -lms 100"}


# ------------------------------------------------------------------------------------------------------
# synthetic workload
# ------------------------------------------------------------------------------------------------------
def synth_ct_into(buf: np.ndarray, rng, k: int):
    """Fill uint64[B][2][W][N] with uniform two's-complement coefficients in [-2^(k-1), 2^(k-1))."""
    B, _, W, n = buf.shape
    chunk = max(1, (1 << 24) // (2 * W * n))
    for s in range(0, B, chunk):
        e = min(B, s + chunk)
        buf[s:e] = rng.integers(0, 1 << 64, size=(e - s, 2, W, n), dtype=np.uint64)
    top = k - 64 * (W - 1)
    if top < 64:
        t = buf[:, :, W - 1, :].view(np.int64)
        t <<= np.int64(64 - top)
        t >>= np.int64(64 - top)


def _poly_rot_i64(a: np.ndarray, rot: int) -> np.ndarray:
    """poly_rot (GPQHE/src/poly.c:263-283) on one signed word per coefficient: r[k] = a[i] (k = i 5^rot mod 2N < N), else r[k-N] = -a[i]."""
    n = a.shape[0]
    g = pow(5, rot, 2 * n)
    k = (np.arange(n, dtype=np.int64) * g) % (2 * n)
    out = np.empty_like(a)
    out[k % n] = np.where(k < n, a, -a)
    return out


def make_secret(rng, n, hw=64):
    s = np.zeros(n, dtype=np.int64)
    pos = rng.choice(n, size=hw, replace=False)
    s[pos] = rng.choice(np.array([-1, 1], dtype=np.int64), size=hw)
    return s


def make_rotation_keys(eng, rng, secret: np.ndarray, rots) -> np.ndarray:
    """REAL switching keys for the rotations `rots`, made by the engine's own he_genswk (hb_he_genswk, he-kem.c:74-119) from a
    sampled secret, error and uniform polynomial -- not uniform residues: a key is smod(., P q_L), which is what the exact-product
    shortcuts of he_gemv rely on (hb_evk_upload records the bound).  Returns uint64[slots][2][dimevk][N], zero where no key was asked."""
    rots = sorted(set(int(r) for r in rots))
    n, WA = eng.n, eng.genswk_words()
    nbits = eng.info.nbitsPqL
    a = rng.integers(0, 1 << 64, size=(len(rots), WA, n), dtype=np.uint64)
    top = nbits - 64 * (WA - 1)
    if top < 64:
        a[:, WA - 1, :] &= np.uint64((1 << top) - 1)
    e = np.rint(rng.normal(0.0, 3.19, size=(len(rots), n))).astype(np.int64)
    sp = np.stack([_poly_rot_i64(secret, r) for r in rots])
    dsk = eng.sk(secret)
    made = eng.he_genswk(a, e, sp, dsk)
    eng.lib.hb_sk_free(eng.ctx, C.byref(dsk))
    keys = np.zeros((eng.slots, 2, eng.info.dimevk, n), dtype=np.uint64)
    for i, r in enumerate(rots):
        keys[r] = made[i]
    return keys


def gemv_rotations(slots: int):
    """rotation amounts he_gemv uses (he-algo.c:57-80): the n1 baby steps and the multiples of n1"""
    n1 = int(np.sqrt(slots))
    if n1 * n1 != slots:
        n1 = int(np.sqrt(2 * slots))
    n2 = slots // n1
    return list(range(n1)) + [i * n1 for i in range(n2)]


def synth_gains(rng, slots, nx=3, nu=2, horizon=4):
    """Ka (mN x n) and Kb (mN x m) zero-padded to slots x slots, as ctr_hempc builds them (src/hempc.c:232-238)."""
    Ka = np.zeros((slots, slots), dtype=np.complex128)
    Kb = np.zeros((slots, slots), dtype=np.complex128)
    Ka[: nu * horizon, :nx] = rng.standard_normal((nu * horizon, nx))
    Kb[: nu * horizon, :nu] = rng.standard_normal((nu * horizon, nu))
    return Ka, Kb


def pinned_array(eng, shape):
    nbytes = int(np.prod(shape)) * 8
    p = eng.lib.hb_host_alloc(eng.ctx, nbytes)
    arr = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint64)), shape=(int(np.prod(shape)),)).reshape(shape)
    return arr, p


# ------------------------------------------------------------------------------------------------------
# roofline accounting: algorithmic bytes per UNIT of work; a unit may span several kernels
# ------------------------------------------------------------------------------------------------------
UNITS = {
    # one he_swk = read d0, d1 + write c0, c1 + the key shared by the batch; RNS intermediates count as on-chip.
    # The per-limb kernel and the per-coefficient tail are the two halves of the SAME key switches: the unit's time is the sum.
    "baby key switches (n1 he_rot of one ciphertext, he-algo.c:66-68)": ("keyswitch_n1", ("k_ks_baby", "k32_ks_baby", "k_ks_finish_hyb", "k_ks_finish_fast", "k_ks_finish")),
    "giant key switches + he_add chain + he_rs (he-algo.c:80-90)": ("keyswitch_ngs", ("k_ks_limb", "k_ks_limb_sum", "k32_ks_limb_sum", "k_ks_finish_sum", "k_ks_finish_sum_hyb", "k_ct_sum_rs", "k_ct_sum", "k_ct_rescale")),
    "ct x pt diagonal products (he_mulpt, he-algo.c:70-79)": ("mulpt", ("k32_poly_ntt_all", "k32_poly_ntt", "k32_mulacc_intt_all", "k32_mulacc_intt", "k32_recon_smod4", "k32_recon_smod",
                                                                         "k_poly_ntt", "k_mulacc_intt", "k_recon_smod", "k_mulpt_limb")),
    "coefficient-wise passes (he_sub, he_add, he_neg, he_moddown)": ("elementwise", ("k_ct_addsub", "k_ct_combine", "k_ct_addpt")),
}


def unit_bytes(kind: str, B: int, n: int, W: int, Wout: int, D_ks: int, D_pt: int, n1: int, n2: int, ngs: int):
    """SURVEY.md 8(d) per-unit figures x the units ONE he_gemv of B instances performs (op-granular: every he_* op reads its
    inputs and writes its output once)."""
    ct = 2 * n * W * 8                                  # one ciphertext at level l
    key = 2 * D_ks * n * 8                              # one evaluation key, shared by the batch
    if kind == "keyswitch_n1":
        return n1 * (B * 2 * ct + key)
    if kind == "keyswitch_ngs":
        return ngs * (B * 2 * ct + key)
    if kind == "mulpt":
        return n1 * n2 * (B * 2 * ct + n * W * 8)       # slots products: ct in, ct out, one plaintext polynomial (shared)
    if kind == "elementwise":
        return B * (3 * ct + (2 * ct + 2 * n * Wout * 8)) / 2     # per gemv: one he_sub (3 ct) + half of the final combination
    return None


def sources_digest() -> str:
    """SHA-256 over the kernel sources: stamps ncu captures (profiles/traffic.json) so that a stale capture is not reported."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "hectr_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".hpp", ".cpp", ".c")):
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def prof_report(eng):
    lib, ctx = eng.lib, eng.ctx
    need = lib.hb_prof_report(ctx, None, 0)
    buf = C.create_string_buffer(need + 1)
    lib.hb_prof_report(ctx, buf, need + 1)
    return json.loads(buf.value.decode())


def run_ours(args):
    import torch
    import torch.distributed as dist
    from hectr_b200 import Engine

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    affinity = "unchanged"
    try:                                   # run (and first-touch the pinned staging buffers) on the CPUs next to this rank's GPU: with one
        import pynvml                      # process per GPU the host copies of 8 ranks otherwise cross the socket interconnect
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
        affinity = f"nvmlDeviceSetCpuAffinity: {len(os.sched_getaffinity(0))} CPUs"
    except Exception as ex:  # pragma: no cover
        affinity = f"unchanged ({type(ex).__name__})"
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")      # keeps NCCL's version banner off stdout (one JSON line only)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from hectr_b200.sharding import gather_results, shard_range
    total = args.instances
    lo, hi = shard_range(total, rank, world)        # contiguous block of controller instances for this GPU
    per = hi - lo
    assert per * world == total, "instances must divide by the number of GPUs"

    eng = Engine(PARAMS["logn"], PARAMS["logq"], PARAMS["slots"], PARAMS["logDelta"], device=local)
    lib, ctx = eng.lib, eng.ctx
    stream = torch.cuda.current_stream()
    lib.hb_ctx_set_stream(ctx, C.c_void_p(stream.cuda_stream))
    l = eng.L
    W, k, n = eng.words(l), eng.logq(l), eng.n
    Wout = eng.words(l - 1)
    rng = np.random.default_rng(0x48454354 + rank)
    krng = np.random.default_rng(7)                  # the same secret and keys on every rank (keys are replicated)
    secret = make_secret(krng, n)
    key_arr = make_rotation_keys(eng, krng, secret, gemv_rotations(eng.slots))
    keys = eng.keys(key_arr)
    Ka, Kb = synth_gains(np.random.default_rng(9), eng.slots)

    names = ("xhat", "xr", "uhat", "ur")
    host, hptr = {}, {}
    for nm in names:
        host[nm], hptr[nm] = pinned_array(eng, (per, 2, W, n))
        synth_ct_into(host[nm], rng, k)
    out_host, out_ptr = [], []
    for _ in range(2):
        a, p = pinned_array(eng, (per, 2, Wout, n))
        out_host.append(a)
        out_ptr.append(p)
    dev = {nm: eng.ct(batch=per) for nm in names}
    Delta = 2.0 ** PARAMS["logDelta"]
    for nm in names:
        lib.hb_ct_upload(ctx, C.byref(dev[nm].h), host[nm].ctypes.data_as(C.c_void_p), l, Delta, 1.0)
    up = eng.ct(batch=per)
    # result gather (the only collective, north_star): two buffer sets, asynchronous, so that the gather of step k overlaps
    # the evaluation of step k+1; every gather is waited for before the timed region closes
    equal = total % world == 0
    nbuf = 2 if (world > 1 and equal) else 1
    gather_in = [torch.empty((per, 2 * Wout * n), dtype=torch.int64, device="cuda") for _ in range(nbuf)]
    gather_out = [torch.empty((total, 2 * Wout * n), dtype=torch.int64, device="cuda") for _ in range(nbuf)] if (world > 1 and equal) else []
    works = [None] * nbuf
    counter = [0]

    def drain():
        for i, w in enumerate(works):
            if w is not None:
                w.wait()          # the current stream waits for the collective; the host does not block
                works[i] = None

    def step():
        eng.ctr_hempc_eval(Ka, Kb, dev["xhat"], dev["xr"], dev["uhat"], dev["ur"], keys, out=up)
        if world > 1:
            b = counter[0] % nbuf
            counter[0] += 1
            if works[b] is not None:
                works[b].wait()
                works[b] = None
            lib.hb_ct_export(ctx, C.byref(up.h), C.c_void_p(gather_in[b].data_ptr()))
            if equal:
                _, works[b] = gather_results(gather_in[b], total, out=gather_out[b], async_op=True)
            else:
                gather_results(gather_in[b], total)

    def barrier():
        drain()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    lib.hb_prof_reset(ctx)
    lib.hb_prof_enable(ctx, 1)
    launches0 = eng.launches()
    lib.hb_timer_start(ctx)
    for _ in range(args.steps):
        step()
    drain()                                   # the last gathers end inside the timed region
    ms = float(lib.hb_timer_stop_ms(ctx))
    barrier()
    launches = eng.launches() - launches0
    lib.hb_prof_enable(ctx, 0)
    prof = prof_report(eng)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = total * args.steps / (ms / 1e3)
    dev_out = up.download() if rank == 0 else None        # the benched result, for the parity spot check below

    # ---- end to end through the C ABI with host buffers: every step uploads its 4 input batches from pinned host memory and
    # downloads its result batch.  Streaming interface: step k+1 is submitted before step k is waited for, so its upload
    # overlaps the evaluation of step k (two device buffer sets inside the library); the timed region ends when the last
    # result is in host memory.
    def submit(i):
        return lib.hb_ctr_hempc_submit_host(ctx, per, l, Ka.ctypes.data_as(C.c_void_p), Kb.ctypes.data_as(C.c_void_p),
                                            C.c_void_p(hptr["xhat"]), C.c_void_p(hptr["xr"]), C.c_void_p(hptr["uhat"]),
                                            C.c_void_p(hptr["ur"]), keys.arr, C.c_void_p(out_ptr[i & 1]))

    def e2e_run(nsteps):
        prev = None
        for i in range(nsteps):
            t = submit(i)
            if prev is not None:
                lib.hb_ctr_hempc_wait(ctx, prev)
            prev = t
        lib.hb_ctr_hempc_wait(ctx, prev)

    e2e_run(3)
    barrier()
    t0 = time.perf_counter()
    e2e_run(args.steps)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = total * args.steps / e2e_s
    h2d = 4 * per * 2 * W * n * 8
    d2h = per * 2 * Wout * n * 8
    e2e_same = bool(np.array_equal(out_host[(args.steps - 1) & 1], dev_out)) if rank == 0 else None
    # one blocking call (hb_ctr_hempc_eval_host: chunks pipelined inside the call), for the latency-bound caller
    def e2e_call():
        lib.hb_ctr_hempc_eval_host(ctx, per, l, Ka.ctypes.data_as(C.c_void_p), Kb.ctypes.data_as(C.c_void_p),
                                   C.c_void_p(hptr["xhat"]), C.c_void_p(hptr["xr"]), C.c_void_p(hptr["uhat"]),
                                   C.c_void_p(hptr["ur"]), keys.arr, C.c_void_p(out_ptr[0]))
    e2e_call()
    barrier()
    t0 = time.perf_counter()
    for _ in range(max(2, args.steps // 4)):
        e2e_call()
    e2e_call_s = (time.perf_counter() - t0) / max(2, args.steps // 4)
    if world > 1:
        t = torch.tensor([e2e_call_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_call_s = float(t.item())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line_tail = finish_line(args, eng, keys, key_arr, secret, Ka, Kb, host, dev_out, prof, per, world, local, l, W, Wout, k, n)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic (uniform random ciphertext words; rotation keys made by hb_he_genswk from a sampled secret; "
                                 "random 8x3 / 8x2 gain blocks in a 16x16 frame as ctr_hempc builds them)",
        "config": {"workload": f"{total} independent cstr-hempc controller instances (BASELINE configs[4] on configs[0] parameters)",
                   "logn": PARAMS["logn"], "logq": PARAMS["logq"], "logDelta": PARAMS["logDelta"], "slots": PARAMS["slots"],
                   "instances": total, "instances_per_gpu": per, "parallelism": f"instances sharded x{world}, keys replicated",
                   "l2": f"inputs larger than L2 ({4 * per * 2 * W * n * 8 >> 20} MiB of ciphertext per GPU per step)"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "api": "hb_ctr_hempc_submit_host / hb_ctr_hempc_wait (pinned host buffers in and out; step k+1 is submitted before step k "
                       "is waited for, so its upload overlaps the evaluation of step k; every launch keeps the full per-GPU batch)",
                "result_equals_device_resident_run": e2e_same, "host_affinity": affinity,
                "results_end_in": "each rank's own pinned host buffer (one process per GPU, no gather on this path)" if world > 1 else "the caller's pinned host buffer",
                "single_blocking_call": {"api": "hb_ctr_hempc_eval_host (chunks pipelined inside one call)",
                                         "value": total / e2e_call_s, "unit": UNIT}},
        "gpu_launches": launches,
        "clocks": clocks,
        **line_tail,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def finish_line(args, eng, keys, key_arr, secret, Ka, Kb, host, dev_out, prof, per, world, local, l, W, Wout, k, n):
    """Everything of the JSON line that is derived after the timed regions (rank 0): roofline, parity spot check, CPU baselines,
    secondary blocks."""
    lib, ctx = eng.lib, eng.ctx
    Delta = 2.0 ** PARAMS["logDelta"]
    peak, peak_src = load_peaks()
    D_ks = lib.hb_dim_keyswitch(ctx, l)
    D_pt = lib.hb_dim_mulpt(ctx, l, Delta)
    S = eng.slots
    n1 = int(np.sqrt(S)) if int(np.sqrt(S)) ** 2 == S else int(np.sqrt(2 * S))
    n2 = S // n1
    # giant steps with a non-zero diagonal (what the plan evaluates), per matrix
    def nz_giant(M):
        cnt = 0
        for i in range(n2):
            if any(np.any(M[np.arange(S), (np.arange(S) + i * n1 + j) % S] != 0) for j in range(n1)):
                cnt += 1
        return cnt or n2
    ngs_avg = (nz_giant(Ka) + nz_giant(Kb)) / 2.0
    steps = args.steps
    tot_ms = sum(v["ms"] for v in prof.values()) or 1.0
    units = {}
    for uname, (kind, kernels) in UNITS.items():
        ms_u = sum(prof[kn]["ms"] for kn in kernels if kn in prof)
        if ms_u == 0:
            continue
        nbytes = unit_bytes(kind, per, n, W, Wout, D_ks, D_pt, n1, n2, ngs_avg)
        per_exec_ms = ms_u / (2 * steps)                       # two he_gemv per iteration
        units[uname] = {"kernels": [kn for kn in kernels if kn in prof], "ms_per_step": round(ms_u / steps, 4), "share_of_step": round(ms_u / tot_ms, 4),
                        "algorithmic_bytes_per_gemv": nbytes, "achieved_gbs": nbytes / (per_exec_ms / 1e3) / 1e9, "frac_of_hbm": nbytes / (per_exec_ms / 1e3) / 1e9 / peak}
    dom = max(prof, key=lambda kname: prof[kname]["ms"])
    dom_unit = next((u for u, (kind, kernels) in UNITS.items() if dom in kernels), None)
    if dom_unit is None:                      # a kernel this table does not know: report the largest known unit instead
        dom_unit = max(units, key=lambda u: units[u]["ms_per_step"])
    du = units[dom_unit]
    traffic, heavy, traffic_note = None, None, None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            tj = json.load(open(tpath))
            if tj.get("sources_digest") != sources_digest():
                traffic_note = "stale: profiles/traffic.json was captured from other kernel sources"
            else:
                rec = tj.get("kernels", {}).get(dom, {})
                traffic, heavy = rec.get(str(per)), rec.get("fmaheavy_pct")
                if traffic is None:
                    traffic_note = f"no ncu capture of {dom} at batch {per}"
        except Exception as ex:
            traffic_note = f"unreadable: {ex}"
    # whole step against the two SURVEY 8(d) byte counts
    it_per_s = per * steps / (tot_ms / 1e3)
    op_granular, fused_bound = 43.3 * 2 ** 20, 576 * 2 ** 10
    roofline = {"bound": "hbm", "kernel": dom, "unit": "GB/s", "peak": peak, "peak_source": peak_src,
                "achieved": du["achieved_gbs"], "frac": du["frac_of_hbm"], "traffic": traffic, "traffic_note": traffic_note,
                "unit_of_work": dom_unit, "unit_kernels": du["kernels"],
                # one launch of the dominant kernel covers 2 / lps matrix-vector products (both of ctr_hempc's when they are paired)
                "algorithmic_bytes_per_launch": du["algorithmic_bytes_per_gemv"] * 2 * steps / prof[dom]["launches"],
                "avg_launch_ms": sum(prof[kn]["ms"] for kn in du["kernels"]) / prof[dom]["launches"],
                "matvecs_per_launch": 2 * steps / prof[dom]["launches"],
                "dominant_kernel_avg_launch_ms": prof[dom]["ms"] / prof[dom]["launches"],
                "share_of_step": prof[dom]["ms"] / tot_ms,
                "units": units,
                "kernels": {kname: {"launches": v["launches"], "ms": round(v["ms"], 3), "share": round(v["ms"] / tot_ms, 4)}
                            for kname, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])},
                "step": {"op_granular_bytes_per_iteration": op_granular, "op_granular_gbs": it_per_s * op_granular / 1e9,
                         "op_granular_frac": it_per_s * op_granular / 1e9 / peak,
                         "fused_bound_bytes_per_iteration": fused_bound, "fused_bound_frac": it_per_s * fused_bound / 1e9 / peak},
                "int_pipe": {"unit": "sm__pipe_fmaheavy_cycles_active % of peak (the pipe IMAD/IMAD.WIDE issue to)", "pct": heavy},
                "note": "achieved = SURVEY 8(d) algorithmic bytes of the unit of work the dominant kernel belongs to (a key switch is a per-limb "
                        "kernel plus a per-coefficient tail: their times are added, the bytes counted once) / CUDA-event time of those kernels "
                        "inside the timed region.  59-bit modular butterflies bind on the integer multiplier, not on HBM (DESIGN.md 4)."}
    out = {"roofline": roofline}
    if not args.no_parity_spot:
        try:
            out["parity_spot"] = parity_spot(eng, key_arr, Ka, Kb, host, dev_out, l, idx=sorted({0, per - 1}))
        except Exception as ex:  # pragma: no cover
            out["parity_spot"] = {"error": repr(ex)}
    if world == 1 and not args.no_cpu_baseline:
        cpus = cpu_baselines(args.cpu_evals)
        out["cpu_baseline"] = cpus.get("Og")
        out["cpu_baseline_O2"] = cpus.get("O2")
        try:
            out["cpu_baseline_all_cores"] = cpu_baseline_all_cores()
        except Exception as ex:  # pragma: no cover
            out["cpu_baseline_all_cores"] = {"error": repr(ex)}
    else:
        out["cpu_baseline"] = None
    if not args.no_ntt:
        try:
            out["ntt"] = ntt_microbench(local, peak)
        except Exception as ex:  # pragma: no cover
            out["ntt"] = {"error": repr(ex)}
        for nm, fn in (("client_inclusive", lambda: client_inclusive_bench(eng, keys, secret, Ka, Kb)),
                       ("config_C", lambda: config_c_bench(local)), ("config_D", lambda: config_d_bench(local)),
                       ("plan_rebuild", lambda: plan_rebuild_bench(eng, keys))):
            try:
                out[nm] = fn()
            except Exception as ex:  # pragma: no cover
                out[nm] = {"error": repr(ex)}
        if world == 1 and not args.no_closed_loop:
            out["closed_loop"] = closed_loop_bench()
    return out


def parity_spot(eng, key_arr, Ka, Kb, host, dev_out, l, idx):
    """Checker leg: instances `idx` of the BENCHED result (device-resident timed run, also compared with the host-pipeline run
    above) are re-evaluated by the oracle -- the unmodified reference's he_sub / he_gemv / ... on the same input words and the same
    keys -- and compared word for word."""
    from oracle import ref as R
    if not R.available():
        return {"checked": 0, "reason": "oracle/_ref not built"}
    ref = R.Ref.get(PARAMS["logn"], PARAMS["logq"], PARAMS["slots"], PARAMS["logDelta"])
    rk = ref.rk_from_array(key_arr)
    Delta = 2.0 ** PARAMS["logDelta"]
    ok = []
    t0 = time.perf_counter()
    for b in idx:
        c = {nm: ref.ct_from_words(host[nm][b], l, nu=Delta, B=1.0) for nm in host}
        xd, ud = ref.sub(c["xhat"], c["xr"]), ref.sub(c["uhat"], c["ur"])
        du = ref.neg(ref.add(ref.gemv(Ka, xd, rk), ref.gemv(Kb, ud, rk)))
        want = ref.ct_to_words(ref.add(ref.moddown(ref.copy_ct(c["uhat"])), du))
        ok.append(bool(np.array_equal(dev_out[b], want)))
    return {"instances": list(map(int, idx)), "bit_exact": ok, "all_equal": all(ok), "oracle": "oracle/_ref/libgpqhe.so (unmodified reference)",
            "oracle_seconds": round(time.perf_counter() - t0, 1)}


# ------------------------------------------------------------------------------------------------------
# secondary metric: N=2^16 batched forward NTT limbs/s (BASELINE configs[1])
# ------------------------------------------------------------------------------------------------------
NTT_RINGS = {13: 218, 14: 438, 15: 881, 16: 1023}     # logn -> the largest log2 q the reference's table admits (capped at 1023)


def ntt_microbench(device: int, peak: float, reps: int = 5):
    """BASELINE configs[1]: batched forward / inverse NTT, N = 2^13..2^16, the reference's prime chain of each ring, 1 / 8 / 24
    primes per launch (16 at N = 2^13: that ring's chain is 16 primes long), >= 512 MiB in + out per launch.  limbs/s, fraction
    of the HBM roofline (16 N bytes per limb) and butterflies per clock per SM (integer ceiling 4.1, profiles/ubench)."""
    from hectr_b200 import Engine
    import torch
    sm = torch.cuda.get_device_properties(device).multi_processor_count
    try:
        import pynvml
        pynvml.nvmlInit()
        mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(pynvml.nvmlDeviceGetHandleByIndex(device), pynvml.NVML_CLOCK_SM))
    except Exception:
        mhz = 1965.0
    table = {}
    for logn, logq in NTT_RINGS.items():
        eng = Engine(logn, logq, 16, 50, device=device)
        lib, ctx = eng.lib, eng.ctx
        n = eng.n
        primes = eng.primes()
        nlimbs = (1 << 26) // n                              # 512 MiB in, 512 MiB out
        rng = np.random.default_rng(3 + logn)
        host = rng.integers(0, 1 << 59, size=(nlimbs, n), dtype=np.uint64)      # below every prime of the chain
        d_in = eng.to_device(host)
        d_out = eng.dev_alloc(host.nbytes)
        for D in (1, 8, 24):
            De = min(D, len(primes))
            res = {}
            for name, fn in (("forward", lambda: lib.hb_ntt_forward(ctx, d_in, d_out, nlimbs, 0, De, 0)),
                             ("inverse", lambda: lib.hb_ntt_inverse(ctx, d_in, d_out, nlimbs, 0, De))):
                for _ in range(3):
                    fn()
                eng.sync()
                lib.hb_timer_start(ctx)
                for _ in range(reps):
                    fn()
                ms = float(lib.hb_timer_stop_ms(ctx)) / reps
                lps = nlimbs / (ms / 1e3)
                gbs = 16.0 * n * lps / 1e9
                res[name] = {"limbs_per_s": lps, "ms_per_launch": ms, "achieved_gbs": gbs, "frac_of_hbm": gbs / peak,
                             "butterflies_per_clk_per_sm": lps * (n // 2) * logn / (sm * mhz * 1e6)}
            table[f"logn{logn}_D{D}"] = {"primes": De, **res}
        eng.dev_free(d_in)
        eng.dev_free(d_out)
        eng.close()
    head = table["logn16_D8"]
    return {"metric": "N=2^16 NTT limbs/s", "logn": 16, "limbs_per_launch": (1 << 26) // (1 << 16), "primes": 8,
            "algorithmic_bytes_per_limb": 16 << 16, "l2": "512 MiB in + 512 MiB out per launch (> L2)",
            "forward": head["forward"], "inverse": head["inverse"], "sm_clock_mhz_assumed": mhz, "table": table,
            "parity": "tests/test_gpu_ntt.py, tests/test_gpu_shapes.py::test_ntt_limb_counts (every ring, 1 / 8 / 24 primes per launch)"}


# ------------------------------------------------------------------------------------------------------
# secondary: BASELINE configs[2] (ct x ct + relinearise + rescale, rotation at N=2^15, q=2^881) and configs[3]
# (rotation-sum matvec at 32/64/128 slots).  Parity: tests/test_gpu_shapes.py::test_config_C_full_level_chain,
# test_config_D_gemv_64_128_slots, tests/test_gpu_parity.py::test_config_D_gemv_32_slots.
# ------------------------------------------------------------------------------------------------------
def _rand_ct(rng, batch, W, n, k):
    buf = np.empty((batch, 2, W, n), dtype=np.uint64)
    synth_ct_into(buf, rng, k)
    return buf


def config_c_bench(device: int, batch: int = 8, reps: int = 3):
    from hectr_b200 import Engine
    eng = Engine(15, 881, 2, 50, device=device)
    lib, ctx = eng.lib, eng.ctx
    rng = np.random.default_rng(5)
    l = eng.L
    W, k = eng.words(l), eng.logq(l)
    secret = make_secret(rng, eng.n)
    keys = eng.keys(make_rotation_keys(eng, rng, secret, [0, 1]))       # key 0 stands in for the relinearisation key (same shape and bound)
    x, y = eng.ct(_rand_ct(rng, batch, W, eng.n, k), l), eng.ct(_rand_ct(rng, batch, W, eng.n, k), l)
    out = eng.ct(batch=batch)
    res = {"logn": 15, "logq": 881, "level": l, "batch": batch, "limbs_keyswitch": int(lib.hb_dim_keyswitch(ctx, l)),
           "limbs_mul": int(lib.hb_dim_mul(ctx, l))}
    for name, fn in (("mul_relin_rs", lambda: (eng.he_mul(x, y, keys, out=out), eng.he_rs(out))),
                     ("rot", lambda: eng.he_rot(x, 1, keys))):
        fn()
        eng.sync()
        lib.hb_timer_start(ctx)
        for _ in range(reps):
            fn()
        ms = float(lib.hb_timer_stop_ms(ctx)) / reps
        res[name] = {"ms_per_batch": ms, "per_s": batch / (ms / 1e3)}
    keys.free()
    eng.close()
    return res


def config_d_bench(device: int, batch: int = 256, reps: int = 2):
    from hectr_b200 import Engine
    res = {"logn": 12, "logq": 109, "batch": batch}
    for slots in (32, 64, 128):
        eng = Engine(12, 109, slots, 50, device=device)
        lib, ctx = eng.lib, eng.ctx
        rng = np.random.default_rng(6)
        l = eng.L
        secret = make_secret(rng, eng.n)
        keys = eng.keys(make_rotation_keys(eng, rng, secret, gemv_rotations(slots)))
        A = rng.standard_normal((slots, slots)) + 1j * rng.standard_normal((slots, slots))
        x = eng.ct(_rand_ct(rng, batch, eng.words(l), eng.n, eng.logq(l)), l)
        out = eng.ct(batch=batch)
        eng.he_gemv(A, x, keys, out=out)          # builds the plan (host encode of the diagonals)
        eng.sync()
        lib.hb_timer_start(ctx)
        for _ in range(reps):
            eng.he_gemv(A, x, keys, out=out)
        ms = float(lib.hb_timer_stop_ms(ctx)) / reps
        res[f"slots{slots}"] = {"ms_per_batch": ms, "gemv_per_s": batch / (ms / 1e3), "matrix": "dense complex"}
        keys.free()
        eng.close()
    return res


def client_inclusive_bench(eng, keys, secret, Ka, Kb, batch: int = 512, reps: int = 3):
    """M1 with the client side: per iteration 5 he_enc_pk (up, xhat, uhat, xr, ur: src/ctr.c:445-481), the evaluation
    and he_dec, all on the device through the C ABI with host arrays in and out.  Host-side sampling and he_ecd / he_dcd are NOT
    included: the samples (v, e0, e1) and the encoded plaintexts are prepared once."""
    lib, ctx = eng.lib, eng.ctx
    rng = np.random.default_rng(11)
    n, L = eng.n, eng.L
    WC = eng.info.words_top
    # a public key consistent with the secret is not needed for timing: any two polynomials mod q_L
    pkw = _rand_ct(rng, 1, WC, n, eng.logq(L))[0]
    pk = eng.pk(pkw[0], pkw[1])
    dsk = eng.sk(secret)
    pts = np.zeros((batch, WC, n), dtype=np.uint64)
    pts[:, 0, :: n // (2 * eng.slots)] = rng.integers(0, 1 << 50, size=(batch, 2 * eng.slots), dtype=np.uint64)
    v = rng.integers(-1, 2, size=(batch, n)).astype(np.int64)
    e0 = np.rint(rng.normal(0, 3.19, size=(batch, n))).astype(np.int64)
    e1 = np.rint(rng.normal(0, 3.19, size=(batch, n))).astype(np.int64)

    def once():
        cts = [eng.he_enc_pk(pts, pk, v, e0, e1) for _ in range(5)]
        up = eng.ctr_hempc_eval(Ka, Kb, cts[1], cts[3], cts[2], cts[4], keys, out=cts[0])
        m = eng.he_dec(up, dsk)
        for c in cts:
            c.free()
        return m

    once()
    t0 = time.perf_counter()
    for _ in range(reps):
        once()
    dt = (time.perf_counter() - t0) / reps
    lib.hb_evk_free(ctx, C.byref(pk))
    lib.hb_sk_free(ctx, C.byref(dsk))
    return {"value": batch / dt, "unit": UNIT, "batch": batch, "per_iteration": "5 hb_he_enc_pk + hb_ctr_hempc_eval + hb_he_dec, host arrays in and out",
            "excluded": "host-side sampling (sample_zo, sample_error) and he_ecd / he_dcd"}


def plan_rebuild_bench(eng, keys, reps: int = 3):
    """A controller that re-linearises every step hands he_gemv a new matrix each call: the GemvPlan (host encode of `slots`
    diagonals, upload, their transforms) is rebuilt.  Cost of that rebuild, per matrix."""
    rng = np.random.default_rng(13)
    l = eng.L
    x = eng.ct(_rand_ct(rng, 1, eng.words(l), eng.n, eng.logq(l)), l)
    out = eng.ct(batch=1)
    ts = []
    for _ in range(reps):
        A = rng.standard_normal((eng.slots, eng.slots)) + 0j
        eng.sync()
        t0 = time.perf_counter()
        eng.he_gemv(A, x, keys, out=out)
        eng.sync()
        t1 = time.perf_counter()
        eng.he_gemv(A, x, keys, out=out)
        eng.sync()
        ts.append((t1 - t0) - (time.perf_counter() - t1))
    x.free()
    out.free()
    return {"ms_per_new_matrix": 1e3 * float(np.median(ts)), "what": "first he_gemv with a new matrix minus a repeat call (batch 1)"}


def closed_loop_bench():
    """BASELINE configs[0]: the reference's own `test-hectr cstr-hempc` binary (40 closed-loop steps, key generation included)
    with the GPQHE evaluation symbols bound to this library (integration/gpqhe_b200_shim.c); client side included."""
    import tempfile
    exe = os.path.join(ROOT, "integration", "_build", "test-hectr-b200")
    if not os.path.exists(exe):
        return {"unavailable": "integration/_build/test-hectr-b200 is not built"}
    d = tempfile.mkdtemp(prefix="hectr_bench_")
    os.makedirs(os.path.join(d, "results"))
    t0 = time.perf_counter()
    p = subprocess.run([exe, "cstr-hempc"], cwd=d, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=900)
    dt = time.perf_counter() - t0
    if p.returncode != 0:
        return {"error": p.stderr.decode()[-500:]}
    txt = p.stdout.decode(errors="replace")
    import re
    m = re.findall(r"(\d+)s\s+(\d+)ms", txt)
    return {"workload": "./test-hectr cstr-hempc (unmodified reference driver + libhectr.so, evaluation through the shim)", "steps": 40,
            "wall_s": dt, "steps_per_s_wall": 40 / dt,
            "note": "wall time includes he_keypair + he_genrk on the CPU (reference code) and 200 he_ecd; parity: tests/test_gpu_closed_loop.py"}


# ------------------------------------------------------------------------------------------------------
# CPU baseline: the reference's own code (oracle/_ref), timed on this box's host cores
# ------------------------------------------------------------------------------------------------------
def _ref_setup(opt="Og"):
    from oracle.ref import Ref
    ref = Ref.get(PARAMS["logn"], PARAMS["logq"], PARAMS["slots"], PARAMS["logDelta"], opt=opt)
    pk, sk = ref.keypair()
    rk = ref.genrk(sk)
    rng = np.random.default_rng(9)
    Ka, Kb = synth_gains(rng, PARAMS["slots"])
    cts = {}
    for nm in ("xhat", "xr", "uhat", "ur"):
        v = np.zeros(PARAMS["slots"], dtype=np.complex128)
        v[:3] = rng.standard_normal(3)
        cts[nm] = ref.enc_pk(ref.encode(v), pk)
    return ref, rk, Ka, Kb, cts


def _ref_eval(ref, rk, Ka, Kb, cts):
    """The ciphertext half of ctr_hempc through the reference's API, in the reference's order (hempc.c:253-266)."""
    xd, ud = ref.sub(cts["xhat"], cts["xr"]), ref.sub(cts["uhat"], cts["ur"])
    a, b = ref.gemv(Ka, xd, rk), ref.gemv(Kb, ud, rk)
    du = ref.neg(ref.add(a, b))
    uc = ref.moddown(ref.copy_ct(cts["uhat"]))
    up = ref.add(uc, du)
    for c in (xd, ud, a, b, du, uc, up):
        ref.free_ct(c)


def cpu_baselines(n_eval=3):
    """The unmodified reference on ONE host core, at the reference's own flags (-Og) and at -O2: one process per
    build (both export the same symbols, so they cannot share a process), run side by side on two cores.  The all-cores figure
    is the reference arm (`bench.py --impl reference`)."""
    from oracle import ref as R
    if not R.available():
        return {"Og": {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}}
    import multiprocessing as mp
    mpctx = mp.get_context("spawn")
    flags = {"Og": "-Og -fomit-frame-pointer -DSUPERCOP (the reference's Makefile)", "O2": "-O2 -fomit-frame-pointer -DSUPERCOP"}
    procs = {}
    for opt in ("Og", "O2"):
        if opt == "O2" and not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libgpqhe_O2.so")):
            continue
        qi, qo = mpctx.Queue(), mpctx.Queue()
        pr = mpctx.Process(target=_ref_worker, args=(qi, qo, opt), daemon=True)
        pr.start()
        procs[opt] = (pr, qi, qo)
    out = {}
    for opt, (pr, qi, qo) in procs.items():
        qo.get()                       # ready (context, keys, inputs)
        qi.put(1)                      # warm
    for opt, (pr, qi, qo) in procs.items():
        qo.get()
        qi.put(n_eval)
    for opt, (pr, qi, qo) in procs.items():
        dt = qo.get()
        qi.put(None)
        pr.join(timeout=10)
        out[opt] = {"value": n_eval / dt, "unit": UNIT, "cores": 1, "kind": "reference", "flags": flags[opt],
                    "sample": f"{n_eval} single-instance ctr_hempc ciphertext evaluations (src/hempc.c:253-266) through "
                              f"oracle/_ref/libgpqhe{'_O2' if opt == 'O2' else ''}.so (unmodified reference), {dt:.1f} s"}
    return out


def cpu_baseline_all_cores(per_proc=1):
    """The unmodified reference (-Og) on ALL host cores: P = os.cpu_count() processes (threads are impossible: process-global
    context), each evaluating `per_proc` instances after one warm-up evaluation; aggregate iterations/s, P stated."""
    from oracle import ref as R
    if not R.available():
        return None
    import multiprocessing as mp
    mpctx = mp.get_context("spawn")
    P = os.cpu_count() or 1
    qs = [(mpctx.Queue(), mpctx.Queue()) for _ in range(P)]
    procs = [mpctx.Process(target=_ref_worker, args=(qi, qo, "Og"), daemon=True) for qi, qo in qs]
    for pr in procs:
        pr.start()
    for _, qo in qs:
        qo.get()
    for qi, _ in qs:
        qi.put(1)
    for _, qo in qs:
        qo.get()
    t0 = time.perf_counter()
    for qi, _ in qs:
        qi.put(per_proc)
    for _, qo in qs:
        qo.get()
    dt = time.perf_counter() - t0
    for qi, _ in qs:
        qi.put(None)
    for pr in procs:
        pr.join(timeout=10)
    return {"value": P * per_proc / dt, "unit": UNIT, "cores": P, "kind": "reference", "flags": "-Og -fomit-frame-pointer -DSUPERCOP",
            "sample": f"{P} processes x {per_proc} ctr_hempc ciphertext evaluation(s), one process per host core, {dt:.1f} s"}


def _ref_worker(q_in, q_out, opt):
    ref, rk, Ka, Kb, cts = _ref_setup(opt)
    q_out.put("ready")
    while True:
        n = q_in.get()
        if n is None:
            break
        t0 = time.perf_counter()
        for _ in range(n):
            _ref_eval(ref, rk, Ka, Kb, cts)
        q_out.put(time.perf_counter() - t0)


def run_reference(args):
    """--impl reference: the unmodified reference on all host cores.  It is single-threaded with process-global state
    (SURVEY F9), so 'all cores' = one process per core, each evaluating its own instance."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref as R
    if not R.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref is not built on this box"}))
        return
    import multiprocessing as mp
    mpctx = mp.get_context("spawn")
    P = args.cpu_procs or os.cpu_count() or 1
    per_step = 1                                    # evaluations per process per step
    qs = [(mpctx.Queue(), mpctx.Queue()) for _ in range(P)]
    procs = [mpctx.Process(target=_ref_worker, args=(qi, qo, args.ref_opt), daemon=True) for qi, qo in qs]
    for p in procs:
        p.start()
    for _, qo in qs:
        qo.get()

    def step():
        t0 = time.perf_counter()
        for qi, _ in qs:
            qi.put(per_step)
        for _, qo in qs:
            qo.get()
        return time.perf_counter() - t0

    for _ in range(args.warmup):
        step()
    t = sum(step() for _ in range(args.steps))
    for qi, _ in qs:
        qi.put(None)
    for p in procs:
        p.join(timeout=10)
    value = P * per_step * args.steps / t
    total = args.instances
    flags = "-Og -fomit-frame-pointer -DSUPERCOP" if args.ref_opt == "Og" else "-O2 -fomit-frame-pointer -DSUPERCOP"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic (reference-encrypted random states, random gain blocks)",
        "config": {"workload": f"{total} independent cstr-hempc controller instances (BASELINE configs[4] on configs[0] parameters)",
                   "logn": PARAMS["logn"], "logq": PARAMS["logq"], "logDelta": PARAMS["logDelta"], "slots": PARAMS["slots"],
                   "instances": total},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": P, "kind": "reference", "flags": flags,
                         "sample": f"each step = {P} processes x {per_step} ctr_hempc ciphertext evaluation(s) of the {total}-instance "
                                   f"workload (bounded sample; the reference is single-threaded, one process per core)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--instances", type=int, default=4096, help="controller instances in the whole job")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ntt", action="store_true", help="skip the secondary blocks (NTT table, configs C / D, client-inclusive, plan rebuild)")
    ap.add_argument("--no-parity-spot", action="store_true")
    ap.add_argument("--no-closed-loop", action="store_true")
    ap.add_argument("--cpu-evals", type=int, default=3)
    ap.add_argument("--cpu-procs", type=int, default=0)
    ap.add_argument("--ref-opt", default="Og", choices=["Og", "O2"])
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
