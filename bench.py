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
q_l (what an RLWE ciphertext looks like), uniform random key limbs, random gain matrices with the hempc block
structure.  The arithmetic is identical for any input values (no data-dependent control flow in the kernels).
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


def synth_keys(eng, rng):
    primes = eng.primes()
    i = eng.info
    keys = np.empty((eng.slots, 2, i.dimevk, eng.n), dtype=np.uint64)
    for d in range(i.dimevk):
        keys[:, :, d, :] = rng.integers(0, int(primes[d]), size=(eng.slots, 2, eng.n), dtype=np.uint64)
    return keys


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


# algorithmic bytes per launch, config A, batch B instances per launch
def algorithmic_bytes(kernel: str, B: int, n: int, W: int, D_ks: int, D_pt: int, n1: int = 4, n2: int = 4):
    """SURVEY.md 8(d) per-unit figures x the units one launch processes.  The grouped he_gemv schedule puts the n1 baby
    rotations (resp. the n2 giant steps) of one matvec into ONE launch, so a launch is n1 (n2) key switches / products."""
    ct_poly = n * W * 8
    hat = B * 2 * D_pt * n * 8
    keyswitch = B * (2 * ct_poly + 2 * ct_poly) + 2 * D_ks * n * 8      # read d0,d1 + write c0,c1 + key shared by the batch
    table = {
        "k_ks_limb": n2 * keyswitch,                  # the n2 giant-step key switches
        "k_ks_baby": n1 * keyswitch,                  # the n1 baby rotations (one shared forward transform)
        "k_ks_finish": n1 * keyswitch,
        "k_ks_finish_fast": n1 * keyswitch,           # second half of the same key switches (n1 == n2 at config A)
        # one ct x pt: ct in + ct out + plaintext (shared)
        "k_mulpt_limb": B * (2 * ct_poly + 2 * ct_poly) + D_pt * n * 8,
        "k_recon_smod": n2 * (B * 2 * ct_poly + hat),
        # memoised gemv: rotated ct in -> NTT limbs out; n1 limb sets in -> n2 limb sets out
        "k_poly_ntt": n1 * (B * 2 * ct_poly + hat),
        "k_mulacc_intt": n1 * hat + n2 * hat,
        "k_ct_addsub": B * 3 * 2 * ct_poly,
        "k_ct_sum": B * (n2 + 1) * 2 * ct_poly,
        "k_ct_rescale": B * 2 * 2 * ct_poly,
    }
    return table.get(kernel)


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
    keys = eng.keys(synth_keys(eng, np.random.default_rng(7)))
    Ka, Kb = synth_gains(np.random.default_rng(9), eng.slots)

    names = ("xhat", "xr", "uhat", "ur")
    host, hptr = {}, {}
    for nm in names:
        host[nm], hptr[nm] = pinned_array(eng, (per, 2, W, n))
        synth_ct_into(host[nm], rng, k)
    out_host, out_ptr = pinned_array(eng, (per, 2, Wout, n))
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
    need = lib.hb_prof_report(ctx, None, 0)
    buf = C.create_string_buffer(need + 1)
    lib.hb_prof_report(ctx, buf, need + 1)
    prof = json.loads(buf.value.decode())
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = total * args.steps / (ms / 1e3)

    # ---- end to end through the C ABI with host buffers: H2D of the 4 input batches + evaluate + D2H of the result ----
    def e2e_step():
        lib.hb_ctr_hempc_eval_host(ctx, per, l, Ka.ctypes.data_as(C.c_void_p), Kb.ctypes.data_as(C.c_void_p),
                                   C.c_void_p(hptr["xhat"]), C.c_void_p(hptr["xr"]), C.c_void_p(hptr["uhat"]),
                                   C.c_void_p(hptr["ur"]), keys.arr, C.c_void_p(out_ptr))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = total * args.steps / e2e_s
    h2d = 4 * per * 2 * W * n * 8
    d2h = per * 2 * Wout * n * 8

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel, timed live with CUDA events in the region above ----
    peak, peak_src = load_peaks()
    D_ks = lib.hb_dim_keyswitch(ctx, l)
    D_pt = lib.hb_dim_mulpt(ctx, l, Delta)
    tot_ms = sum(v["ms"] for v in prof.values()) or 1.0
    dom = max(prof, key=lambda kname: prof[kname]["ms"])
    avg_ms = prof[dom]["ms"] / prof[dom]["launches"]
    abytes = algorithmic_bytes(dom, per, n, W, D_ks, D_pt)
    achieved = abytes / (avg_ms / 1e3) / 1e9 if abytes else None
    traffic, heavy = None, None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            rec = json.load(open(tpath)).get(dom, {})
            traffic, heavy = rec.get(str(per)), rec.get("fmaheavy_pct")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": abytes, "avg_launch_ms": avg_ms,
                "share_of_step": prof[dom]["ms"] / tot_ms,
                "kernels": {kname: {"launches": v["launches"], "ms": round(v["ms"], 3), "share": round(v["ms"] / tot_ms, 4)}
                            for kname, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])},
                "int_pipe": {"unit": "sm__pipe_fmaheavy_cycles_active % of peak (the pipe IMAD/IMAD.WIDE issue to)", "pct": heavy,
                             "source": "profiles/r01/ncu_summary.md (ncu --set full capture of this kernel at this batch)"},
                "note": "integer-pipe bound path (59-bit modular butterflies); HBM fraction is reported as the contract asks"}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_single(n_eval=args.cpu_evals)
    ntt = None
    if not args.no_ntt:
        try:
            ntt = ntt_microbench(local, peak)
        except Exception as ex:  # pragma: no cover
            ntt = {"error": str(ex)}

    secondary = {}
    if not args.no_ntt:
        for nm, fn in (("config_C", config_c_bench), ("config_D", config_d_bench)):
            try:
                secondary[nm] = fn(local)
            except Exception as ex:  # pragma: no cover
                secondary[nm] = {"error": str(ex)}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic (uniform random ciphertext words and key limbs, random 8x3 / 8x2 gain blocks)",
        "config": {"workload": f"{total} independent cstr-hempc controller instances (BASELINE configs[4] on configs[0] parameters)",
                   "logn": PARAMS["logn"], "logq": PARAMS["logq"], "logDelta": PARAMS["logDelta"], "slots": PARAMS["slots"],
                   "instances": total, "instances_per_gpu": per, "parallelism": f"instances sharded x{world}, keys replicated",
                   "l2": f"inputs larger than L2 ({4 * per * 2 * W * n * 8 >> 20} MiB of ciphertext per GPU per step)"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "api": "hb_ctr_hempc_eval_host (pinned host buffers in, pinned host buffer out; 512-instance chunks pipelined over upload / evaluate / download streams)"},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": roofline,
        "cpu_baseline": cpu,
        "ntt": ntt,
        **secondary,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------------
# secondary metric: N=2^16 batched forward NTT limbs/s (BASELINE configs[1])
# ------------------------------------------------------------------------------------------------------
def ntt_microbench(device: int, peak: float, logn: int = 16, nlimbs: int = 1024, D: int = 8, reps: int = 5):
    from hectr_b200 import Engine
    eng = Engine(logn, 109, 16, 50, device=device)
    lib, ctx = eng.lib, eng.ctx
    n = eng.n
    primes = eng.primes()
    rng = np.random.default_rng(3)
    host = np.empty((nlimbs, n), dtype=np.uint64)
    for j in range(nlimbs):
        host[j] = rng.integers(0, int(primes[j % D]), size=n, dtype=np.uint64)
    d_in = eng.to_device(host)
    d_out = eng.dev_alloc(host.nbytes)
    res = {}
    for name, fn in (("forward", lambda: lib.hb_ntt_forward(ctx, d_in, d_out, nlimbs, 0, D, 0)),
                     ("inverse", lambda: lib.hb_ntt_inverse(ctx, d_in, d_out, nlimbs, 0, D))):
        for _ in range(3):
            fn()
        eng.sync()
        lib.hb_timer_start(ctx)
        for _ in range(reps):
            fn()
        ms = float(lib.hb_timer_stop_ms(ctx)) / reps
        gbs = 16.0 * n * nlimbs / (ms / 1e3) / 1e9
        res[name] = {"limbs_per_s": nlimbs / (ms / 1e3), "ms_per_launch": ms, "achieved_gbs": gbs, "frac_of_hbm": gbs / peak}
    eng.dev_free(d_in)
    eng.dev_free(d_out)
    eng.close()
    return {"metric": "N=2^16 NTT limbs/s", "logn": logn, "limbs_per_launch": nlimbs, "primes": D,
            "algorithmic_bytes_per_limb": 16 * n, "l2": f"{host.nbytes >> 20} MiB in + out per launch (> L2)", **res}


# ------------------------------------------------------------------------------------------------------
# secondary: BASELINE configs[2] (ct x ct + relinearise + rescale, rotation at N=2^15, q=2^881) and configs[3]
# (rotation-sum matvec at 32/64/128 slots).  Parity for both is in tests/test_gpu_parity.py; these are throughput only.
# ------------------------------------------------------------------------------------------------------
def _rand_ct(rng, batch, W, n, k):
    buf = np.empty((batch, 2, W, n), dtype=np.uint64)
    synth_ct_into(buf, rng, k)
    return buf


def _rand_keys(eng, rng, nkeys):
    primes = eng.primes()
    keys = np.empty((nkeys, 2, eng.info.dimevk, eng.n), dtype=np.uint64)
    for d in range(eng.info.dimevk):
        keys[:, :, d, :] = rng.integers(0, int(primes[d]), size=(nkeys, 2, eng.n), dtype=np.uint64)
    return keys


def config_c_bench(device: int, batch: int = 8, reps: int = 3):
    from hectr_b200 import Engine
    eng = Engine(15, 881, 16, 50, device=device)
    lib, ctx = eng.lib, eng.ctx
    rng = np.random.default_rng(5)
    l = eng.L
    W, k = eng.words(l), eng.logq(l)
    keys = eng.keys(_rand_keys(eng, rng, 2))
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
        keys = eng.keys(_rand_keys(eng, rng, slots))
        A = rng.standard_normal((slots, slots)) + 1j * rng.standard_normal((slots, slots))
        x = eng.ct(_rand_ct(rng, batch, eng.words(l), eng.n, eng.logq(l)), l)
        out = eng.ct(batch=batch)
        eng.he_gemv(A, x, keys, out=out)          # builds the plan (host encode of the diagonals)
        eng.sync()
        lib.hb_timer_start(ctx)
        for _ in range(reps):
            eng.he_gemv(A, x, keys, out=out)
        ms = float(lib.hb_timer_stop_ms(ctx)) / reps
        res[f"slots{slots}"] = {"ms_per_batch": ms, "gemv_per_s": batch / (ms / 1e3)}
        keys.free()
        eng.close()
    return res


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


def cpu_baseline_single(n_eval=3):
    from oracle import ref as R
    if not R.available():
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    ref, rk, Ka, Kb, cts = _ref_setup()
    _ref_eval(ref, rk, Ka, Kb, cts)   # warm
    t0 = time.perf_counter()
    for _ in range(n_eval):
        _ref_eval(ref, rk, Ka, Kb, cts)
    dt = time.perf_counter() - t0
    return {"value": n_eval / dt, "unit": UNIT, "cores": 1, "kind": "reference",
            "sample": f"{n_eval} single-instance ctr_hempc ciphertext evaluations through oracle/_ref/libgpqhe.so "
                      f"(unmodified reference, -Og as its Makefile builds it), {dt:.1f} s",
            "flags": "-Og -fomit-frame-pointer -DSUPERCOP"}


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
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--instances", type=int, default=4096, help="controller instances in the whole job")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ntt", action="store_true")
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
