#!/usr/bin/env python
"""Condense `ncu --page raw --csv` exports into the per-kernel table the design notes cite, and write
profiles/traffic.json (DRAM bytes per launch of each kernel, read by bench.py for `roofline.traffic`).

    python profiles/summarize.py gpurun_out/r01/prof_hempc.raw.csv gpurun_out/r01/prof_ntt16.raw.csv --batch 4096 > profiles/r01/ncu_summary.md
"""
import argparse
import csv
import json
import os
import re
import sys

ALT = {   # metric name variants across ncu versions
    "dram__throughput.avg.pct_of_peak_sustained_elapsed": ["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"],
    "smsp__issue_active.avg.pct_of_peak_sustained_elapsed": ["smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_issued.avg.pct_of_peak_sustained_active",
                                                             "smsp__inst_issued.avg.pct_of_peak_sustained_active"],
}

COLS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram rd"),
    ("dram__bytes_write.sum", "dram wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
    ("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "fmaheavy %"),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed", "alu %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_elapsed", "lsu %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_elapsed", "issue %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
]


def to_bytes(v, unit):
    f = float(v.replace(",", ""))
    u = unit.lower()
    mult = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}.get(u, 1)
    return f * mult


def to_ms(v, unit):
    f = float(v.replace(",", ""))
    return f * {"ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3}.get(unit.lower().replace("usecond", "us").replace("msecond", "ms").replace("nsecond", "ns").replace("second", "s"), 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("files", nargs="+")
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--traffic", default=os.path.join(os.path.dirname(os.path.abspath(__file__)), "traffic.json"))
    a = ap.parse_args()
    traffic = {}
    print("| kernel | " + " | ".join(c[1] for c in COLS) + " | top stall (warps per issue) |")
    print("|---|" + "---|" * (len(COLS) + 1))
    for path in a.files:
        rows = list(csv.reader(open(path)))
        hdr, units = rows[0], rows[1]
        idx = {h: i for i, h in enumerate(hdr)}
        seen = set()
        for r in rows[2:]:
            name = re.sub(r"^void\s+", "", r[idx["Kernel Name"]])
            name = re.sub(r"\(.*$", "", name).replace("hbdev::", "").replace("hb32::", "")
            if name in seen:
                continue
            seen.add(name)
            cells = []
            rd = wr = 0.0
            heavy = None
            for key, label in COLS:
                if key not in idx:
                    key = next((alt for alt in ALT.get(key, []) if alt in idx), key)
                if key not in idx:
                    cells.append("-")
                    continue
                v, u = r[idx[key]], units[idx[key]]
                try:
                    if label == "time":
                        cells.append(f"{to_ms(v, u):.3f} ms")
                    elif label.startswith("dram r") or label.startswith("dram w"):
                        b = to_bytes(v, u)
                        if label == "dram rd":
                            rd = b
                        else:
                            wr = b
                        cells.append(f"{b / 1e6:.0f} MB")
                    elif label.endswith("%"):
                        cells.append(f"{float(v.replace(',', '')):.1f}")
                        if label == "fmaheavy %":
                            heavy = float(v.replace(",", ""))
                    else:
                        cells.append(v)
                except ValueError:
                    cells.append(v)
            # first stall reason: the largest smsp__average_warps_issue_stalled_*_per_issue_active column
            stall = None
            for hname, i in idx.items():
                m = re.match(r"smsp__average_warps?_(?:latency_)?issue_stalled_(.+?)_per_(?:issue_active|warp_active).*", hname)
                if m and not m.group(1).endswith("not_issued"):
                    try:
                        val = float(r[i].replace(",", ""))
                    except ValueError:
                        continue
                    if stall is None or val > stall[1]:
                        stall = (m.group(1), val)
            cells.append(f"{stall[0]} ({stall[1]:.2f})" if stall else "-")
            print(f"| `{name}` | " + " | ".join(cells) + " |")
            base = name.split("<")[0]
            traffic.setdefault(base, {})[str(a.batch)] = rd + wr
            if heavy is not None:
                traffic[base]["fmaheavy_pct"] = heavy      # the pipe IMAD / IMAD.WIDE issue to: the binding unit of this path
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    merged = {"sources_digest": bench.sources_digest(), "kernels": {}}
    if os.path.exists(a.traffic):                      # keep captures of other kernels made from the same sources
        try:
            old = json.load(open(a.traffic))
            if old.get("sources_digest") == merged["sources_digest"]:
                merged["kernels"].update(old.get("kernels", {}))
        except Exception:
            pass
    for kname, rec in traffic.items():
        merged["kernels"].setdefault(kname, {}).update(rec)
    json.dump(merged, open(a.traffic, "w"), indent=1, sort_keys=True)
    print(f"\n(traffic per launch written to {os.path.relpath(a.traffic)})", file=sys.stderr)


if __name__ == "__main__":
    main()
