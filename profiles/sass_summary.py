#!/usr/bin/env python
"""Summarise an `ncu --page source --csv --print-source sass` export (gz ok): stall reasons and opcode mix per kernel."""
import collections
import csv
import gzip
import sys

csv.field_size_limit(10 ** 9)


def main(path, top=12):
    f = gzip.open(path, "rt") if path.endswith(".gz") else open(path)
    hdr, kern = None, None
    stalls = collections.defaultdict(lambda: collections.Counter())
    ops = collections.defaultdict(lambda: collections.Counter())
    samples_by_op = collections.defaultdict(lambda: collections.Counter())
    hot = collections.defaultdict(list)
    for row in csv.reader(f):
        if not row:
            continue
        if row[0] == "Kernel Name":
            kern = row[1][:60]
            continue
        if row[0] == "Address":
            hdr = row
            continue
        if hdr is None or len(row) != len(hdr):
            continue
        d = dict(zip(hdr, row))
        src = d["Source"].strip()
        op = src.split()[0] if not src.startswith("@") else src.split()[1]
        for pre in ("IMAD.WIDE", "IMAD.MOV", "IMAD.IADD", "IMAD.X", "IMAD.SHL", "IMAD.HI"):      # the IMAD family shares the fmaheavy pipe
            if op.startswith(pre):
                op = pre
                break
        else:
            op = op.split(".")[0]
        n = int(d["Instructions Executed"] or 0)
        ops[kern][op] += n
        s = int(d["# Samples"] or 0)
        samples_by_op[kern][op] += s
        hot[kern].append((s, src))
        for k, v in d.items():
            if k.startswith("stall_") and "Not Issued" not in k:
                stalls[kern][k] += int(v or 0)
    for k in ops:
        tot = sum(ops[k].values())
        print(f"== {k}: {tot} warp instructions")
        print("  opcode mix :", ", ".join(f"{o} {100 * n / tot:.1f}%" for o, n in ops[k].most_common(top)))
        ts = sum(stalls[k].values()) or 1
        print("  stalls     :", ", ".join(f"{o[6:]} {100 * n / ts:.1f}%" for o, n in stalls[k].most_common(8)))
        tsa = sum(samples_by_op[k].values()) or 1
        print("  samples@op :", ", ".join(f"{o} {100 * n / tsa:.1f}%" for o, n in samples_by_op[k].most_common(top)))
        for s, src in sorted(hot[k], reverse=True)[:8]:
            print(f"     {100 * s / tsa:5.2f}%  {src[:90]}")


if __name__ == "__main__":
    for p in sys.argv[1:]:
        main(p)
