#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` dump by SASS opcode: executed warp instructions, share, samples.
usage: ncu -i X.ncu-rep --page source --csv > src.csv; python tools/ncu_opcode_hist.py src.csv [kernel-substring]"""
import csv
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    want = sys.argv[2] if len(sys.argv) > 2 else None
    kernel = None
    hdr = None
    agg = {}
    for row in csv.reader(open(path)):
        if not row:
            continue
        if row[0] == "Kernel Name":
            kernel = row[1]
            hdr = None
            agg[kernel] = defaultdict(lambda: [0, 0, 0])
            continue
        if row[0] == "Address":
            hdr = {k: i for i, k in enumerate(row)}
            continue
        if hdr is None or kernel is None:
            continue
        src = row[hdr["Source"]].strip()
        toks = src.split()
        if not toks:
            continue
        op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
        op = op.rstrip(";")
        base = op.split(".")[0]
        a = agg[kernel][base]
        a[0] += int(row[hdr["Instructions Executed"]] or 0)
        a[1] += int(row[hdr["Thread Instructions Executed"]] or 0)
        a[2] += int(row[hdr["# Samples"]] or 0)
    for k, d in agg.items():
        if want and want not in k:
            continue
        tot = sum(v[0] for v in d.values()) or 1
        ts = sum(v[2] for v in d.values()) or 1
        print(f"== {k[:100]}\n   total warp-inst {tot}")
        for op, v in sorted(d.items(), key=lambda kv: -kv[1][0])[:40]:
            print(f"   {op:12s} {v[0]:12d} {100.0 * v[0] / tot:6.2f}%  lanes/inst {v[1] / max(v[0], 1):5.1f}  samples {100.0 * v[2] / ts:5.1f}%")


if __name__ == "__main__":
    main()
