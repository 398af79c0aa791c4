#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` export: executed warp-instructions by SASS opcode and
the top stall-sample instructions.  Usage: ncu_sass_hist.py file.csv [top_n]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
iS, iE, iT, iSm = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
ops = collections.Counter(); tot = 0; samples = []
for r in rows[2:]:
    if len(r) <= iSm: continue
    try: e = int(r[iE]); s = int(r[iSm])
    except ValueError: continue
    toks = r[iS].split()
    op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
    op = op.split(".")[0]
    ops[op] += e; tot += e
    samples.append((s, e, r[iS].strip()))
print("total warp-instructions executed:", tot)
for op, e in ops.most_common(topn):
    print(f"  {op:10s} {e:12d} {100.0*e/tot:5.1f}%")
stot = sum(s for s, _, _ in samples)
print("top stall-sample instructions (of %d samples):" % stot)
for s, e, src in sorted(samples, reverse=True)[:topn]:
    print(f"  {s:6d} {100.0*s/max(stot,1):5.1f}%  exec={e:9d}  {src[:90]}")
