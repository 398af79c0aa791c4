#!/usr/bin/env python
"""Join an ncu report's per-SASS-instruction counters with the source lines of the built library:
   python tools/ncu_by_line.py REPORT.ncu-rep KERNEL_REGEX MANGLED_SUBSTRING [LIB.so] > profiles/....txt
Prints instructions / stall samples per source line, the opcode mix and the stall-reason totals (needs ncu, cuobjdump,
nvdisasm on PATH; runs on the CPU box)."""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

rep, kregex, mangled = sys.argv[1:4]
lib = sys.argv[4] if len(sys.argv) > 4 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                          "image_codecs_b200", "libimage_codecs_b200.so")
d = tempfile.mkdtemp()
subprocess.run(f"cd {d} && cuobjdump -xelf all {lib} > /dev/null 2>&1", shell=True)
cub = max([f for f in os.listdir(d) if f.endswith(".cubin")], key=lambda f: os.path.getsize(os.path.join(d, f)))
txt = subprocess.run(f"nvdisasm -g -c {d}/{cub}", shell=True, capture_output=True, text=True).stdout.split("\n")
i0 = [i for i, l in enumerate(txt) if l.startswith(".text.") and mangled in l and l.rstrip().endswith(":")][0]
addr2line, cur = {}, None
for l in txt[i0 + 1:]:
    if l.startswith("//--------------------- .text."):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        addr2line[int(m.group(1), 16)] = (cur, m.group(2))
# NCU_SKIP=n: skip the first n launches that match (e.g. the other instantiation of a template)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kregex}", "--launch-skip",
                      os.environ.get("NCU_SKIP", "0"), "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = [i for i, r in enumerate(rows) if "Address" in r and "Source" in r][0]
hdr = rows[h]
ia, ie, ism, ite = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
st = [(i, c[6:]) for i, c in enumerate(hdr) if c.startswith("stall_") and not c.endswith("(Not Issued)")]
sn = [(i, c[6:-13]) for i, c in enumerate(hdr) if c.startswith("stall_") and c.endswith("(Not Issued)")]
base = None
ins, tins, smp = collections.Counter(), collections.Counter(), collections.Counter()
ops, lineops = collections.Counter(), collections.defaultdict(collections.Counter)
stall, stall_ni, linestall = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
for r in rows[h + 1:]:
    if h and len(r) <= ie or r[0].startswith("Kernel Name"):
        break
    try:
        a = int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
    except ValueError:
        continue
    if base is None:
        base = a
    li, sass = addr2line.get(a - base, (None, ""))
    n = float(r[ie] or 0)
    ins[li] += n
    tins[li] += float(r[ite] or 0)
    smp[li] += float(r[ism] or 0)
    op = (sass.split()[1] if sass.startswith("@") else (sass.split()[0] if sass else "?")).split(".")[0]
    ops[op] += n
    lineops[li][op] += n
    for i, k in st:
        v = float(r[i] or 0)
        stall[k] += v
        linestall[li][k] += v
    for i, k in sn:
        stall_ni[k] += float(r[i] or 0)
ti, ts = sum(ins.values()), sum(smp.values())
print(f"# {rep}  kernel /{kregex}/  first launch in the report")
print(f"warp instructions executed {ti:.0f}   average active threads per instruction {sum(tins.values()) / max(ti, 1):.1f}")
print("\nopcode mix (% of warp instructions):")
print("  " + "  ".join(f"{k} {100 * v / ti:.1f}" for k, v in ops.most_common(24)))
print("\nwarp-state samples (all):")
print("  " + "  ".join(f"{k} {100 * v / max(sum(stall.values()), 1):.1f}%" for k, v in stall.most_common(10)))
print("cycles in which the scheduler issued nothing, by the state of the sampled warp:")
print("  " + "  ".join(f"{k} {100 * v / max(sum(stall_ni.values()), 1):.1f}%" for k, v in stall_ni.most_common(8)))
print("\nper source line (instructions %, stall samples %, active threads, top opcodes, top stall reasons):")
for li, n in ins.most_common(45):
    print("  %5.2f%%  %5.2f%%  %4.1f  %-36s %s  %s" % (100 * n / ti, 100 * smp[li] / max(ts, 1), tins[li] / max(n, 1), str(li),
          " ".join(f"{k}:{100 * v / ti:.1f}" for k, v in lineops[li].most_common(3)),
          " ".join(f"{k}:{int(v)}" for k, v in linestall[li].most_common(3))))
