#!/usr/bin/env python
"""Print the interesting fields of a bench.py JSON line read from stdin (helper for gpurun one-liners)."""
import json
import sys

import os

tag = sys.argv[1] if len(sys.argv) > 1 else ""
# the argument may be a file name (never block on an empty stdin inside a gpurun call)
text = open(tag).read() if os.path.isfile(tag) else ("" if sys.stdin.isatty() else sys.stdin.read())
d = json.loads(text.strip().splitlines()[-1])
e = d.get("e2e") or {}
print(tag, "value=%.0f" % d["value"], "e2e=%.0f (%.1f ms)" % (e.get("value", 0), e.get("ms_per_step", 0)),
      "ent=%.2fms fused=%.2fms frac=%.3f" % (d["stages"]["entropy_ms"], d["stages"]["idct_color_ms"], d["roofline"]["frac"]),
      "cpu=", (d.get("cpu_baseline") or {}).get("value"))
