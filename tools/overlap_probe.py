#!/usr/bin/env python
"""Diagnostic: does the device-resident decode gain from running sub-batches concurrently on separate streams?
K decoder objects (each its own stream) take total/K images each; run() on all, then sync() on all; wall clock
around the lot (inputs resident, as bench.py's `value`).  tools/overlap_probe.py TOTAL K [K...]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import image_codecs_b200 as ic
from tools import testenc as te

total = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
ks = [int(a) for a in sys.argv[2:]] or [1, 2, 4]
distinct = te.encode_many(64, 1920, 1080, kind=-1, base_seed=1000, quality=90, subsampling="420", restart_interval=16)
for k in ks:
    per = total // k
    decs = []
    for j in range(k):
        d = ic.BatchDecoder(device=0, profile=False)
        d.prepare([distinct[(j * per + i) % len(distinct)] for i in range(per)], 3)
        decs.append(d)
    best = 1e9
    for rep in range(6):
        for d in decs:
            d.sync()
        t0 = time.perf_counter()
        for d in decs:
            d.run()
        for d in decs:
            d.sync()
        ms = 1e3 * (time.perf_counter() - t0)
        if rep >= 2:
            best = min(best, ms)
    print("sub-batches %d x %d images: %.2f ms  %.0f MP/s" % (k, per, best, total * 1920 * 1080 / 1e3 / best), flush=True)
    for d in decs:
        d.close()
