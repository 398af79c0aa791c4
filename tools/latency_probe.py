#!/usr/bin/env python
"""Diagnostic: where the time of ONE per-image load goes (config C1: 512x512 4:2:0 q90 jpeg_write.c stream)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import image_codecs_b200 as ic
from tools import testenc as te

w = int(sys.argv[1]) if len(sys.argv) > 1 else 512
h = int(sys.argv[2]) if len(sys.argv) > 2 else 512
stream = ic.write_jpg_to_bytes(te.synth_rgb(w, h, 0, 12345), 90)
rst = te.encode(te.synth_rgb(w, h, 0, 12345), quality=90, subsampling="420", restart_interval=4)
for name, s in (("jpeg_write (no RSTn, self-sync)", stream), ("restart interval 4", rst)):
    d = ic.BatchDecoder(device=0, profile=True, host_threads=1)
    out = np.empty((h, w, 3), dtype=np.uint8)
    t = {k: [] for k in ("prepare", "run", "sync", "fetch", "total")}
    for it in range(60):
        t0 = time.perf_counter()
        d.prepare([s], 3)
        t1 = time.perf_counter()
        d.run()
        t2 = time.perf_counter()
        d.sync()
        t3 = time.perf_counter()
        a = d.fetch(0)
        t4 = time.perf_counter()
        if it >= 10:
            for k, v in zip(t, (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t4 - t0)):
                t[k].append(1e3 * v)
    tm = d.timings()
    print(name, len(s), "bytes:", {k: round(float(np.median(v)), 3) for k, v in t.items()},
          {k: round(v, 3) for k, v in tm.items() if k.endswith("_ms") or k in ("launches", "selfsync_rounds")})
    lat = []
    for it in range(100):
        t0 = time.perf_counter()
        ic.stbi_load_from_memory(s, 3)
        lat.append(1e3 * (time.perf_counter() - t0))
    print("   stbi_load_from_memory p50 %.3f ms" % float(np.median(lat[10:])))
    d.close()
