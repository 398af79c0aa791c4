#!/usr/bin/env python
"""Diagnostic: host-buffer batched decode (stbi_jpeg_decode_batch) wall time vs accumulated stage timings."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import image_codecs_b200 as ic
from image_codecs_b200.api import PinnedBuffer
from tools import testenc as te

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 128
distinct = te.encode_many(32, 1920, 1080, kind=-1, base_seed=1000, quality=90, subsampling="420", restart_interval=16)
streams = [distinct[i % len(distinct)] for i in range(n)]
in_bytes = sum(len(s) for s in streams)
pin_in = PinnedBuffer(in_bytes + 64)
bufs = (C.c_void_p * n)()
lens = (C.c_int * n)()
off = 0
for i, s in enumerate(streams):
    pin_in.array[off:off + len(s)] = np.frombuffer(s, dtype=np.uint8)
    bufs[i] = pin_in.ptr + off
    lens[i] = len(s)
    off += len(s)
per = 1920 * 1080 * 3
pin_out = PinnedBuffer(per * n)
outs = (C.c_void_p * n)()
caps = (C.c_size_t * n)()
for i in range(n):
    outs[i] = pin_out.ptr + i * per
    caps[i] = per
st = (C.c_int * n)()
for prof in ((False,) if os.environ.get('QUICK') else (True, False)):
    d = ic.BatchDecoder(device=0, profile=prof, chunk_images=chunk, host_threads=int(os.environ.get('HT', '0')),
                        trace=int(os.environ.get('TRACE', '0')), pipeline_depth=int(os.environ.get('DEPTH', '0')), no_download_merge=int(os.environ.get('DLMODE', '0')))
    d.decode_into(bufs, lens, n, 3, outs, caps, None, None, None, st)
    t0 = time.perf_counter()
    rc = d.decode_into(bufs, lens, n, 3, outs, caps, None, None, None, st)
    wall = 1e3 * (time.perf_counter() - t0)
    print("profile", prof, "ok", rc, "wall %.1f ms" % wall, "MP/s %.0f" % (n * 1920 * 1080 / 1e3 / wall),
          {k: round(v, 1) for k, v in d.timings().items() if k.endswith("_ms")}, "D2H GB/s if alone %.1f" % (per * n / 1e6 / wall))
    d.close()
