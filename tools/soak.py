#!/usr/bin/env python
"""Soak: create / use / destroy batch objects in a loop (pinned and pageable buffers, per-image loads, a failing file in
every batch) and watch device memory, host RSS and the thread count for growth.   python tools/soak.py [seconds]"""
import ctypes as C
import os
import sys
import threading
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import image_codecs_b200 as ic
from image_codecs_b200.api import PinnedBuffer, lib
from tools import testenc as te

secs = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
streams = te.encode_many(6, 1920, 1080, kind=-1, base_seed=9, quality=90, subsampling="420", restart_interval=16)
streams += te.encode_many(3, 1280, 720, kind=-1, base_seed=10, quality=90, subsampling="444")
streams += te.encode_many(2, 800, 600, kind=-1, base_seed=11, quality=85, subsampling="420", progressive=1)
streams.append(streams[0][: len(streams[0]) // 2] + b"\xff\xd9")  # damaged: decodes (zeros) or fails, either way alone
streams.append(b"not a jpeg")


def rss_mb():
    for line in open("/proc/self/status"):
        if line.startswith("VmRSS"):
            return int(line.split()[1]) / 1024.0
    return 0.0


def dev_free_mb():
    f, t = C.c_size_t(), C.c_size_t()
    lib().jcu_mem_info(0, C.byref(f), C.byref(t))
    return f.value / 2 ** 20


t_end = time.time() + secs
it = 0
first = None
while time.time() < t_end:
    d = ic.BatchDecoder(device=0, profile=False, chunk_images=16 if it % 2 else 0)
    outs = d.decode(streams * (4 if it % 3 else 12), 3 if it % 2 else 0)  # numpy (pageable) outputs
    assert sum(o is None for o in outs) in (4, 8, 12, 24), [o is None for o in outs][:14]
    d.close()
    got = ic.stbi_load_from_memory(streams[it % 9], 3)[0]
    assert got is not None
    it += 1
    if it in (5, 10) or it % 25 == 0:
        snap = (rss_mb(), dev_free_mb(), threading.active_count(), len(os.listdir("/proc/self/task")))
        if it == 10:
            first = snap
        print("iteration %d: host RSS %.0f MB, device free %.0f MB, python threads %d, OS threads %d" % ((it,) + snap), flush=True)
last = (rss_mb(), dev_free_mb(), threading.active_count(), len(os.listdir("/proc/self/task")))
print("done: %d iterations; since iteration 10: RSS %+.0f MB, device free %+.0f MB, OS threads %+d" %
      (it, last[0] - first[0], last[1] - first[1], last[3] - first[3]))
