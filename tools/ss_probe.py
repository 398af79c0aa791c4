#!/usr/bin/env python
"""Diagnostic: refine rounds and time of the self-synchronising decoder per image kind / format."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import image_codecs_b200 as ic
from tools import testenc as te

d = ic.BatchDecoder(device=0, profile=True)
for name, kw in [("420", dict(subsampling="420")), ("444", dict(subsampling="444")), ("gray", dict(colorspace=te.CS_GRAY)),
                 ("420q100opt", dict(subsampling="420", quality=100, optimize=True)), ("420q50", dict(subsampling="420", quality=50))]:
    for kind in (0, 1, 2):
        kw2 = dict(quality=90)
        kw2.update(kw)
        data = te.encode_many(4, 1920, 1080, kind=kind, base_seed=5, **kw2)
        d.prepare(data, 3)
        d.run(); d.sync(); d.run(); d.sync()
        tm = d.timings()
        print(name, "kind", kind, "bytes", len(data[0]), "scans", tm["selfsync_scans"], "rounds", tm["selfsync_rounds"],
              "entropy_ms %.2f" % tm["entropy_ms"])
