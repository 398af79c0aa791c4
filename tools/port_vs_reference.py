#!/usr/bin/env python
"""CPU only: the plain-C restatement (oracle/jpeg_port.c) against the compiled reference (oracle/_ref) on the damaged
files of tools/fuzz_probe.py -- verdict, failure string and pixels.   python tools/port_vs_reference.py SEED [PER_FORMAT]
(the reference reads uninitialised memory on some of these files and may crash: run one seed per process)"""
import collections
import sys

import numpy as np

sys.path.insert(0, ".")
from oracle import oracle_lib
from tools import fuzz_probe as fp

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
per = int(sys.argv[2]) if len(sys.argv) > 2 else 40
ref, port = oracle_lib.load("reference"), oracle_lib.load("port")
cats = collections.Counter()
for tag, data in fp.gen_cases(seed, per):
    a = ref.load_from_memory(data, 3)[0]
    ra = ref.failure_reason() if a is None else None
    b = port.load_from_memory(data, 3)[0]
    rb = port.failure_reason() if b is None else None
    region = tag.split("[")[1][:-1]
    if (a is None) != (b is None):
        cats[f"{region}: verdict"] += 1
        print(f"{tag}: reference {'fails: ' + str(ra) if a is None else 'decodes'}, port {'fails: ' + str(rb) if b is None else 'decodes'}")
    elif a is None:
        cats["same" if ra == rb else f"{region}: reason"] += 1
        if ra != rb:
            print(f"{tag}: reference '{ra}', port '{rb}'")
    elif not np.array_equal(a, b):
        cats[f"{region}: pixels"] += 1
    else:
        cats["same"] += 1
print(f"seed {seed}: {dict(cats)}")
