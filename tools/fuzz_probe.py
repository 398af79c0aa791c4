"""Exploration: random damage (bit flips, byte overwrites, truncation, dropped / duplicated spans) on every test
format; GPU decode vs the oracle (accept/reject, failure string, pixels).
Usage: python tools/fuzz_probe.py [seed] [per_format] [modes]"""
import sys
import collections
import numpy as np

sys.path.insert(0, ".")
import image_codecs_b200 as ic
from oracle import oracle_lib
from tests import streams as S

def gen_cases(seed, per):
    rng = np.random.default_rng(seed)
    cases = []
    for f, _ in S.FORMATS:
        for (w, h) in ((97, 61), (64, 48)):
            raw = S.stream(f, w, h, int(rng.integers(0, 3)))
            n = len(raw)
            for t in range(per):
                b = bytearray(raw)
                kind = int(rng.integers(0, 5))
                pos = int(rng.integers(2, n - 2))
                if kind == 0:
                    b[pos] ^= 1 << int(rng.integers(0, 8))
                elif kind == 1:
                    b[pos] = int(rng.integers(0, 256))
                elif kind == 2:
                    b = b[:pos]
                elif kind == 3:
                    del b[pos : pos + int(rng.integers(1, 9))]
                else:
                    span = int(rng.integers(1, 9))
                    b[pos:pos] = b[pos : pos + span]
                sos = raw.index(b"\xff\xda")
                markers = lambda d: [d[i + 1] for i in range(len(d) - 1) if d[i] == 0xFF and d[i + 1] not in (0, 0xFF)]
                region = "header" if pos < sos + 14 else ("markers-changed" if markers(b) != markers(raw) else "entropy")
                cases.append((f"{f} {w}x{h} k{kind}@{pos} [{region}]", bytes(b)))
    return cases


if __name__ == "__main__":
    seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    per = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    modes = [int(m) for m in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0, 1, 2]
    o = oracle_lib.load()
    cases = gen_cases(seed, per)
    for mode in modes:
        d = ic.BatchDecoder(device=0, entropy_mode=mode)
        cats = collections.Counter()
        shown = 0
        for c0 in range(0, len(cases), 64):
            chunk = cases[c0 : c0 + 64]
            outs = d.decode([c[1] for c in chunk], 3)
            for i, ((tag, data), got) in enumerate(zip(chunk, outs)):
                want, *_ = o.load_from_memory(data, 3)
                msg = None
                if (want is None) != (got is None):
                    why = o.failure_reason() if want is None else "gpu fails"
                    cats[f"{tag.split('[')[1][:-1]}: oracle {'fails' if want is None else 'decodes'} / {why}"] += 1
                    msg = f"oracle {'fails' if want is None else 'decodes'}, gpu {'fails' if got is None else 'decodes'} ({why})"
                elif want is not None and not np.array_equal(want, got):
                    cats[f"{tag.split('[')[1][:-1]}: pixels"] += 1
                    dd = np.argwhere(want != got)
                    msg = f"{len(dd)} bytes differ, first {dd[0]}"
                else:
                    cats["same"] += 1
                if msg and shown < 30 and ("[entropy]" in tag or "fails" in msg):
                    shown += 1
                    print(f"mode{mode} {tag}: {msg}")
        print(f"mode {mode}: {dict(cats)}")
        d.close()
