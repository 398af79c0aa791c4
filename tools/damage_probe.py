"""Exploration: decode truncated / bit-flipped streams on the GPU and compare with the oracle (accept/reject + pixels)."""
import sys
import numpy as np
sys.path.insert(0, '.')
import image_codecs_b200 as ic
from oracle import oracle_lib
from tests import streams as S

o = oracle_lib.load()
rng = np.random.default_rng(11)
fmts = ["420", "420_rst16", "444", "gray", "420_prog", "420_prog_rst5", "444_prog_ss", "gray_prog", "420_nonint", "420_opt", "422"]
cases = []
for f in fmts:
    for (w, h) in ((233, 177), (64, 48)):
        raw = S.stream(f, w, h, 1)
        n = len(raw)
        for frac in (0.5, 0.8, 0.97):
            cases.append((f"{f} {w}x{h} cut{frac}", bytes(raw[: int(n * frac)])))
        for t in range(6):
            b = bytearray(raw)
            pos = int(rng.integers(n // 3, n - 2))
            b[pos] ^= 1 << int(rng.integers(0, 8))
            cases.append((f"{f} {w}x{h} flip{t}@{pos}", bytes(b)))
for mode in (0, 1, 2):
    d = ic.BatchDecoder(device=0, entropy_mode=mode)
    outs = d.decode([c[1] for c in cases], 3)
    bad = 0
    for (tag, data), got in zip(cases, outs):
        want, *_ = o.load_from_memory(data, 3)
        if (want is None) != (got is None):
            bad += 1
            print(f"mode{mode} {tag}: oracle {'fails' if want is None else 'decodes'} ({o.failure_reason() if want is None else ''}), gpu {'fails' if got is None else 'decodes'}")
        elif want is not None and not np.array_equal(want, got):
            bad += 1
            dd = np.argwhere(want != got)
            print(f"mode{mode} {tag}: {len(dd)} bytes differ, first {dd[0]}")
    print(f"mode {mode}: {bad} of {len(cases)} damaged streams differ")
    d.close()
