import sys
import numpy as np
sys.path.insert(0, '.')
import image_codecs_b200 as ic
from oracle import oracle_lib
from tests import streams as S
o = oracle_lib.load()
cases = [("gray_prog", 64, 48, 356), ("420_prog", 64, 48, 505)]
d = ic.BatchDecoder(device=0, entropy_mode=1)
for f, w, h, pos in cases:
    raw = S.stream(f, w, h, 1)
    for bit in range(8):
        b = bytearray(raw); b[pos] ^= 1 << bit; data = bytes(b)
        want, *_ = o.load_from_memory(data, 1)
        got = d.decode([data], 1)[0]
        if want is None or got is None:
            print(f, pos, bit, 'oracle', want is None, 'gpu', got is None); continue
        df = np.argwhere(want != got)
        if len(df):
            y0, x0 = df[:, 0].min(), df[:, 1].min()
            print(f, pos, bit, len(df), 'differ; block at', y0, x0, 'to', df[:, 0].max(), df[:, 1].max())
            print(' want\n', want[y0:y0+8, x0:x0+8, 0]); print(' got\n', got[y0:y0+8, x0:x0+8, 0])
d.close()
