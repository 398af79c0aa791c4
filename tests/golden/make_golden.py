#!/usr/bin/env python
"""Generate tests/golden/*.json|npz from the REFERENCE build of the oracle (oracle/_ref).

Run in the build container (where /root/reference exists and `make -C oracle ref` has been run):

    python tests/golden/make_golden.py

The reference ships no golden vectors of its own, so these fixtures are
outputs of the reference's own jpeg.c (compiled unmodified behind oracle/ref_prelude.h) on this
repository's seeded test streams.  They pin
  * the test-side stream generator (stream sha256: the encoder must stay deterministic),
  * whichever oracle is available at test time (decoded sha256 for req_comp 0/1/3/4),
  * the four probe hashes recorded in SURVEY.md 8(c) (jpeg_write.c q90 LCG image, FNV-1a 64),
and they travel to the GPU box, where /root/reference does not exist.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import oracle_lib  # noqa: E402
from tests import streams as S  # noqa: E402


def sha(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()[:24]


def lcg_image(w, h):
    """The SURVEY 8(c) probe image: s = s*1664525+1013904223 per pixel, n = (s>>24)&31."""
    img = np.empty((h, w, 3), dtype=np.uint8)
    s = 12345
    for y in range(h):
        row = img[y]
        for x in range(w):
            s = (s * 1664525 + 1013904223) & 0xFFFFFFFF
            n = (s >> 24) & 31
            row[x, 0] = ((x * 255 // w) ^ n) & 255
            row[x, 1] = ((y * 255 // h) + n) & 255
            row[x, 2] = (((x + y) & 255) ^ (n << 1)) & 255
    return img


def main():
    o = oracle_lib.load("reference")
    out = {"generator": "tests/golden/make_golden.py", "oracle": "reference jpeg.c via oracle/_ref (generic-C kernels)",
           "matrix": [], "probe": {}}
    for fmt, w, h, kind in S.small_matrix():
        data = S.stream(fmt, w, h, kind)
        ent = {"fmt": fmt, "w": w, "h": h, "kind": kind, "stream_len": len(data), "stream_sha": sha(data), "decoded": {}}
        for rc in (0, 1, 3, 4):
            px, x, y, c = o.load_from_memory(data, rc)
            assert px is not None, (fmt, w, h, rc, o.failure_reason())
            ent["decoded"][str(rc)] = {"x": x, "y": y, "comp": c, "sha": sha(px.tobytes())}
        out["matrix"].append(ent)

    # SURVEY 8(c) probe: reference encoder + reference decoder, 512x512 q90
    img = lcg_image(512, 512)
    data = o.write_jpg(img, 90)
    px, x, y, c = o.load_from_memory(data, 3)
    out["probe"] = {"w": 512, "h": 512, "quality": 90, "stream_len": len(data),
                    "stream_fnv1a64": oracle_lib.fnv1a64(data), "decoded_fnv1a64": oracle_lib.fnv1a64(px.tobytes()),
                    "survey_expected": {"stream_len": 104701, "stream_fnv1a64": "b1ad09619f6beb7b",
                                        "decoded_fnv1a64": "cb27b971830dade0"}}
    np.savez_compressed(os.path.join(HERE, "probe512.npz"), stream=np.frombuffer(data, np.uint8))

    # a few complete vectors (stream + pixels) small enough to commit
    vec = {}
    for fmt, w, h in [("420", 17, 9), ("420_rst1", 64, 48), ("444", 17, 9), ("gray", 17, 9), ("420_prog", 64, 48),
                      ("422", 17, 9), ("cmyk", 17, 9), ("ycck", 17, 9), ("rgb", 17, 9)]:
        data = S.stream(fmt, w, h, 0)
        for rc in (0, 3):
            px, _, _, _ = o.load_from_memory(data, rc)
            vec[f"{fmt}_{w}x{h}_rc{rc}"] = px
        vec[f"{fmt}_{w}x{h}_stream"] = np.frombuffer(data, np.uint8)
    np.savez_compressed(os.path.join(HERE, "vectors.npz"), **vec)

    # stage-level known answers straight from the reference's static kernels
    rng = np.random.default_rng(20261104)
    coef = np.zeros((256, 64), dtype=np.int16)
    for i in range(256):
        k = int(rng.integers(1, 64))
        idx = rng.choice(64, size=k, replace=False)
        amp = [8, 64, 512, 1024][i % 4]
        coef[i, idx] = rng.integers(-amp, amp + 1, size=k)
    idct = np.stack([o.idct_block(c) for c in coef])
    ycc = rng.integers(0, 256, size=(4096, 3)).astype(np.uint8)
    rgb = o.ycbcr_row(ycc[:, 0].copy(), ycc[:, 1].copy(), ycc[:, 2].copy(), 3)
    near = rng.integers(0, 256, size=97).astype(np.uint8)
    far = rng.integers(0, 256, size=97).astype(np.uint8)
    np.savez_compressed(os.path.join(HERE, "kernels.npz"), coef=coef, idct=idct, ycc=ycc, rgb=rgb, near=near, far=far,
                        hv2=o.resample_row(2, near, far, 2), h2=o.resample_row(1, near, far, 2),
                        v2=o.resample_row(0, near, far, 1))
    json.dump(out, open(os.path.join(HERE, "golden.json"), "w"), indent=1)
    print("wrote", len(out["matrix"]), "matrix entries; probe", out["probe"])


if __name__ == "__main__":
    main()
