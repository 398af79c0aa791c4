"""ctypes binding for tools/jpeg_testenc.c -- the test-side stream generator.

Test/bench tooling only (the reference's jpeg_write.c cannot emit DRI/RSTn, gray,
4:4:4 at q<=90 or progressive streams, so configs C2-C5 need a generator).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libjpeg_testenc.so")

CS_YCBCR, CS_GRAY, CS_RGB, CS_CMYK, CS_YCCK, CS_YCBCR_NOJFIF = range(6)

SUBSAMPLING = {
    "444": ((1, 1, 1), (1, 1, 1)),
    "422": ((2, 1, 1), (1, 1, 1)),
    "420": ((2, 1, 1), (2, 1, 1)),
    "440": ((1, 1, 1), (2, 1, 1)),
    "411": ((4, 1, 1), (1, 1, 1)),
    "410": ((4, 1, 1), (2, 1, 1)),
    "441": ((1, 1, 1), (4, 1, 1)),
}


class Params(C.Structure):
    _fields_ = [
        ("width", C.c_int), ("height", C.c_int), ("in_channels", C.c_int), ("colorspace", C.c_int),
        ("quality", C.c_int), ("h", C.c_int * 4), ("v", C.c_int * 4), ("restart_interval", C.c_int),
        ("progressive", C.c_int), ("optimize", C.c_int), ("dqt16", C.c_int), ("noninterleaved", C.c_int),
        ("n_script", C.c_int), ("script", (C.c_int * 6) * 32), ("dri_after_tables", C.c_int),
        ("pad_fill_bytes", C.c_int),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            subprocess.check_call(["make", "-C", HERE, "-s"])
        _lib = C.CDLL(LIB)
        _lib.tje_encode.restype = C.c_int
        _lib.tje_encode.argtypes = [C.POINTER(Params), C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
        _lib.tje_free.argtypes = [C.c_void_p]
        _lib.tje_synth_rgb.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint]
        _lib.tje_encode_many.restype = C.c_int
        _lib.tje_encode_many.argtypes = [C.POINTER(Params), C.c_int, C.c_int, C.c_uint, C.c_int,
                                         C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
    return _lib


def make_params(width: int, height: int, *, quality: int = 90, subsampling: str = "420", colorspace: int = CS_YCBCR,
                restart_interval: int = 0, progressive: int = 0, optimize: bool = False, dqt16: bool = False,
                noninterleaved: bool = False, script: Optional[Sequence[Sequence[int]]] = None,
                sampling: Optional[Tuple[Sequence[int], Sequence[int]]] = None, in_channels: int = 3,
                dri_after_tables: bool = False, pad_fill_bytes: int = 0) -> Params:
    p = Params()
    p.width, p.height, p.in_channels, p.colorspace, p.quality = width, height, in_channels, colorspace, quality
    hs, vs = sampling if sampling is not None else SUBSAMPLING[subsampling]
    ncomp = {CS_GRAY: 1, CS_CMYK: 4, CS_YCCK: 4}.get(colorspace, 3)
    for i in range(4):
        p.h[i] = hs[i] if i < len(hs) and i < ncomp else 1
        p.v[i] = vs[i] if i < len(vs) and i < ncomp else 1
    if ncomp == 1:
        p.h[0] = p.v[0] = 1 if sampling is None else p.h[0]
    if ncomp == 4 and sampling is None:
        p.h[3], p.v[3] = p.h[0], p.v[0]
    p.restart_interval = restart_interval
    p.progressive = 3 if script is not None else progressive
    p.optimize = int(optimize)
    p.dqt16 = int(dqt16)
    p.noninterleaved = int(noninterleaved)
    p.dri_after_tables = int(dri_after_tables)
    p.pad_fill_bytes = pad_fill_bytes
    if script is not None:
        p.n_script = len(script)
        for i, row in enumerate(script):
            for j, v in enumerate(row):
                p.script[i][j] = v
    return p


def encode(pixels: np.ndarray, **kw) -> bytes:
    """pixels: (h, w, c) uint8.  Keyword arguments as make_params."""
    pixels = np.ascontiguousarray(pixels, dtype=np.uint8)
    if pixels.ndim == 2:
        pixels = pixels[:, :, None]
    h, w, c = pixels.shape
    p = make_params(w, h, in_channels=c, **kw)
    out = C.c_void_p()
    n = C.c_size_t()
    if not lib().tje_encode(C.byref(p), pixels.ctypes.data, C.byref(out), C.byref(n)):
        raise RuntimeError("tje_encode failed")
    data = C.string_at(out, n.value)
    lib().tje_free(out)
    return data


def synth_rgb(w: int, h: int, kind: int = 0, seed: int = 12345) -> np.ndarray:
    img = np.empty((h, w, 3), dtype=np.uint8)
    lib().tje_synth_rgb(img.ctypes.data, w, h, kind, seed)
    return img


def encode_many(count: int, width: int, height: int, *, kind: int = -1, base_seed: int = 1000,
                nthreads: Optional[int] = None, **kw) -> List[bytes]:
    """Synthesise `count` seeded images (seed = base_seed + i) and encode them with a thread pool."""
    p = make_params(width, height, in_channels=3, **kw)
    outs = (C.c_void_p * count)()
    lens = (C.c_size_t * count)()
    nthreads = nthreads or os.cpu_count() or 1
    ok = lib().tje_encode_many(C.byref(p), count, kind, base_seed, nthreads, outs, lens)
    res = []
    for i in range(count):
        res.append(C.string_at(outs[i], lens[i]) if outs[i] else b"")
        if outs[i]:
            lib().tje_free(outs[i])
    if not ok:
        raise RuntimeError("tje_encode_many failed")
    return res
