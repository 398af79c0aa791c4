"""ctypes loader for the checkers under oracle/ -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Nothing in image_codecs_b200/ does (the product path has no CPU fallback).

Two interchangeable checkers, same C interface (oracle/ref_unity.c, oracle/jpeg_port.c):

  kind="reference"  oracle/_ref/liboracle_ref.so       the reference's own codec/jpeg.c, common.c,
                                                       convert.c, image_api.c compiled unmodified
                                                       (generic-C kernels; canonical)
  kind="reference_sse2"  oracle/_ref/liboracle_ref_sse2.so   same with STBI_SSE2 (timed CPU baseline)
  kind="port"       oracle/liboracle_port.so           this repo's CPU restatement of the path

`load()` prefers the reference build and falls back to the port when oracle/_ref is absent.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

_PATHS = {
    "reference": os.path.join(HERE, "_ref", "liboracle_ref.so"),
    "reference_sse2": os.path.join(HERE, "_ref", "liboracle_ref_sse2.so"),
    "port": os.path.join(HERE, "liboracle_port.so"),
}

u8p = C.POINTER(C.c_ubyte)


def fnv1a64(data: bytes) -> str:
    """FNV-1a 64 over bytes, hex -- the hash SURVEY.md 8(c) records probe results with."""
    h = 0xCBF29CE484222325
    arr = np.frombuffer(data, dtype=np.uint8)
    # chunked pure-python would be slow for MBs; do it in C-ish numpy-free loop only for small
    # inputs, else use the vectorised-by-blocks fallback below.
    for b in arr.tobytes():
        h ^= b
        h = (h * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return f"{h:016x}"


class Oracle:
    def __init__(self, kind: str):
        path = _PATHS[kind]
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.kind = kind
        self.path = path
        lib = C.CDLL(path)
        self.lib = lib
        lib.stbi_load_from_memory.restype = C.c_void_p
        lib.stbi_load_from_memory.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                              C.POINTER(C.c_int), C.c_int]
        lib.stbi_info_from_memory.restype = C.c_int
        lib.stbi_info_from_memory.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                              C.POINTER(C.c_int)]
        lib.stbi_failure_reason.restype = C.c_char_p
        lib.stbi_image_free.argtypes = [C.c_void_p]
        lib.oracle_idct_block.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        lib.oracle_ycbcr_row.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        lib.oracle_resample_row.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        self.has_capture = hasattr(lib, "oracle_decode_capture")  # the reference builds only (intermediate planes)
        if self.has_capture:
            lib.oracle_plane_geometry.restype = C.c_int
            lib.oracle_plane_geometry.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                                  C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                                  C.POINTER(C.c_int)]
            lib.oracle_decode_capture.restype = C.c_void_p
            lib.oracle_decode_capture.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                                  C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p),
                                                  C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        lib.oracle_time_decode.restype = C.c_double
        lib.oracle_time_decode.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int,
                                           C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_int)]
        lib.oracle_free.argtypes = [C.c_void_p]
        self.has_writer = hasattr(lib, "oracle_write_jpg")
        if self.has_writer:
            lib.oracle_write_jpg.restype = C.c_void_p
            lib.oracle_write_jpg.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int)]

    # ---- whole-image entry points (reference convert.c:254, image_api.c:119) ----
    def load_from_memory(self, data: bytes, req_comp: int = 0) -> Tuple[Optional[np.ndarray], int, int, int]:
        x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
        p = self.lib.stbi_load_from_memory(data, len(data), C.byref(x), C.byref(y), C.byref(c), req_comp)
        if not p:
            return None, 0, 0, 0
        n = req_comp if req_comp else c.value
        out = np.ctypeslib.as_array(C.cast(p, u8p), shape=(y.value, x.value, n)).copy()
        self.lib.stbi_image_free(p)
        return out, x.value, y.value, c.value

    def info_from_memory(self, data: bytes) -> Optional[Tuple[int, int, int]]:
        x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
        ok = self.lib.stbi_info_from_memory(data, len(data), C.byref(x), C.byref(y), C.byref(c))
        return (x.value, y.value, c.value) if ok else None

    def failure_reason(self) -> Optional[str]:
        r = self.lib.stbi_failure_reason()
        return r.decode() if r else None

    # ---- stage-level kernels ----
    def idct_block(self, coef: np.ndarray) -> np.ndarray:
        d = np.ascontiguousarray(coef, dtype=np.int16).reshape(64).copy()
        out = np.zeros((8, 8), dtype=np.uint8)
        self.lib.oracle_idct_block(out.ctypes.data, 8, d.ctypes.data)
        return out

    def ycbcr_row(self, y: np.ndarray, cb: np.ndarray, cr: np.ndarray, step: int = 3) -> np.ndarray:
        n = y.shape[0]
        out = np.zeros(n * step + 4, dtype=np.uint8)
        y = np.ascontiguousarray(y, np.uint8)
        cb = np.ascontiguousarray(cb, np.uint8)
        cr = np.ascontiguousarray(cr, np.uint8)
        self.lib.oracle_ycbcr_row(out.ctypes.data, y.ctypes.data, cb.ctypes.data, cr.ctypes.data, n, step)
        return out[: n * step].reshape(n, step)

    def resample_row(self, kind: int, near: np.ndarray, far: np.ndarray, hs: int) -> np.ndarray:
        w = near.shape[0]
        near = np.ascontiguousarray(near, np.uint8)
        far = np.ascontiguousarray(far, np.uint8)
        out = np.zeros(w * max(hs, 1) + 8, dtype=np.uint8)
        self.lib.oracle_resample_row(kind, out.ctypes.data, near.ctypes.data, far.ctypes.data, w, hs)
        return out[: w * hs]

    def plane_geometry(self, data: bytes):
        x, y = C.c_int(0), C.c_int(0)
        w2 = (C.c_int * 4)()
        h2 = (C.c_int * 4)()
        cx = (C.c_int * 4)()
        cy = (C.c_int * 4)()
        n = self.lib.oracle_plane_geometry(data, len(data), C.byref(x), C.byref(y), w2, h2, cx, cy)
        if not n:
            return None
        return dict(n=n, x=x.value, y=y.value, w2=list(w2)[:n], h2=list(h2)[:n], cx=list(cx)[:n], cy=list(cy)[:n])

    def decode_capture(self, data: bytes, req_comp: int = 0):
        """Full decode; returns (pixels, coef[list of int16 (bh,bw,64)], planes[list of u8 (h2,w2)], seen)."""
        g = self.plane_geometry(data)
        if g is None:
            return None
        coefs, planes, seens = [], [], []
        cp = (C.c_void_p * 4)()
        pp = (C.c_void_p * 4)()
        sp = (C.c_void_p * 4)()
        for i in range(g["n"]):
            bw, bh = g["w2"][i] // 8, g["h2"][i] // 8
            coefs.append(np.zeros((bh, bw, 64), dtype=np.int16))
            planes.append(np.zeros((g["h2"][i], g["w2"][i]), dtype=np.uint8))
            seens.append(np.zeros((bh, bw), dtype=np.uint8))
            cp[i] = coefs[i].ctypes.data
            pp[i] = planes[i].ctypes.data
            sp[i] = seens[i].ctypes.data
        x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
        p = self.lib.oracle_decode_capture(data, len(data), C.byref(x), C.byref(y), C.byref(c), req_comp, cp, pp, sp)
        if not p:
            return None
        n = req_comp if req_comp else c.value
        out = np.ctypeslib.as_array(C.cast(p, u8p), shape=(y.value, x.value, n)).copy()
        self.lib.stbi_image_free(p)
        return out, coefs, planes, seens

    # ---- encoder (reference jpeg_write.c:368) ----
    def write_jpg(self, rgb: np.ndarray, quality: int) -> bytes:
        if not self.has_writer:
            raise RuntimeError("this oracle has no jpeg_write.c")
        h, w, comp = rgb.shape
        rgb = np.ascontiguousarray(rgb, np.uint8)
        n = C.c_int(0)
        p = self.lib.oracle_write_jpg(w, h, comp, rgb.ctypes.data, quality, C.byref(n))
        if not p:
            raise RuntimeError("oracle_write_jpg failed")
        data = C.string_at(p, n.value)
        self.lib.oracle_free(p)
        return data

    # ---- timed CPU baseline ----
    def time_decode(self, streams: Sequence[bytes], req_comp: int, reps: int, nthreads: int):
        n = len(streams)
        bufs = (C.c_char_p * n)(*streams)
        lens = (C.c_int * n)(*[len(s) for s in streams])
        px = C.c_longlong(0)
        fails = C.c_int(0)
        secs = self.lib.oracle_time_decode(bufs, lens, n, req_comp, reps, nthreads, C.byref(px), C.byref(fails))
        return secs, px.value, fails.value


_cache = {}


def load(kind: Optional[str] = None) -> Oracle:
    """kind=None: the reference build if present, else the port."""
    kinds: List[str] = [kind] if kind else ["reference", "port"]
    last = None
    for k in kinds:
        if k in _cache:
            return _cache[k]
        try:
            _cache[k] = Oracle(k)
            return _cache[k]
        except (FileNotFoundError, OSError) as e:  # pragma: no cover - depends on what is built
            last = e
    raise RuntimeError(f"no oracle library available ({last}); run `make -C oracle`")


def available(kind: str) -> bool:
    return os.path.exists(_PATHS[kind])
