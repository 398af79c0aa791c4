"""ctypes binding of include/image_api.h and include/jpeg_cuda_abi.h."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_NAME = "libimage_codecs_b200.so"

u8p = C.POINTER(C.c_ubyte)


class LibraryMissing(RuntimeError):
    """The CUDA extension is not built: there is no fallback, run `python -c 'import __graft_entry__ as g; g.build()'`."""


def lib_path() -> str:
    # IMAGE_CODECS_B200_LIB: an alternative build of the same library (kernel tuning experiments, tools/build_variant.sh)
    return os.environ.get("IMAGE_CODECS_B200_LIB") or os.path.join(HERE, _LIB_NAME)


class Params(C.Structure):
    _fields_ = [("device", C.c_int), ("host_threads", C.c_int), ("chunk_images", C.c_int), ("profile", C.c_int),
                ("entropy_mode", C.c_int), ("pipeline_depth", C.c_int), ("device_mem_mb", C.c_int),
                ("no_download_merge", C.c_int), ("trace", C.c_int), ("force_generic", C.c_int),
                ("rows_per_band", C.c_int), ("debug_guards", C.c_int), ("selfsync_sub_bytes", C.c_int)]


class Timings(C.Structure):
    _fields_ = [("parse_ms", C.c_float), ("h2d_ms", C.c_float), ("entropy_ms", C.c_float), ("idct_ms", C.c_float),
                ("d2h_ms", C.c_float), ("total_ms", C.c_float), ("entropy_bits", C.c_double),
                ("coef_bytes", C.c_double), ("out_bytes", C.c_double), ("launches", C.c_int),
                ("selfsync_scans", C.c_int), ("selfsync_rounds", C.c_int), ("reserved", C.c_int * 3)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


_lib = None


def lib() -> C.CDLL:
    """Load the C-ABI library; raises LibraryMissing (never falls back to anything else)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise LibraryMissing(f"{path} not found -- build it with `make -C {os.path.join(HERE, 'csrc')}`")
    L = C.CDLL(path)
    ip = C.POINTER(C.c_int)
    L.stbi_load_from_memory.restype = C.c_void_p
    L.stbi_load_from_memory.argtypes = [C.c_char_p, C.c_int, ip, ip, ip, C.c_int]
    L.stbi_load.restype = C.c_void_p
    L.stbi_load.argtypes = [C.c_char_p, ip, ip, ip, C.c_int]
    L.stbi_info_from_memory.restype = C.c_int
    L.stbi_info_from_memory.argtypes = [C.c_char_p, C.c_int, ip, ip, ip]
    L.stbi_info.restype = C.c_int
    L.stbi_info.argtypes = [C.c_char_p, ip, ip, ip]
    L.stbi_failure_reason.restype = C.c_char_p
    L.stbi_image_free.argtypes = [C.c_void_p]
    L.stbi_set_flip_vertically_on_load.argtypes = [C.c_int]
    L.stbi_jpeg_inspect.restype = C.c_int
    L.stbi_jpeg_inspect.argtypes = [C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_longlong)]
    L.stbi_jpeg_inspect_scans.restype = C.c_int
    L.stbi_jpeg_inspect_scans.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.c_int]
    L.stbi_jpeg_cuda_device_count.restype = C.c_int
    L.stbi_jpeg_batch_create.restype = C.c_void_p
    L.stbi_jpeg_batch_create.argtypes = [C.POINTER(Params)]
    L.stbi_jpeg_batch_destroy.argtypes = [C.c_void_p]
    pp = C.POINTER(C.c_void_p)
    L.stbi_jpeg_decode_batch.restype = C.c_int
    L.stbi_jpeg_decode_batch.argtypes = [C.c_void_p, pp, ip, C.c_int, C.c_int, pp, C.POINTER(C.c_size_t), ip, ip, ip,
                                         ip]
    L.stbi_jpeg_batch_prepare.restype = C.c_int
    L.stbi_jpeg_batch_prepare.argtypes = [C.c_void_p, pp, ip, C.c_int, C.c_int, ip, ip, ip, ip]
    L.stbi_jpeg_batch_run.restype = C.c_int
    L.stbi_jpeg_batch_run.argtypes = [C.c_void_p]
    L.stbi_jpeg_batch_sync.restype = C.c_int
    L.stbi_jpeg_batch_sync.argtypes = [C.c_void_p]
    L.stbi_jpeg_batch_fetch.restype = C.c_int
    L.stbi_jpeg_batch_fetch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
    L.stbi_jpeg_batch_fetch_all.restype = C.c_int
    L.stbi_jpeg_batch_fetch_all.argtypes = [C.c_void_p, pp, C.POINTER(C.c_size_t)]
    L.stbi_jpeg_batch_device_pixels.restype = C.c_void_p
    L.stbi_jpeg_batch_device_pixels.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_size_t)]
    L.stbi_jpeg_batch_status.restype = C.c_int
    L.stbi_jpeg_batch_status.argtypes = [C.c_void_p, C.c_int]
    L.stbi_jpeg_batch_reason.restype = C.c_char_p
    L.stbi_jpeg_batch_reason.argtypes = [C.c_void_p, C.c_int]
    L.stbi_jpeg_batch_debug_guards.restype = C.c_int
    L.stbi_jpeg_batch_debug_guards.argtypes = [C.c_void_p]
    L.stbi_jpeg_batch_timings_get.restype = C.c_int
    L.stbi_jpeg_batch_timings_get.argtypes = [C.c_void_p, C.POINTER(Timings)]
    # thin CUDA ABI (stage-level tests, pinned memory)
    L.jcu_host_alloc.restype = C.c_void_p
    L.jcu_host_alloc.argtypes = [C.c_size_t]
    L.jcu_host_free.argtypes = [C.c_void_p]
    L.jcu_ctx_create.restype = C.c_int
    L.jcu_ctx_create.argtypes = [C.c_int, C.c_int, C.c_int, pp]
    L.jcu_ctx_destroy.argtypes = [C.c_void_p]
    L.jcu_test_idct_blocks.restype = C.c_int
    L.jcu_test_idct_blocks.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.jcu_read_coefs.restype = C.c_int
    L.jcu_read_coefs.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p]
    L.jcu_error_string.restype = C.c_char_p
    L.jcu_error_string.argtypes = [C.c_void_p]
    L.stbi_jpeg_multi_create.restype = C.c_void_p
    L.stbi_jpeg_multi_create.argtypes = [ip, C.c_int, C.POINTER(Params)]
    L.stbi_jpeg_multi_destroy.argtypes = [C.c_void_p]
    L.stbi_jpeg_multi_device_count.restype = C.c_int
    L.stbi_jpeg_multi_device_count.argtypes = [C.c_void_p]
    L.stbi_jpeg_multi_last_split.restype = C.c_int
    L.stbi_jpeg_multi_last_split.argtypes = [C.c_void_p, ip, C.c_int]
    L.stbi_jpeg_decode_batch_multi.restype = C.c_int
    L.stbi_jpeg_decode_batch_multi.argtypes = [C.c_void_p, pp, ip, C.c_int, C.c_int, pp, C.POINTER(C.c_size_t), ip, ip, ip,
                                               ip]
    L.stbi_jpeg_set_default_device.argtypes = [C.c_int]
    _lib = L
    return L


def failure_reason() -> Optional[str]:
    r = lib().stbi_failure_reason()
    return r.decode() if r else None


def cuda_device_count() -> int:
    return lib().stbi_jpeg_cuda_device_count()


def set_flip_vertically_on_load(flag: bool) -> None:
    lib().stbi_set_flip_vertically_on_load(int(bool(flag)))


def stbi_load_from_memory(data: bytes, req_comp: int = 0) -> Tuple[Optional[np.ndarray], int, int, int]:
    """(pixels[h, w, n] | None, x, y, comp) -- comp = components in the file (convert.c:254, jpeg.c:2436)."""
    x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
    p = lib().stbi_load_from_memory(data, len(data), C.byref(x), C.byref(y), C.byref(c), req_comp)
    if not p:
        return None, 0, 0, 0
    n = req_comp if req_comp else c.value
    out = np.ctypeslib.as_array(C.cast(p, u8p), shape=(y.value, x.value, n)).copy()
    lib().stbi_image_free(p)
    return out, x.value, y.value, c.value


def stbi_load(filename: str, req_comp: int = 0) -> Tuple[Optional[np.ndarray], int, int, int]:
    x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
    p = lib().stbi_load(os.fsencode(filename), C.byref(x), C.byref(y), C.byref(c), req_comp)
    if not p:
        return None, 0, 0, 0
    n = req_comp if req_comp else c.value
    out = np.ctypeslib.as_array(C.cast(p, u8p), shape=(y.value, x.value, n)).copy()
    lib().stbi_image_free(p)
    return out, x.value, y.value, c.value


def encode_batch(images: Sequence[np.ndarray], quality: int = 90, device: int = 0) -> Tuple[List[Optional[bytes]], float]:
    """stbi_jpeg_encode_batch: images are (h, w, comp) uint8 arrays (comp 1..4).  Returns (files, kernel_ms); the
    bytes of each file equal what the reference's stbi_write_jpg_to_func (jpeg_write.c:368) writes."""
    L = lib()
    n = len(images)
    arrs = [np.ascontiguousarray(a if a.ndim == 3 else a[:, :, None], dtype=np.uint8) for a in images]
    px = (C.c_void_p * n)(*[a.ctypes.data for a in arrs])
    xs = (C.c_int * n)(*[a.shape[1] for a in arrs])
    ys = (C.c_int * n)(*[a.shape[0] for a in arrs])
    cs = (C.c_int * n)(*[a.shape[2] for a in arrs])
    outs = (C.c_void_p * n)()
    lens = (C.c_int * n)()
    ms = C.c_float(0)
    L.stbi_jpeg_encode_batch.restype = C.c_int
    L.stbi_jpeg_encode_batch.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                                         C.c_void_p, C.POINTER(C.c_float)]
    L.stbi_jpeg_encode_batch(device, px, xs, ys, cs, n, quality, outs, lens, C.byref(ms))
    files: List[Optional[bytes]] = []
    for i in range(n):
        if outs[i]:
            files.append(C.string_at(outs[i], lens[i]))
            L.stbi_image_free(C.c_void_p(outs[i]))
        else:
            files.append(None)
    return files, float(ms.value)


def write_jpg_to_bytes(image: np.ndarray, quality: int = 90) -> Optional[bytes]:
    """stbi_write_jpg_to_func with a collecting callback (the reference's public writer entry point)."""
    L = lib()
    a = np.ascontiguousarray(image if image.ndim == 3 else image[:, :, None], dtype=np.uint8)
    chunks: List[bytes] = []
    CB = C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_int)
    cb = CB(lambda ctx, data, size: chunks.append(C.string_at(data, size)))
    L.stbi_write_jpg_to_func.restype = C.c_int
    L.stbi_write_jpg_to_func.argtypes = [CB, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
    ok = L.stbi_write_jpg_to_func(cb, None, a.shape[1], a.shape[0], a.shape[2], a.ctypes.data, quality)
    return b"".join(chunks) if ok else None


def stbi_info_from_memory(data: bytes) -> Optional[Tuple[int, int, int]]:
    x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
    ok = lib().stbi_info_from_memory(data, len(data), C.byref(x), C.byref(y), C.byref(c))
    return (x.value, y.value, c.value) if ok else None


_INSPECT_KEYS = ["width", "height", "comps_in_file", "progressive", "n_scans", "n_segments", "path", "color_mode",
                 "n_out", "decode_n", "hmax", "vmax", "mcu_x", "mcu_y", "coef_blocks", "entropy_bytes"]


def inspect(data: bytes, req_comp: int = 0) -> Optional[dict]:
    """Host-only full marker parse (no GPU).  None on failure (see failure_reason())."""
    info = (C.c_longlong * 16)()
    if not lib().stbi_jpeg_inspect(data, len(data), req_comp, info):
        return None
    return dict(zip(_INSPECT_KEYS, list(info)))


def inspect_scans(data: bytes) -> List[int]:
    """Dependency level ("wave") of every scan of the stream (host only); [] if it does not parse."""
    cap = 256
    waves = (C.c_int * cap)()
    n = lib().stbi_jpeg_inspect_scans(data, len(data), waves, cap)
    return list(waves[:min(n, cap)])


class PinnedBuffer:
    """cudaHostAlloc'ed host memory exposed as a numpy uint8 array."""

    def __init__(self, nbytes: int):
        self.nbytes = int(nbytes)
        self.ptr = lib().jcu_host_alloc(max(self.nbytes, 1))
        if not self.ptr:
            raise MemoryError("jcu_host_alloc failed (no CUDA device?)")
        self.array = np.ctypeslib.as_array(C.cast(self.ptr, u8p), shape=(max(self.nbytes, 1),))

    def free(self):
        if self.ptr:
            self.array = None
            lib().jcu_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class BatchDecoder:
    """stbi_jpeg_batch_* (include/image_api.h): one object drives one device."""

    def __init__(self, device: int = 0, host_threads: int = 0, chunk_images: int = 0, profile: bool = True,
                 entropy_mode: int = 0, **tuning):
        """tuning: pipeline_depth, device_mem_mb, no_download_merge, trace, force_generic, rows_per_band, debug_guards,
        selfsync_sub_bytes (stbi_jpeg_batch_params in include/image_api.h)."""
        p = Params()
        p.device, p.host_threads, p.chunk_images, p.profile, p.entropy_mode = (device, host_threads, chunk_images,
                                                                               int(profile), entropy_mode)
        for k, v in tuning.items():
            if k not in dict(Params._fields_):
                raise TypeError(f"unknown batch parameter {k}")
            setattr(p, k, int(v))
        self._h = lib().stbi_jpeg_batch_create(C.byref(p))
        if not self._h:
            raise RuntimeError(f"stbi_jpeg_batch_create failed: {failure_reason()} (no CPU fallback exists)")
        self.count = 0
        self._keep = None

    def close(self):
        if self._h:
            lib().stbi_jpeg_batch_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def _inputs(streams: Sequence):
        n = len(streams)
        bufs = (C.c_void_p * n)()
        lens = (C.c_int * n)()
        keep = []
        for i, s in enumerate(streams):
            if isinstance(s, np.ndarray):
                bufs[i] = s.ctypes.data
                lens[i] = s.nbytes
                keep.append(s)
            else:
                b = C.c_char_p(bytes(s))
                keep.append(b)
                bufs[i] = C.cast(b, C.c_void_p)
                lens[i] = len(s)
        return bufs, lens, keep

    # ---- staged form ----
    def prepare(self, streams: Sequence, req_comp: int = 0):
        n = len(streams)
        bufs, lens, keep = self._inputs(streams)
        xs, ys, cs, st = ((C.c_int * n)() for _ in range(4))
        rc = lib().stbi_jpeg_batch_prepare(self._h, bufs, lens, n, req_comp, xs, ys, cs, st)
        if rc < 0:
            raise RuntimeError(f"stbi_jpeg_batch_prepare failed: {failure_reason()}")
        self.count = n
        self.req_comp = req_comp
        self.xs, self.ys, self.comps = list(xs), list(ys), list(cs)
        return rc

    def run(self):
        if not lib().stbi_jpeg_batch_run(self._h):
            raise RuntimeError(f"stbi_jpeg_batch_run failed: {failure_reason()}")

    def sync(self):
        if not lib().stbi_jpeg_batch_sync(self._h):
            raise RuntimeError(f"stbi_jpeg_batch_sync failed: {failure_reason()}")

    def status(self, i: int) -> bool:
        return lib().stbi_jpeg_batch_status(self._h, i) == 1

    def reason(self, i: int) -> Optional[str]:
        r = lib().stbi_jpeg_batch_reason(self._h, i)
        return r.decode() if r else None

    def n_out(self, i: int) -> int:
        return self.req_comp if self.req_comp else self.comps[i]

    def fetch(self, i: int) -> Optional[np.ndarray]:
        if not self.xs[i]:
            return None
        n = self.n_out(i)
        out = np.empty((self.ys[i], self.xs[i], n), dtype=np.uint8)
        if not lib().stbi_jpeg_batch_fetch(self._h, i, out.ctypes.data, out.nbytes):
            return None
        return out

    def device_pixels(self, i: int) -> Tuple[int, int]:
        nb = C.c_size_t(0)
        p = lib().stbi_jpeg_batch_device_pixels(self._h, i, C.byref(nb))
        return (p or 0), nb.value

    def guards_ok(self) -> int:
        """Test hook (IMAGE_CODECS_B200_GUARD=1): 1 = no kernel wrote past the end of a device buffer, 0 = one did."""
        return lib().stbi_jpeg_batch_debug_guards(self._h)

    def timings(self) -> dict:
        t = Timings()
        lib().stbi_jpeg_batch_timings_get(self._h, C.byref(t))
        return t.as_dict()

    # ---- one call, host buffers in and out ----
    def decode(self, streams: Sequence, req_comp: int = 0) -> List[Optional[np.ndarray]]:
        """Decode a list of streams; returns one array (h, w, n) per image or None for failures."""
        n = len(streams)
        bufs, lens, keep = self._inputs(streams)
        xs, ys, cs, st = ((C.c_int * n)() for _ in range(4))
        rc = lib().stbi_jpeg_decode_batch(self._h, bufs, lens, n, req_comp, None, None, xs, ys, cs, st)
        if rc < 0:
            raise RuntimeError(f"stbi_jpeg_decode_batch (sizing) failed: {failure_reason()}")
        outs = (C.c_void_p * n)()
        caps = (C.c_size_t * n)()
        arrays: List[Optional[np.ndarray]] = []
        for i in range(n):
            if st[i] == 1:
                ch = req_comp if req_comp else cs[i]
                a = np.empty((ys[i], xs[i], ch), dtype=np.uint8)
                outs[i] = a.ctypes.data
                caps[i] = a.nbytes
                arrays.append(a)
            else:
                arrays.append(None)
        rc = lib().stbi_jpeg_decode_batch(self._h, bufs, lens, n, req_comp, outs, caps, xs, ys, cs, st)
        if rc < 0:
            raise RuntimeError(f"stbi_jpeg_decode_batch failed: {failure_reason()}")
        self.count = min(n, 64)
        self.last_status = list(st)
        return [a if st[i] == 1 else None for i, a in enumerate(arrays)]

    def decode_into(self, bufs, lens, n, req_comp, outs, caps, xs=None, ys=None, cs=None, st=None) -> int:
        """Raw pointer form for bench.py: ctypes arrays prepared by the caller (pinned buffers)."""
        return lib().stbi_jpeg_decode_batch(self._h, bufs, lens, n, req_comp, outs, caps, xs, ys, cs, st)


class MultiDecoder:
    """stbi_jpeg_multi_* (include/image_api.h): every listed device behind one call, split and threaded in C."""

    def __init__(self, devices: Optional[Sequence[int]] = None, host_threads: int = 0, chunk_images: int = 0, **tuning):
        p = Params()
        p.host_threads, p.chunk_images = host_threads, chunk_images
        for k, v in tuning.items():
            setattr(p, k, int(v))
        if devices is None:
            self._h = lib().stbi_jpeg_multi_create(None, 0, C.byref(p))
        else:
            arr = (C.c_int * len(devices))(*devices)
            self._h = lib().stbi_jpeg_multi_create(arr, len(devices), C.byref(p))
        if not self._h:
            raise RuntimeError(f"stbi_jpeg_multi_create failed: {failure_reason()} (no CPU fallback exists)")
        self.n_devices = lib().stbi_jpeg_multi_device_count(self._h)

    def close(self):
        if self._h:
            lib().stbi_jpeg_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def last_split(self) -> List[int]:
        first = (C.c_int * (self.n_devices + 1))()
        lib().stbi_jpeg_multi_last_split(self._h, first, self.n_devices + 1)
        return list(first)

    def decode(self, streams: Sequence, req_comp: int = 0) -> List[Optional[np.ndarray]]:
        n = len(streams)
        bufs, lens, keep = BatchDecoder._inputs(streams)
        xs, ys, cs, st = ((C.c_int * n)() for _ in range(4))
        rc = lib().stbi_jpeg_decode_batch_multi(self._h, bufs, lens, n, req_comp, None, None, xs, ys, cs, st)
        if rc < 0:
            raise RuntimeError(f"stbi_jpeg_decode_batch_multi (sizing) failed: {failure_reason()}")
        outs = (C.c_void_p * n)()
        caps = (C.c_size_t * n)()
        arrays: List[Optional[np.ndarray]] = []
        for i in range(n):
            if st[i] == 1:
                ch = req_comp if req_comp else cs[i]
                a = np.empty((ys[i], xs[i], ch), dtype=np.uint8)
                outs[i], caps[i] = a.ctypes.data, a.nbytes
                arrays.append(a)
            else:
                arrays.append(None)
        rc = lib().stbi_jpeg_decode_batch_multi(self._h, bufs, lens, n, req_comp, outs, caps, xs, ys, cs, st)
        if rc < 0:
            raise RuntimeError(f"stbi_jpeg_decode_batch_multi failed: {failure_reason()}")
        return [a if st[i] == 1 else None for i, a in enumerate(arrays)]

    def decode_into(self, bufs, lens, n, req_comp, outs, caps, xs=None, ys=None, cs=None, st=None) -> int:
        return lib().stbi_jpeg_decode_batch_multi(self._h, bufs, lens, n, req_comp, outs, caps, xs, ys, cs, st)
