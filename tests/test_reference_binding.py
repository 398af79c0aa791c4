"""The reference-side binding, COMPILED:

* oracle/_ref/liboracle_ref_b200stub.so = the reference's own common.c / image_api.c / convert.c around a jpeg.c whose
  stbi__jpeg_load (jpeg.c:2442-2452) is the stub INTEGRATION.md section 2 prints, linked against the product library.
  Its stbi_load_from_memory is the REFERENCE's (convert.c:254): stbi__load_main (image_api.c:3-13) dispatches to the
  stub, the stub crosses the C-ABI.  Pixels, sizes and failure strings must equal the unmodified reference's.
* tests/native/consumer.c = a plain-C program that includes include/image_api.h and links the product.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

import image_codecs_b200 as ic
from tests import streams as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STUB = os.path.join(ROOT, "oracle", "_ref", "liboracle_ref_b200stub.so")
CONSUMER = os.path.join(ROOT, "tests", "native", "consumer")


def _fnv(a: np.ndarray) -> str:
    h = 0xCBF29CE484222325
    for b in a.tobytes():
        h = ((h ^ b) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return f"{h:016x}"


def test_integration_md_prints_the_stub_that_is_compiled():
    """The stub in INTEGRATION.md and the one oracle/ref_stub_unity.c compiles are the same text."""
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    src = open(os.path.join(ROOT, "oracle", "ref_stub_unity.c")).read()
    body = src[src.index("static stbi_jpeg_batch *stbi__b200;"):src.index("/* ---- END of the stub ---- */")].strip()
    assert body in doc, "INTEGRATION.md section 2 no longer shows the compiled stub"


def test_c_consumer_builds_and_reads_headers_without_a_gpu(tmp_path):
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "native")])
    f = tmp_path / "a.jpg"
    f.write_bytes(S.stream("420_rst16", 233, 177, 0))
    out = subprocess.run([CONSUMER, "--info", str(f)], capture_output=True, text=True, check=True).stdout
    assert out.strip() == "info 233 177 3"


@pytest.mark.gpu
def test_reference_dispatch_reaches_the_gpu_through_the_stub(oracle):
    if not os.path.exists(STUB):
        pytest.skip("oracle/_ref/liboracle_ref_b200stub.so not built (needs /root/reference at build time)")
    lib = C.CDLL(STUB)
    ip = C.POINTER(C.c_int)
    lib.refstub_load_from_memory.restype = C.c_void_p
    lib.refstub_load_from_memory.argtypes = [C.c_char_p, C.c_int, ip, ip, ip, C.c_int]
    lib.refstub_failure_reason.restype = C.c_char_p
    lib.refstub_free.argtypes = [C.c_void_p]
    lib.refstub_set_flip.argtypes = [C.c_int]
    cases = [c for c in S.small_matrix() if c[1:3] in ((233, 177), (64, 48), (17, 9))]
    for f, w, h, k in cases:
        data = S.stream(f, w, h, k)
        for rc in (0, 3, 4):
            want, ox, oy, oc = oracle.load_from_memory(data, rc)
            x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
            p = lib.refstub_load_from_memory(data, len(data), C.byref(x), C.byref(y), C.byref(c), rc)
            assert p, (f, w, h, rc, lib.refstub_failure_reason())
            n = rc if rc else c.value
            got = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_ubyte)), shape=(y.value, x.value, n)).copy()
            lib.refstub_free(p)
            assert (x.value, y.value, c.value) == (ox, oy, oc)
            assert np.array_equal(got, want), (f, w, h, rc)
    # the reference's own flip-on-load (convert.c:97-101) still runs around the stub
    data = S.stream("420", 233, 177, 0)
    want, *_ = oracle.load_from_memory(data, 3)
    lib.refstub_set_flip(1)
    x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
    p = lib.refstub_load_from_memory(data, len(data), C.byref(x), C.byref(y), C.byref(c), 3)
    lib.refstub_set_flip(0)
    got = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_ubyte)), shape=(y.value, x.value, 3)).copy()
    lib.refstub_free(p)
    assert np.array_equal(got, want[::-1])
    # failures keep the reference's strings: a foreign file never reaches the stub, a damaged JPEG fails inside it
    assert not lib.refstub_load_from_memory(b"GIF89a", 6, C.byref(x), C.byref(y), C.byref(c), 3)
    assert lib.refstub_failure_reason() == b"unknown image type"
    bad = bytearray(S.stream("420", 64, 48, 0))
    bad[bad.index(b"\xff\xc0") + 4] = 12
    assert not lib.refstub_load_from_memory(bytes(bad), len(bad), C.byref(x), C.byref(y), C.byref(c), 3)
    assert lib.refstub_failure_reason() == b"only 8-bit"


@pytest.mark.gpu
def test_c_consumer_decodes_on_the_gpu(oracle, tmp_path):
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "native")])
    for fmt, w, h, rc in (("420_rst16", 233, 177, 3), ("gray", 129, 65, 0), ("444", 64, 48, 4)):
        data = S.stream(fmt, w, h, 1)
        f = tmp_path / f"{fmt}.jpg"
        f.write_bytes(data)
        want, x, y, c = oracle.load_from_memory(data, rc)
        hh = _fnv(want)
        out = subprocess.run([CONSUMER, str(f), str(rc)], capture_output=True, text=True, check=True).stdout.split("\n")
        assert out[0] == f"info {x} {y} {c}"
        assert out[1] == f"load {x} {y} {c} {hh}"
        assert out[2] == f"batch 3 {hh} {hh} {hh}"
        assert out[3].startswith("multi ") and out[3].endswith(f" 3 {hh} {hh} {hh}"), out[3]
