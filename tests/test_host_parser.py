"""CPU tests of the host C side: marker parser accept/reject parity with the reference
(SURVEY.md B.5, 8(b) error strings), header queries, and the no-CPU-fallback rule."""
import numpy as np
import pytest

import image_codecs_b200 as ic
from tests import streams as S


def _mutations(d: bytes):
    sof = max(d.find(b"\xff\xc0"), d.find(b"\xff\xc2"))
    sos = d.find(b"\xff\xda")
    dqt = d.find(b"\xff\xdb")
    dht = d.find(b"\xff\xc4")
    assert min(sof, sos, dqt, dht) > 0
    m = {
        "empty": b"",
        "not_jpeg": b"not a jpeg at all",
        "soi_only": d[:2],
        "truncated_header": d[:sof + 6],
        "truncated_scan": d[: sos + 40],
        "no_eoi": d[:-2],
        "sof3_lossless": d[:sof] + b"\xff\xc3" + d[sof + 2:],
        "sof9_arith": d[:sof] + b"\xff\xc9" + d[sof + 2:],
        "precision_12": d[:sof + 4] + b"\x0c" + d[sof + 5:],
        "height_0": d[:sof + 5] + b"\x00\x00" + d[sof + 7:],
        "width_0": d[:sof + 7] + b"\x00\x00" + d[sof + 9:],
        "ncomp_2": d[:sof + 9] + b"\x02" + d[sof + 10:],
        "bad_sof_len": d[:sof + 2] + b"\x00\x09" + d[sof + 4:],
        "bad_h": d[:sof + 11] + b"\x51" + d[sof + 12:],
        "bad_v": d[:sof + 11] + b"\x15" + d[sof + 12:],
        "bad_tq": d[:sof + 12] + b"\x04" + d[sof + 13:],
        "bad_dri_len": d[:2] + b"\xff\xdd\x00\x05\x00\x10\x00" + d[2:],
        "bad_dqt_type": d[:dqt + 4] + b"\x20" + d[dqt + 5:],
        "bad_dqt_table": d[:dqt + 4] + b"\x04" + d[dqt + 5:],
        "bad_dht_header": d[:dht + 4] + b"\x24" + d[dht + 5:],
        "bad_code_lengths": d[:dht + 5] + b"\xff" * 3 + d[dht + 8:],
        "bad_sos_count": d[:sos + 4] + b"\x05" + d[sos + 5:],
        "bad_sos_len": d[:sos + 2] + b"\x00\x0d" + d[sos + 4:],
        "bad_sos_comp_id": d[:sos + 5] + b"\x63" + d[sos + 6:],
        "bad_dc_huff": d[:sos + 6] + b"\x40" + d[sos + 7:],
        "bad_ac_huff": d[:sos + 6] + b"\x04" + d[sos + 7:],
        "bad_sos_se": d[:sos + 12] + b"\x40" + d[sos + 13:],
        "app_len_1": d[:2] + b"\xff\xe1\x00\x01" + d[2:],
        "com_len_0": d[:2] + b"\xff\xfe\x00\x00" + d[2:],
        "junk_before_sof_ok": d[:2] + b"\x00\x01\x02" + d[2:],
        "fill_ff_before_marker_ok": d[:sof] + b"\xff\xff\xff" + d[sof:],
        "app14_adobe_ok": d[:2] + b"\xff\xee\x00\x0eAdobe\x00\x64\x00\x00\x00\x00\x01" + d[2:],
        "com_ok": d[:2] + b"\xff\xfe\x00\x05abc" + d[2:],
        "dnl_ok": None,  # filled below
        "dnl_bad_height": None,
    }
    h = int.from_bytes(d[sof + 5:sof + 7], "big")
    m["dnl_ok"] = d[:-2] + b"\xff\xdc\x00\x04" + h.to_bytes(2, "big") + d[-2:]
    m["dnl_bad_height"] = d[:-2] + b"\xff\xdc\x00\x04" + (h + 1).to_bytes(2, "big") + d[-2:]
    return m


@pytest.mark.parametrize("fmt", ["420", "444_rst7", "gray", "420_prog"])
def test_accept_reject_parity_with_reference(oracle, fmt):
    """Same verdict and same short failure string as the reference for malformed / unusual headers."""
    base = S.stream(fmt, 64, 48, 0)
    for name, data in _mutations(base).items():
        want_px, wx, wy, wc = oracle.load_from_memory(data, 3)
        want_reason = oracle.failure_reason() if want_px is None else None
        info = ic.inspect(data, 3)
        if name == "truncated_scan":
            # the reference "succeeds" on a truncated scan (jpeg.c:1184) only when a marker follows;
            # here both must at least agree on the verdict
            pass
        assert (info is not None) == (want_px is not None), (fmt, name, ic.failure_reason(), want_reason)
        if info is None:
            assert ic.failure_reason() == want_reason, (fmt, name)
        else:
            assert (info["width"], info["height"], info["comps_in_file"]) == (wx, wy, wc), (fmt, name)


def test_info_matches_reference(oracle):
    for fmt, w, h, k in S.small_matrix():
        d = S.stream(fmt, w, h, k)
        assert ic.stbi_info_from_memory(d) == oracle.info_from_memory(d), (fmt, w, h)
    assert ic.stbi_info_from_memory(b"\x89PNG\r\n") is None


def test_bad_req_comp_and_unknown_type_strings():
    d = S.stream("420", 64, 48, 0)
    assert ic.inspect(d, 5) is None and ic.failure_reason() == "bad req_comp"
    assert ic.inspect(d, -1) is None and ic.failure_reason() == "bad req_comp"
    assert ic.inspect(b"GIF89a", 0) is None and ic.failure_reason() == "unknown image type"


def test_geometry_and_path_selection():
    """Frame geometry (jpeg.c:1616-1637) and the output-kernel choice made on the host."""
    i = ic.inspect(S.stream("420_rst16", 233, 177, 0), 3)
    assert (i["hmax"], i["vmax"], i["mcu_x"], i["mcu_y"]) == (2, 2, 15, 12)
    assert i["n_segments"] == (15 * 12 + 15) // 16 and i["coef_blocks"] == 15 * 12 * 6
    assert i["path"] == 1  # fused 4:2:0
    assert ic.inspect(S.stream("444", 64, 48, 0), 3)["path"] == 2
    g = ic.inspect(S.stream("gray", 64, 48, 0), 0)
    assert (g["path"], g["n_out"], g["comps_in_file"]) == (3, 1, 1)
    # a YCbCr file asked for 1 channel decodes Y only (decode_n = 1, jpeg.c:2246)
    y = ic.inspect(S.stream("420", 64, 48, 0), 1)
    assert (y["decode_n"], y["n_out"], y["comps_in_file"]) == (1, 1, 3)
    p = ic.inspect(S.stream("420_prog", 64, 48, 0), 3)
    assert p["progressive"] == 1 and p["n_scans"] > 1
    assert ic.inspect(S.stream("422", 64, 48, 0), 3)["path"] == 4  # fused 4:2:2 strip kernel
    assert ic.inspect(S.stream("440", 64, 48, 0), 3)["path"] == 0  # generic two-pass path
    c = ic.inspect(S.stream("cmyk", 17, 9, 0), 0)
    assert (c["comps_in_file"], c["n_out"], c["color_mode"]) == (3, 3, 3)


def test_segment_offsets_follow_restart_markers():
    d = S.stream("420_rst1", 64, 48, 0)
    i = ic.inspect(d, 3)
    assert i["n_segments"] == 4 * 3
    n_rst = sum(d.count(bytes([0xFF, 0xD0 + k])) for k in range(8))
    assert n_rst >= i["n_segments"] - 1


def test_no_cpu_fallback_without_gpu():
    """The product path must fail loudly when no CUDA device is usable."""
    if ic.cuda_device_count() > 0:
        pytest.skip("a GPU is present")
    px, x, y, c = ic.stbi_load_from_memory(S.stream("420", 64, 48, 0), 3)
    assert px is None and ic.failure_reason() == "no cuda device"
    with pytest.raises(RuntimeError):
        ic.BatchDecoder(device=0)


def test_product_never_imports_oracle():
    import os
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "image_codecs_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".c", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle|liboracle|oracle_lib|#include\s+\"[^\"]*oracle", src, re.M), f


def test_scan_waves_follow_data_dependencies():
    """Scans are scheduled by dependency level, not file order: scans on disjoint (component, band) sets share a
    launch.  libjpeg's default progressive script (tools/jpeg_testenc.c) is 10 scans in 5 levels: a first AC pass
    may store up to 15 positions past its band on a damaged stream (jpeg.c:461-463 does not check), so the luma
    band 6-63 waits for the band 1-5."""
    from image_codecs_b200 import api
    prog = S.stream("420_prog", 64, 48)
    assert api.inspect_scans(prog) == [0, 1, 1, 1, 2, 3, 1, 2, 2, 4]  # the DC refinement only waits for the first DC scan
    assert api.inspect_scans(S.stream("420", 64, 48)) == [0]
    # spectral selection only: DC of all components, then two bands per component: the components are independent of each other
    ss = api.inspect_scans(S.stream("444_prog_ss", 64, 48))
    assert ss[0] == 0 and set(ss[1:]) == {1, 2} and len(ss) == 7


def test_info_from_callbacks_reads_no_further_than_the_header(oracle):
    """stbi_info_from_callbacks (reference image_api.c:126): same answer as the reference's, short reads tolerated,
    and a source with more behind the frame header is not drained (the reference stops behind the SOF)."""
    import ctypes as C

    L = ic.lib()
    ip = C.POINTER(C.c_int)
    L.stbi_info_from_callbacks.restype = C.c_int
    L.stbi_info_from_callbacks.argtypes = [C.c_void_p, C.c_void_p, ip, ip, ip]
    ref = oracle.lib if oracle.kind.startswith("reference") else None
    if ref is not None:
        ref.stbi_info_from_callbacks.restype = C.c_int
        ref.stbi_info_from_callbacks.argtypes = [C.c_void_p, C.c_void_p, ip, ip, ip]
    for fmt, w, h in (("420", 233, 177), ("gray_prog", 64, 48), ("cmyk", 40, 31)):
        data = S.stream(fmt, w, h, 0) + bytes(1 << 20)
        for max_read in (1 << 30, 7):
            src = S.CallbackSource(data, max_read)
            x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
            assert L.stbi_info_from_callbacks(C.byref(src.struct), None, C.byref(x), C.byref(y), C.byref(c)) == 1
            assert (x.value, y.value) == (w, h)
            assert src.pos <= 2 * 4096 + 4096, src.pos  # one doubling step past the header, not the whole megabyte
            if ref is not None:
                src2 = S.CallbackSource(data, max_read)
                rx, ry, rc = C.c_int(0), C.c_int(0), C.c_int(0)
                assert ref.stbi_info_from_callbacks(C.byref(src2.struct), None, C.byref(rx), C.byref(ry), C.byref(rc)) == 1
                assert (rx.value, ry.value, rc.value) == (x.value, y.value, c.value)
    src = S.CallbackSource(b"no image here" * 100)
    x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
    assert L.stbi_info_from_callbacks(C.byref(src.struct), None, C.byref(x), C.byref(y), C.byref(c)) == 0
