"""GPU parity tests proper: CUDA path vs the oracle, through the C-ABI (bit-exact)."""
import collections
import os

import numpy as np
import pytest

import image_codecs_b200 as ic
from tests import streams as S

pytestmark = pytest.mark.gpu


def _check(oracle, got, data, req_comp, tag):
    want, x, y, c = oracle.load_from_memory(data, req_comp)
    assert want is not None, f"oracle failed on {tag}: {oracle.failure_reason()}"
    assert got is not None, f"gpu failed on {tag}"
    assert got.shape == want.shape, tag
    if not np.array_equal(got, want):
        d = np.argwhere(got != want)
        raise AssertionError(f"{tag}: {len(d)} differing bytes, first at {d[0]}: got {got[tuple(d[0])]} want {want[tuple(d[0])]}")


def test_idct_kernel_matches_reference(oracle):
    lib = ic.lib()
    import ctypes as C

    ctx = C.c_void_p()
    assert lib.jcu_ctx_create(0, 0, 0, C.byref(ctx)) == 0
    rng = np.random.default_rng(1)
    n = 4096
    coef = np.zeros((n, 64), dtype=np.int16)
    # sparse low-frequency blocks, dense blocks, extreme values, DC-only columns
    for i in range(n):
        k = rng.integers(1, 64)
        idx = rng.choice(64, size=k, replace=False)
        amp = [8, 64, 512, 2048, 32767][i % 5]
        coef[i, idx] = rng.integers(-amp, amp + 1, size=k)
    qt = rng.integers(1, 64, size=64).astype(np.uint16)
    out = np.zeros((n, 64), dtype=np.uint8)
    assert lib.jcu_test_idct_blocks(ctx, coef.ctypes.data, qt.ctypes.data, n, out.ctypes.data) == 0
    lib.jcu_ctx_destroy(ctx)
    deq = (coef.astype(np.int32) * qt.astype(np.int32)).astype(np.int16)
    for i in range(n):
        want = oracle.idct_block(deq[i])
        assert np.array_equal(out[i].reshape(8, 8), want), f"block {i}"


def test_ycbcr_kernel_matches_reference(oracle):
    """Colour arithmetic alone, through the TEST library (csrc/test_hooks.cu; the product exports no such hook)."""
    import ctypes as C

    hooks = C.CDLL(os.path.join(os.path.dirname(ic.lib_path()), "libimage_codecs_b200_testhooks.so"))
    hooks.jcu_test_ycbcr.restype = C.c_int
    hooks.jcu_test_ycbcr.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    # every (Cb, Cr) pair at 16 luma levels + every luma at 256 chroma pairs
    y, cb, cr = np.meshgrid(np.arange(0, 256, 17), np.arange(256), np.arange(256), indexing="ij")
    y, cb, cr = (a.astype(np.uint8).ravel() for a in (y, cb, cr))
    out = np.zeros((y.size, 3), dtype=np.uint8)
    assert hooks.jcu_test_ycbcr(0, y.ctypes.data, cb.ctypes.data, cr.ctypes.data, y.size, 3, out.ctypes.data) == 0
    want = oracle.ycbcr_row(y, cb, cr, 3)
    assert np.array_equal(out, want)


@pytest.mark.parametrize("req_comp", [0, 3])
def test_batch_matrix_bit_exact(oracle, decoder, req_comp):
    cases = S.small_matrix()
    datas = [S.stream(f, w, h, k) for f, w, h, k in cases]
    outs = decoder.decode(datas, req_comp)
    for (f, w, h, k), data, got in zip(cases, datas, outs):
        _check(oracle, got, data, req_comp, f"{f} {w}x{h} kind{k} req{req_comp}")


@pytest.mark.parametrize("req_comp", [1, 2, 4])
def test_batch_req_comp_variants(oracle, decoder, req_comp):
    cases = [c for c in S.small_matrix() if c[1:3] in ((233, 177), (17, 9))]
    datas = [S.stream(f, w, h, k) for f, w, h, k in cases]
    outs = decoder.decode(datas, req_comp)
    for (f, w, h, k), data, got in zip(cases, datas, outs):
        _check(oracle, got, data, req_comp, f"{f} {w}x{h} kind{k} req{req_comp}")


def test_generic_path_matches_too(oracle):
    """The two-pass generic path must agree with the fused kernels' formats as well."""
    d = ic.BatchDecoder(device=0, force_generic=1)
    cases = [c for c in S.small_matrix() if c[0] in ("420", "444", "gray", "420_rst16", "420_prog")]
    datas = [S.stream(f, w, h, k) for f, w, h, k in cases]
    for rc in (0, 4):
        outs = d.decode(datas, rc)
        for (f, w, h, k), data, got in zip(cases, datas, outs):
            _check(oracle, got, data, rc, f"generic {f} {w}x{h} req{rc}")
    d.close()


def test_per_image_api(oracle):
    for fmt, w, h in [("420", 233, 177), ("gray", 64, 48), ("420_prog", 129, 65)]:
        data = S.stream(fmt, w, h, 0)
        for rc in (0, 1, 3, 4):
            got, x, y, c = ic.stbi_load_from_memory(data, rc)
            want, ox, oy, oc = oracle.load_from_memory(data, rc)
            assert (x, y, c) == (ox, oy, oc)
            assert np.array_equal(got, want), (fmt, rc)
    assert ic.stbi_load_from_memory(b"not a jpeg", 0)[0] is None
    assert ic.failure_reason() == "unknown image type"
    assert ic.stbi_load_from_memory(S.stream("420", 64, 48, 0), 7)[0] is None
    assert ic.failure_reason() == "bad req_comp"


def test_batch_edge_cases(oracle):
    """Empty batch, empty and foreign buffers, header rejects and device-level failures between good images (a bad
    neighbour must not disturb the others), 1x1 / 1xN / Nx1 images, a batch that spans several pipeline chunks."""
    d = ic.BatchDecoder(device=0, chunk_images=4)
    assert d.decode([], 3) == []
    good = [S.stream("420", 1, 1, 0), S.stream("444", 1, 37, 1), S.stream("gray", 41, 1, 2), S.stream("420_rst1", 17, 9, 0),
            S.stream("420_prog", 8, 8, 1), S.stream("422", 15, 15, 2)]
    sof = good[3].index(b"\xff\xc0")
    bad_header = bytearray(good[3])
    bad_header[sof + 4] = 12  # 12-bit precision: "only 8-bit"
    truncated = good[4][:40]
    mixed = [good[0], b"", good[1], b"GIF89a-not-a-jpeg", good[2], bytes(bad_header), good[3], truncated, good[4], good[5]] * 3
    for rc in (0, 3):
        outs = d.decode(mixed, rc)
        assert len(outs) == len(mixed)
        for i, (data, got) in enumerate(zip(mixed, outs)):
            want, *_ = oracle.load_from_memory(data, rc) if data else (None,)
            if want is None:
                assert got is None, i
            else:
                assert got is not None and np.array_equal(got, want), i
    d.close()
    for data, why in ((b"", "unknown image type"), (bytes(bad_header), "only 8-bit")):
        assert ic.stbi_load_from_memory(data, 3)[0] is None and ic.failure_reason() == why


@pytest.mark.parametrize("rows_per_band", [0, 1, 5, 64])
def test_fused_420_strips_and_bands(oracle, rows_per_band):
    """4:2:0 strip kernel: images wider than one 640-pixel strip (2 and 3 strips, a 1-MCU last strip,
    widths that are not multiples of 16 or 8), bands of 1 / 5 / all MCU rows, RGB and RGBA, baseline and
    progressive (decode-time vs output-time dequantisation)."""
    if True:
        cases = [("420_rst16", 1300, 100, 0), ("420", 650, 50, 1), ("420", 641, 33, 2), ("420_rst1", 1924, 70, 0),
                 ("420_prog", 1283, 47, 1), ("420_opt", 656, 129, 2), ("420", 1280, 16, 0), ("420", 1288, 200, 1)]
        datas = [S.stream(f, w, h, k) for f, w, h, k in cases]
        d = ic.BatchDecoder(device=0, rows_per_band=rows_per_band)
        for rc in (3, 4):
            outs = d.decode(datas, rc)
            for (f, w, h, k), data, got in zip(cases, datas, outs):
                _check(oracle, got, data, rc, f"strip {f} {w}x{h} req{rc} rpb{rows_per_band}")
        d.close()


def test_fused_422_strips_and_edges(oracle):
    """4:2:2 strip kernel (jpeg_fused422.cuh; stbi__resample_row_h_2, jpeg.c:1784-1812): images wider than one
    480-pixel strip (2, 3 and 5 strips, a 1-MCU last strip), every residue of the width modulo 16 around a strip
    boundary -- the asymmetric out[2w-2] sample (jpeg.c:1805) then falls on every position of a pixel group, in the
    halo block and in the strip's first block --, w_lores 1 and 2, more MCU rows than one band, RGB and RGBA, baseline,
    progressive (output-time dequantisation), restart markers, and the flipped store order."""
    from tools import testenc as te

    cases = []
    for w in list(range(465, 500)) + [1, 2, 3, 4, 15, 16, 17, 31, 32, 33, 959, 960, 961, 1443, 2400]:
        cases.append((w, 37 if w < 600 else 19, dict()))
    cases += [(1000, 300, dict()), (523, 140, dict(progressive=1)), (700, 90, dict(restart_interval=3)),
              (481, 129, dict(progressive=1, restart_interval=7)), (1920, 1080, dict(restart_interval=16))]
    datas = [te.encode(te.synth_rgb(w, h, i % 3, 600 + i), quality=90, subsampling="422", **kw) for i, (w, h, kw) in enumerate(cases)]
    d = ic.BatchDecoder(device=0)
    for rc in (3, 4, 0):
        outs = d.decode(datas, rc)
        for (w, h, kw), data, got in zip(cases, datas, outs):
            _check(oracle, got, data, rc, f"422 {w}x{h} {kw} req{rc}")
    ic.set_flip_vertically_on_load(True)
    try:
        outs = d.decode(datas[:40], 3)
        for (w, h, kw), data, got in zip(cases, datas, outs):
            want, *_ = oracle.load_from_memory(data, 3)
            assert got is not None and np.array_equal(got, want[::-1]), f"flipped 422 {w}x{h}"
    finally:
        ic.set_flip_vertically_on_load(False)
    d.close()


def test_full_size_1080p_420_rst16(oracle, decoder):
    """BASELINE.json configs[1] geometry: 1920x1080 4:2:0 q90 restart interval 16 (two distinct images)."""
    from tools import testenc as te

    datas = te.encode_many(2, 1920, 1080, kind=-1, base_seed=4242, quality=90, subsampling="420", restart_interval=16)
    outs = decoder.decode(datas, 3)
    for i, (data, got) in enumerate(zip(datas, outs)):
        _check(oracle, got, data, 3, f"1080p #{i}")


def test_full_batch_config_c2_every_image(oracle):
    """BASELINE.json configs[1] at its FULL size: 1024 x 1920x1080 4:2:0 q90 restart-16 in one batched call (64 distinct
    streams, cycled, as bench.py builds the batch).  Every one of the 1024 results -- 6.4 GB of pixels -- is compared
    with the reference's decode of its stream, so a fault that only shows at a particular place of a large batch
    (a CTA index, a chunk boundary of the pipeline, a slab offset beyond 4 GB) cannot hide behind a sample."""
    import ctypes as C
    from tools import testenc as te

    distinct = te.encode_many(64, 1920, 1080, kind=-1, base_seed=1000, quality=90, subsampling="420", restart_interval=16)
    wants = [oracle.load_from_memory(s_, 3)[0] for s_ in distinct]
    n, per = 1024, 1920 * 1080 * 3
    streams = [distinct[i % 64] for i in range(n)]
    out = np.empty(n * per, dtype=np.uint8)
    bufs, lens, keep = ic.BatchDecoder._inputs(streams)
    outs = (C.c_void_p * n)(*[out.ctypes.data + i * per for i in range(n)])
    caps = (C.c_size_t * n)(*([per] * n))
    st = (C.c_int * n)()
    d = ic.BatchDecoder(device=0, profile=False)
    rc = d.decode_into(bufs, lens, n, 3, outs, caps, None, None, None, st)
    d.close()
    assert rc == n and all(st[i] == 1 for i in range(n)), rc
    bad = [i for i in range(n) if not np.array_equal(out[i * per:(i + 1) * per].reshape(1080, 1920, 3), wants[i % 64])]
    assert not bad, bad[:8]


@pytest.mark.parametrize("config", ["c3", "c4"])
def test_full_batch_configs_c3_c4_every_image(oracle, config):
    """BASELINE.json configs[2] and [3] at their full batch size (256 x 1080p): c3 = 4:4:4 colour and 8-bit grey
    alternating, no restart markers (self-synchronising decoder at its large-batch subsequence length); c4 =
    progressive 4:2:0 without restart markers (all scans of an image pipelined in one launch).  Every image compared
    with the reference's decode of its stream."""
    from tools import testenc as te

    if config == "c3":
        col = te.encode_many(8, 1920, 1080, kind=-1, base_seed=1000, quality=90, subsampling="444")
        gry = te.encode_many(8, 1920, 1080, kind=-1, base_seed=1500, quality=90, colorspace=te.CS_GRAY)
        distinct = [x for pair in zip(col, gry) for x in pair]
    else:
        distinct = te.encode_many(16, 1920, 1080, kind=-1, base_seed=1000, quality=90, subsampling="420", progressive=1)
    wants = [oracle.load_from_memory(s_, 3)[0] for s_ in distinct]
    d = ic.BatchDecoder(device=0, profile=False)
    outs = d.decode([distinct[i % 16] for i in range(256)], 3)
    tm = d.timings()
    d.close()
    bad = [i for i, got in enumerate(outs) if got is None or not np.array_equal(got, wants[i % 16])]
    assert not bad, bad[:8]
    if config == "c3":
        assert tm["selfsync_scans"] == 256, tm


def test_full_size_2160p_config_c5(oracle, decoder):
    """BASELINE.json configs[4] geometry: 3840x2160 4:2:0 q90 (12 strips x 135 MCU rows, 194 400 blocks per image,
    reference driver jpeg.c:2224-2440) with restart interval 16 AND without restart markers (self-synchronising
    path), an odd size one pixel past the MCU grid (3841x2161: a 1-MCU last strip, clamped chroma edges on both
    axes), a progressive file, and RGBA output -- every pixel against the compiled reference."""
    from tools import testenc as te

    datas, tags = [], []
    for i, d in enumerate(te.encode_many(2, 3840, 2160, kind=-1, base_seed=5150, quality=90, subsampling="420", restart_interval=16)):
        datas.append(d)
        tags.append(f"2160p rst16 #{i}")
    for i, d in enumerate(te.encode_many(2, 3840, 2160, kind=-1, base_seed=5152, quality=90, subsampling="420")):
        datas.append(d)
        tags.append(f"2160p no-rst #{i}")
    datas.append(te.encode(te.synth_rgb(3841, 2161, 1, 5154), quality=90, subsampling="420", restart_interval=16))
    tags.append("3841x2161 rst16")
    datas.append(te.encode(te.synth_rgb(3841, 2161, 0, 5155), quality=90, subsampling="420"))
    tags.append("3841x2161 no-rst")
    datas.append(te.encode(te.synth_rgb(3840, 2160, 1, 5156), quality=90, subsampling="420", progressive=1, restart_interval=16))
    tags.append("2160p progressive rst16")
    datas.append(te.encode(te.synth_rgb(3840, 2160, 0, 5157), quality=90, subsampling="420", progressive=1))
    tags.append("2160p progressive no-rst")
    for rc in (3, 4):
        outs = decoder.decode(datas, rc)
        for tag, data, got in zip(tags, datas, outs):
            _check(oracle, got, data, rc, f"{tag} req{rc}")
    tm = decoder.timings()
    assert tm["selfsync_scans"] >= 3, tm


@pytest.mark.parametrize("entropy_mode", [1, 2])
def test_restartless_streams_serial_and_selfsync(oracle, entropy_mode):
    """Baseline scans without restart markers: entropy_mode 1 = one thread per scan (serial, as the reference),
    2 = self-synchronising speculative decoder forced even for tiny scans.  Both must be bit-exact."""
    cases = [c for c in S.small_matrix()
             if c[0] in ("420", "444", "gray", "420_opt", "422", "440", "411", "420_nonint", "cmyk", "ycck", "420_q100",
                         "420_q10_dqt16", "rgb", "ycc_nojfif", "410", "441")]
    datas = [S.stream(f, w, h, k) for f, w, h, k in cases]
    d = ic.BatchDecoder(device=0, entropy_mode=entropy_mode)
    for rc in (0, 3):
        outs = d.decode(datas, rc)
        for (f, w, h, k), data, got in zip(cases, datas, outs):
            _check(oracle, got, data, rc, f"mode{entropy_mode} {f} {w}x{h} kind{k} req{rc}")
        tm = d.timings()
        assert (tm["selfsync_scans"] > 0) == (entropy_mode == 2), tm
    d.close()


def test_later_scan_of_a_component_wins(oracle):
    """T.81 allows a component to appear in several scans of a baseline file; the reference decodes them
    in file order, so the later one wins (jpeg.c:1716).  Two restart-less Cb scans long enough for the
    self-synchronising path, in both orders: a scan that shares a component with another scan of its image must keep
    its place in the wave order (it stays on the serial path), whatever the entropy mode."""
    from tools import testenc as te

    def split(data):
        idx = [i for i in range(2, len(data) - 1) if data[i] == 0xFF and data[i + 1] == 0xDA]
        eoi = data.rindex(b"\xff\xd9")
        parts = [data[idx[k]:(idx[k + 1] if k + 1 < len(idx) else eoi)] for k in range(len(idx))]
        return data[:idx[0]], parts, data[eoi:]

    a = te.encode(te.synth_rgb(640, 360, 0, 11), quality=90, subsampling="420", noninterleaved=True)
    b = te.encode(te.synth_rgb(640, 360, 2, 12), quality=90, subsampling="420", noninterleaved=True)
    (ha, pa, ea), (_, pb, _) = split(a), split(b)
    assert len(pa) == 3 and len(pb) == 3 and len(pb[1]) > 4096
    later_wins = ha + pa[0] + pa[1] + pa[2] + pb[1] + ea   # Cb of image b overwrites Cb of image a
    earlier_loses = ha + pb[1] + pa[0] + pa[1] + pa[2] + ea  # ... and here it is overwritten: equals image a
    for mode in (0, 1, 2):
        d = ic.BatchDecoder(device=0, entropy_mode=mode)
        outs = d.decode([later_wins, earlier_loses, a], 3)
        _check(oracle, outs[0], later_wins, 3, f"mode{mode} later Cb scan wins")
        _check(oracle, outs[1], earlier_loses, 3, f"mode{mode} earlier Cb scan is overwritten")
        assert np.array_equal(outs[1], outs[2]) and not np.array_equal(outs[0], outs[2])
        d.close()


def test_selfsync_full_size_444_and_gray(oracle, decoder):
    """BASELINE.json configs[2]: 1080p 4:4:4 colour and 8-bit grayscale without restart markers (auto mode picks
    the self-synchronising path for scans this long), plus 4:2:0 and a dense-noise image."""
    from tools import testenc as te

    datas = []
    datas += te.encode_many(2, 1920, 1080, kind=-1, base_seed=77, quality=90, subsampling="444")
    datas += te.encode_many(2, 1920, 1080, kind=-1, base_seed=78, quality=90, colorspace=te.CS_GRAY)
    datas += te.encode_many(2, 1920, 1080, kind=-1, base_seed=79, quality=90, subsampling="420")
    datas += te.encode_many(1, 1000, 700, kind=-1, base_seed=80, quality=100, subsampling="420", optimize=True)
    outs = decoder.decode(datas, 3)
    for i, (data, got) in enumerate(zip(datas, outs)):
        _check(oracle, got, data, 3, f"selfsync 1080p #{i}")
    tm = decoder.timings()
    assert tm["selfsync_scans"] >= 7 and 1 <= tm["selfsync_rounds"] <= 2000, tm  # dense q100 noise re-syncs slowly
    print("self-sync:", tm["selfsync_scans"], "scans,", tm["selfsync_rounds"], "refine rounds")


@pytest.mark.parametrize("sub_bytes", [32, 64, 128, 256])
def test_selfsync_subsequence_sizes(oracle, sub_bytes):
    """The self-synchronising decoder cuts a scan into subsequences of 32 ... 256 unstuffed bytes (short ones for small
    batches: the latency of every phase is the subsequence length; 256 when the batch fills the GPU).  Every size must
    give the reference's pixels: tiny forced scans (fewer bytes than one subsequence), blocks that span several
    subsequences (q100 dense noise), 4:4:4 / grey / 4:2:0 / 4:1:1 at sizes around 1080p."""
    from tools import testenc as te

    datas = [S.stream(f, w, h, k) for f, w, h, k in S.small_matrix() if f in ("420", "444", "gray", "411", "420_q100", "cmyk")]
    datas += te.encode_many(1, 1920, 1080, kind=-1, base_seed=91, quality=90, subsampling="444")
    datas += te.encode_many(1, 1283, 727, kind=-1, base_seed=92, quality=90, colorspace=te.CS_GRAY)
    datas += te.encode_many(1, 1000, 700, kind=-1, base_seed=93, quality=100, subsampling="420", optimize=True)
    datas += te.encode_many(1, 640, 360, kind=1, base_seed=94, quality=30, subsampling="420")  # few bits per block
    d = ic.BatchDecoder(device=0, entropy_mode=2, selfsync_sub_bytes=sub_bytes)
    for rc in (3, 0):
        outs = d.decode(datas, rc)
        for i, (data, got) in enumerate(zip(datas, outs)):
            _check(oracle, got, data, rc, f"sub{sub_bytes} #{i} req{rc}")
    assert d.timings()["selfsync_scans"] >= len(datas), d.timings()
    d.close()


@pytest.mark.parametrize("entropy_mode", [0, 1])
def test_progressive_restartless_scans(oracle, entropy_mode):
    """BASELINE.json configs[3]: progressive streams without restart markers.  entropy_mode 0 = warp-cooperative
    AC scan decoder (jpeg_entropy_prog.cuh), 1 = one thread per scan.  Full-size 1080p (smooth and dense-noise
    images, 4:2:0 / 4:4:4 / grey), deep successive-approximation scripts (long correction-bit runs, several
    refinement passes per band), spectral selection only, and the small-size matrix."""
    from tools import testenc as te

    datas, tags = [], []
    for kw in (dict(subsampling="420", progressive=1), dict(subsampling="444", progressive=2),
               dict(colorspace=te.CS_GRAY, progressive=1)):
        for i, d in enumerate(te.encode_many(2, 1920, 1080, kind=-1, base_seed=501, quality=90, **kw)):
            datas.append(d)
            tags.append(f"1080p {kw} #{i}")
    deep = [(7, 0, 0, 0, 2), (1, 1, 63, 0, 3), (2, 1, 63, 0, 2), (4, 1, 63, 0, 2), (1, 1, 63, 3, 2), (1, 1, 63, 2, 1),
            (2, 1, 63, 2, 1), (4, 1, 63, 2, 1), (7, 0, 0, 2, 1), (7, 0, 0, 1, 0), (1, 1, 63, 1, 0), (2, 1, 63, 1, 0),
            (4, 1, 63, 1, 0)]
    bands = [(7, 0, 0, 0, 0), (1, 1, 2, 0, 1), (1, 3, 9, 0, 1), (1, 10, 63, 0, 1), (2, 1, 63, 0, 0), (4, 1, 63, 0, 0),
             (1, 1, 9, 1, 0), (1, 10, 63, 1, 0)]
    for q in (50, 90, 100):
        for kind in (0, 1, 2):
            img = te.synth_rgb(640, 360, kind, 900 + q + kind)
            for name, script in (("deep", deep), ("bands", bands)):
                datas.append(te.encode(img, subsampling="420", quality=q, script=script))
                tags.append(f"{name} q{q} kind{kind}")
    for f, w, h, k in S.small_matrix():
        if "prog" in f:
            datas.append(S.stream(f, w, h, k))
            tags.append(f"{f} {w}x{h} kind{k}")
    d = ic.BatchDecoder(device=0, entropy_mode=entropy_mode)
    for rc in (3, 0):
        outs = d.decode(datas, rc)
        for tag, data, got in zip(tags, datas, outs):
            _check(oracle, got, data, rc, f"mode{entropy_mode} {tag} req{rc}")
    d.close()


def test_encoder_matches_reference_writer(oracle):
    """the GPU encoder must produce the byte stream of the reference's stbi_write_jpg_to_func
    (jpeg_write.c:368): float AAN DCT without FMA contraction, quantiser tables per quality, 4:2:0 at quality <= 90
    and 4:4:4 above, edge replication for sizes that are not MCU multiples, 1-4 input components, byte stuffing."""
    if not oracle.has_writer:
        pytest.skip("oracle without jpeg_write.c")
    from tools import testenc as te

    cases = []
    for (w, h) in ((64, 48), (233, 177), (17, 9), (1, 1), (16, 16), (640, 360), (1920, 1080)):
        for kind in (0, 1):
            img = te.synth_rgb(w, h, kind, 100 + w + kind)
            for q in ((90, 95, 30) if w < 1000 else (90,)):
                cases.append((f"{w}x{h} kind{kind} q{q} rgb", img, q))
    rng = np.random.default_rng(3)
    noise = rng.integers(0, 256, (200, 312, 3), dtype=np.uint8)  # dense noise: long codes, many 0xFF bytes
    for q in (1, 50, 90, 100, 0):
        cases.append((f"noise q{q}", noise, q))
    base = te.synth_rgb(100, 75, 0, 9)
    cases.append(("grey comp1", np.ascontiguousarray(base[:, :, :1]), 90))
    cases.append(("grey+alpha comp2", np.ascontiguousarray(base[:, :, :2]), 92))
    cases.append(("rgba comp4", np.concatenate([base, base[:, :, :1]], axis=2), 75))
    cases.append(("white", np.full((40, 56, 3), 255, np.uint8), 90))
    cases.append(("black", np.zeros((33, 47, 3), np.uint8), 100))
    for q in sorted({c[2] for c in cases}):
        group = [c for c in cases if c[2] == q]
        files, ms = ic.encode_batch([c[1] for c in group], q)
        for (tag, img, _), got in zip(group, files):
            want = oracle.write_jpg(img, q)
            assert got is not None, tag
            assert len(got) == len(want), f"{tag}: {len(got)} bytes, reference {len(want)}"
            if got != want:
                first = next(i for i in range(len(want)) if got[i] != want[i])
                raise AssertionError(f"{tag}: first difference at byte {first} of {len(want)}")
    # the reference-named entry point, and a round trip through the decoder
    img = te.synth_rgb(233, 177, 1, 5)
    data = ic.write_jpg_to_bytes(img, 90)
    assert data == oracle.write_jpg(img, 90)
    back, x, y, c = ic.stbi_load_from_memory(data, 3)
    want, *_ = oracle.load_from_memory(data, 3)
    assert (x, y, c) == (233, 177, 3) and np.array_equal(back, want)


def test_load_from_file_leaves_the_stream_behind_the_image(oracle, tmp_path):
    """stbi_load_from_file / stbi_load_from_file_16 (reference convert.c:199-226) 'unget' what the reader buffered but
    did not use: the FILE is left just behind the EOI marker, so images stored back to back load one after the other."""
    import ctypes as C

    a, b = S.stream("420_rst16", 97, 61, 1), S.stream("gray_prog", 40, 33, 2)
    path = tmp_path / "two.jpg"
    path.write_bytes(a + b + b"trailing")
    libc = C.CDLL(None)
    libc.fopen.restype = C.c_void_p
    libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
    libc.ftell.restype = C.c_long
    libc.ftell.argtypes = [C.c_void_p]
    libc.fclose.argtypes = [C.c_void_p]
    L = ic.lib()
    ip = C.POINTER(C.c_int)
    for fn, ctype, scale in (("stbi_load_from_file", C.c_uint8, 1), ("stbi_load_from_file_16", C.c_uint16, 257)):
        f = getattr(L, fn)
        f.restype = C.c_void_p
        f.argtypes = [C.c_void_p, ip, ip, ip, C.c_int]
        L.stbi_image_free.argtypes = [C.c_void_p]
        fp = libc.fopen(str(path).encode(), b"rb")
        assert fp
        try:
            end = 0
            for data in (a, b):
                x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
                p_ = f(fp, C.byref(x), C.byref(y), C.byref(c), 3)
                assert p_, ic.failure_reason()
                got = np.ctypeslib.as_array(C.cast(p_, C.POINTER(ctype)), shape=(y.value, x.value, 3)).copy()
                L.stbi_image_free(p_)
                want, *_ = oracle.load_from_memory(data, 3)
                assert np.array_equal(got, want.astype(got.dtype) * scale)
                end += len(data)
                assert libc.ftell(fp) == end
            # what follows is no image: the call fails like the reference does on the same bytes
            x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
            assert not f(fp, C.byref(x), C.byref(y), C.byref(c), 3)
        finally:
            libc.fclose(fp)


def test_load_from_callbacks_stops_behind_the_image(oracle):
    """stbi_load_from_callbacks (reference convert.c:261): pixels as from memory, with short reads; a stream with
    more behind the image is read only until the image is complete (the reference: at most 128 bytes further)."""
    import ctypes as C

    L = ic.lib()
    ip = C.POINTER(C.c_int)
    L.stbi_load_from_callbacks.restype = C.c_void_p
    L.stbi_load_from_callbacks.argtypes = [C.c_void_p, C.c_void_p, ip, ip, ip, C.c_int]
    L.stbi_image_free.argtypes = [C.c_void_p]
    for fmt, w, h in (("420_rst16", 233, 177), ("444_prog_ss", 129, 65), ("gray", 31, 9)):
        data = S.stream(fmt, w, h, 1)
        want, *_ = oracle.load_from_memory(data, 3)
        for max_read in (1 << 30, 1000):
            src = S.CallbackSource(data + bytes(4 << 20), max_read)
            x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
            p_ = L.stbi_load_from_callbacks(C.byref(src.struct), None, C.byref(x), C.byref(y), C.byref(c), 3)
            assert p_, ic.failure_reason()
            got = np.ctypeslib.as_array(C.cast(p_, C.POINTER(C.c_uint8)), shape=(y.value, x.value, 3)).copy()
            L.stbi_image_free(p_)
            assert np.array_equal(got, want)
            assert src.pos <= max(2 * len(data), 8192), (src.pos, len(data))


def test_flip_on_load_is_a_store_epilogue(oracle):
    """stbi_set_flip_vertically_on_load (reference convert.c:37-61, 97-101): the output kernels store bottom-up, so
    the batched calls -- host buffers and device-resident results alike -- return flipped rows without a host pass."""
    cases = [("420", 233, 177), ("420", 64, 16), ("444", 129, 65), ("gray", 31, 9), ("440", 50, 37), ("ycck", 48, 33),
             ("420_prog", 100, 71)]
    datas = [S.stream(f, w, h, i % 3) for i, (f, w, h) in enumerate(cases)]
    try:
        ic.set_flip_vertically_on_load(True)
        d = ic.BatchDecoder(device=0)
        for rc in (0, 1, 3, 4):
            outs = d.decode(datas, rc)
            for (f, w, h), s_, o in zip(cases, datas, outs):
                want, *_ = oracle.load_from_memory(s_, rc)
                assert o is not None and np.array_equal(o, want[::-1]), f"{f} {w}x{h} req_comp {rc}"
        # device-resident result: read the pixel slab where the kernels left it
        import torch

        d.prepare(datas, 3)
        d.run()
        d.sync()
        for i, s_ in enumerate(datas):
            ptr, nbytes = d.device_pixels(i)
            want, *_ = oracle.load_from_memory(s_, 3)
            assert nbytes == want.size
            view = type("DevView", (), {"__cuda_array_interface__": {
                "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}})()
            host = torch.as_tensor(view, device="cuda:0").cpu().numpy()
            assert np.array_equal(host.reshape(want.shape), want[::-1]), cases[i]
        d.close()
    finally:
        ic.set_flip_vertically_on_load(False)


def test_16_bit_and_float_variants(oracle):
    """stbi_load_16* / stbi_loadf* (reference convert.c:213-252, 286-337): the 8-bit decode widened exactly as the
    reference widens its own (v * 257; pow(v / 255, gamma) * scale, alpha linear), with the flip flag and a changed
    gamma/scale, and the reference's failure strings."""
    import ctypes as C

    ours = ic.lib()
    ref = None
    if oracle.kind.startswith("reference"):
        try:
            oracle.lib.stbi_loadf_from_memory, oracle.lib.oracle_set_flip
            ref = oracle.lib
        except AttributeError:
            ref = None
    ip = C.POINTER(C.c_int)
    for L in filter(None, (ours, ref)):
        L.stbi_load_16_from_memory.restype = C.c_void_p
        L.stbi_load_16_from_memory.argtypes = [C.c_char_p, C.c_int, ip, ip, ip, C.c_int]
        L.stbi_loadf_from_memory.restype = C.c_void_p
        L.stbi_loadf_from_memory.argtypes = [C.c_char_p, C.c_int, ip, ip, ip, C.c_int]
        L.stbi_ldr_to_hdr_gamma.argtypes = [C.c_float]
        L.stbi_ldr_to_hdr_scale.argtypes = [C.c_float]
        L.stbi_image_free.argtypes = [C.c_void_p]
        L.stbi_failure_reason.restype = C.c_char_p

    def call(L, fn, data, rc, dtype):
        x, y, c = C.c_int(0), C.c_int(0), C.c_int(0)
        p = getattr(L, fn)(data, len(data), C.byref(x), C.byref(y), C.byref(c), rc)
        if not p:
            return None, L.stbi_failure_reason()
        n = rc if rc else c.value
        ct = C.c_uint16 if dtype == np.uint16 else C.c_float
        out = np.ctypeslib.as_array(C.cast(p, C.POINTER(ct)), shape=(y.value, x.value, n)).copy()
        L.stbi_image_free(p)
        return out, (x.value, y.value, c.value)

    # fused 4:2:0, fused 1x1 (1 and 3 components) and the generic two-pass output path (4:2:2, CMYK)
    streams = [S.stream("420", 233, 177, 1), S.stream("gray", 64, 48, 0), S.stream("444_prog_ss", 129, 65, 2),
               S.stream("422", 75, 53, 1), S.stream("cmyk", 40, 31, 0)]
    for flip in (0, 1):
        ic.set_flip_vertically_on_load(bool(flip))
        if ref is not None:
            ref.oracle_set_flip(flip)
        for data in streams:
            for rc in (0, 1, 2, 3, 4):
                base, *_ = ic.stbi_load_from_memory(data, rc)
                got16, meta = call(ours, "stbi_load_16_from_memory", data, rc, np.uint16)
                assert np.array_equal(got16, base.astype(np.uint16) * 257)
                gotf, _ = call(ours, "stbi_loadf_from_memory", data, rc, np.float32)
                assert gotf.shape == base.shape
                if ref is not None:
                    w16, wmeta = call(ref, "stbi_load_16_from_memory", data, rc, np.uint16)
                    assert meta == wmeta and np.array_equal(got16, w16)
                    wf, _ = call(ref, "stbi_loadf_from_memory", data, rc, np.float32)
                    assert np.array_equal(gotf.view(np.uint32), wf.view(np.uint32)), f"float widening differs (req_comp {rc})"
    ic.set_flip_vertically_on_load(False)
    if ref is not None:
        ref.oracle_set_flip(0)
        for L in (ours, ref):
            L.stbi_ldr_to_hdr_gamma(1.8)
            L.stbi_ldr_to_hdr_scale(2.5)
        a, _ = call(ours, "stbi_loadf_from_memory", streams[0], 4, np.float32)
        b, _ = call(ref, "stbi_loadf_from_memory", streams[0], 4, np.float32)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        for L in (ours, ref):
            L.stbi_ldr_to_hdr_gamma(2.2)
            L.stbi_ldr_to_hdr_scale(1.0)
        bad = streams[0][:300]
        assert call(ours, "stbi_loadf_from_memory", bad, 3, np.float32) == call(ref, "stbi_loadf_from_memory", bad, 3, np.float32)
        assert call(ours, "stbi_load_16_from_memory", bad, 3, np.uint16) == call(ref, "stbi_load_16_from_memory", bad, 3, np.uint16)


def _marker_list(b):
    return [b[i + 1] for i in range(len(b) - 1) if b[i] == 0xFF and b[i + 1] not in (0x00, 0xFF)]


@pytest.mark.parametrize("entropy_mode", [0, 1, 2])
def test_damaged_streams_decode_like_the_reference(oracle, entropy_mode):
    """Truncated, bit-flipped, shortened and lengthened files: whatever the reference makes of them (zero bits past
    the end of a scan, coefficients stored past a band, wrapped values, 'bad huffman code', and 'unknown marker' when
    a restart segment is decoded before its bytes are used up -- jpeg.c:1186, 1722-1735) the CUDA path must make too:
    same verdict, same failure string, same bytes.  Damage that changes the marker structure itself (a lost table or
    RSTn) is left out: the reference then reads tables / sample planes it never initialised.  One narrow window is
    tolerated and counted: a stuffed 0xFF within two bytes of where the reference's bit buffer happens to stand
    when a progressive scan (or one decoded by the generic per-segment kernel) runs out of MCUs -- its fill level
    depends on the whole refill history; baseline scans replay the reference's reader there and are exact
    (DESIGN.md 7)."""
    rng = np.random.default_rng(11)
    fmts = ["420", "420_rst16", "420_rst1", "444", "444_rst7", "gray", "gray_rst5", "420_prog", "420_prog_rst5", "444_prog_ss",
            "gray_prog", "420_nonint", "420_nonint_rst11", "420_opt", "422", "420_rst3_fill", "420_dri_late"]
    cases = []
    for f in fmts:
        for (w, h) in ((233, 177), (64, 48)):
            raw = S.stream(f, w, h, 1)
            n = len(raw)
            sos = raw.index(b"\xff\xda")
            for frac in (0.5, 0.8, 0.97):
                cases.append((f"{f} {w}x{h} cut{frac}", bytes(raw[: int(n * frac)])))
            for t in range(14):
                b = bytearray(raw)
                kind = t % 4
                pos = int(rng.integers(max(n // 3, sos + 16), n - 2))
                if kind < 2:
                    b[pos] ^= 1 << int(rng.integers(0, 8))
                elif kind == 2:
                    del b[pos : pos + int(rng.integers(1, 9))]
                else:
                    b[pos:pos] = b[pos : pos + int(rng.integers(1, 9))]
                if _marker_list(b) == _marker_list(raw):
                    cases.append((f"{f} {w}x{h} damage{kind}.{t}@{pos}", bytes(b)))
    # one restart interval too many: a complete last interval swallows the extra RSTn and the marker loop skips what
    # follows (jpeg.c:1180-1189, 1722-1735); an incomplete one runs into it ("unknown marker")
    for f, w, h in (("420_rst1", 64, 48), ("gray_rst5", 80, 40), ("420_rst16", 233, 177), ("420_prog_rst5", 80, 80)):
        raw = S.stream(f, w, h, 1)
        rst = [i for i in range(len(raw) - 1) if raw[i] == 0xFF and 0xD0 <= raw[i + 1] <= 0xD7]
        for k in (0, len(rst) // 2, len(rst) - 2):
            b = bytearray(raw)
            b[rst[k + 1] : rst[k + 1]] = raw[rst[k] : rst[k + 1]]  # RSTn + its interval, twice
            cases.append((f"{f} {w}x{h} interval {k} doubled", bytes(b)))
    d = ic.BatchDecoder(device=0, entropy_mode=entropy_mode)
    outs = d.decode([c[1] for c in cases], 3)
    reasons = collections.Counter()
    window = 0  # verdicts that hinge on the reference's read-ahead at the end of a scan
    for (tag, data), got in zip(cases, outs):
        want, *_ = oracle.load_from_memory(data, 3)
        if want is None:
            why = oracle.failure_reason()
            reasons[why] += 1
            if got is not None and why == "unknown marker":
                window += 1
                continue
            assert got is None, f"mode{entropy_mode} {tag}: reference fails ({why}), gpu decodes"
        else:
            if got is None:
                ic.stbi_load_from_memory(data, 3)
                assert ic.failure_reason() == "unknown marker", f"mode{entropy_mode} {tag}: reference decodes, gpu fails"
                window += 1
                continue
            assert np.array_equal(got, want), f"mode{entropy_mode} {tag}: pixels differ"
    assert reasons["unknown marker"] > 10 and reasons["bad huffman code"] > 0, reasons  # both rules were exercised
    assert window <= len(cases) // 50, (window, len(cases))
    d.close()
    # the failure string itself, through the per-image entry point
    for tag, data in cases[:: max(1, len(cases) // 40)]:
        want, *_ = oracle.load_from_memory(data, 3)
        got, *_ = ic.stbi_load_from_memory(data, 3)
        if want is None and got is None:
            assert ic.failure_reason() == oracle.failure_reason(), tag


def test_damaged_fixtures_found_by_fuzzing(oracle):
    """tests/golden/damaged/: files tools/fuzz_probe.py found where the verdict hinges on the reference's bit reader.
    * 420_rst1 ... leftover_exactly_24_bits: a restart interval decoded with exactly 24 real bits left before the RSTn.
      The reference refills only below 24 buffered bits (jpeg.c:1185), so whether it meets the RSTn depends on how many of
      the 24 sat in its buffer already; k_entropy_baseline replays its reader for this one value (entb_replay) and
      must fail with "unknown marker" like the reference.
    * 420_prog ... refinement_code_straddles_the_end: an AC refinement scan whose last Huffman code is longer than the
      bits left before EOI.  stbi__jpeg_huff_decode's refill meets the marker and adds nothing, so the reference
      fails with "bad huffman code" (jpeg.c:218, 499); the progressive kernels feed zero bits there and decode
      (known gap) -- the test pins the reference's verdict and records ours."""
    here = os.path.join(os.path.dirname(__file__), "golden", "damaged")
    data = open(os.path.join(here, "420_rst1_97x61_leftover_exactly_24_bits.jpg"), "rb").read()
    assert oracle.load_from_memory(data, 3)[0] is None and oracle.failure_reason() == "unknown marker"
    for mode in (0, 2):
        d = ic.BatchDecoder(device=0, entropy_mode=mode)
        assert d.decode([data], 3)[0] is None, f"mode {mode}: the reference fails with 'unknown marker'"
        d.close()
    assert ic.stbi_load_from_memory(data, 3)[0] is None and ic.failure_reason() == "unknown marker"
    data = open(os.path.join(here, "420_prog_97x61_refinement_code_straddles_the_end.jpg"), "rb").read()
    assert oracle.load_from_memory(data, 3)[0] is None and oracle.failure_reason() == "bad huffman code"
    got = ic.stbi_load_from_memory(data, 3)[0]
    print("refinement code straddling the end of the data: gpu", "fails too" if got is None else "decodes (known gap)")


@pytest.mark.parametrize("entropy_mode", [0, 2])
def test_no_kernel_writes_past_a_device_buffer(oracle, entropy_mode):
    """Guard mode (compute-sanitizer is not available on every box): every device buffer is allocated at exactly
    the size the host asked for plus a 4 KB canary; after decoding the whole format matrix (all entropy paths:
    restart segments, progressive, serial and self-synchronising restart-less scans; all output kernels) and a
    set of corrupted streams, every canary must be intact and the pixels still bit-exact."""
    if True:
        d = ic.BatchDecoder(device=0, entropy_mode=entropy_mode, debug_guards=1)  # the pipeline objects inherit the switch
        cases = S.small_matrix()
        datas = [S.stream(f, w, h, k) for f, w, h, k in cases]
        for rc in (0, 3, 4):
            outs = d.decode(datas, rc)
            assert d.guards_ok() == 1, f"req_comp {rc}"
            for (f, w, h, k), data, got in zip(cases, datas, outs):
                _check(oracle, got, data, rc, f"guard {f} {w}x{h} kind{k} req{rc}")
        # truncated and bit-flipped entropy data must not make a kernel run off its buffers either
        rng = np.random.default_rng(5)
        bad = []
        for f, w, h, k in cases[:12]:
            raw = bytearray(S.stream(f, w, h, k))
            cut = bytes(raw[: max(200, len(raw) * 2 // 3)])
            for _ in range(8):
                raw[int(rng.integers(len(raw) // 2, len(raw) - 2))] ^= 1 << int(rng.integers(0, 8))
            bad += [cut, bytes(raw)]
        d.decode(bad, 3)
        assert d.guards_ok() == 1, "corrupted streams"
        d.close()


def _forge_sof_size(raw: bytes, w: int, h: int) -> bytes:
    b = bytearray(raw)
    sof = b.index(b"\xff\xc0")
    b[sof + 5 : sof + 9] = bytes([h >> 8, h & 255, w >> 8, w & 255])
    return bytes(b)


def test_oversized_image_fails_alone_and_chunks_follow_the_memory_budget(oracle):
    """a ~300-byte file may declare any size, and each forged image asks the device for that much memory
    (the real attack: 20000x30000, gigabytes).  It must not take the batch down.  Scaled to a 48 MB chunk budget:
    2000x1500 forgeries (18 MB each) decode -- the budget cuts the chunk after two of them and the rest follows in
    the next chunk -- a 4000x3000 forgery (73 MB) exceeds the budget by itself and fails alone, and every ordinary
    image of the batch decodes bit-exactly."""
    good = [S.stream(f, w, h, 0) for f, w, h in (("420", 233, 177), ("444", 64, 48), ("gray", 129, 65), ("420_rst16", 256, 128))]
    big = _forge_sof_size(S.stream("420", 17, 9, 0), 2000, 1500)
    huge = _forge_sof_size(S.stream("420", 17, 9, 0), 4000, 3000)
    batch = [good[0], big, good[1], huge, good[2], big, big, good[3], big]
    d = ic.BatchDecoder(device=0, device_mem_mb=48, chunk_images=16)
    outs = d.decode(batch, 3)
    for i, (data, got) in enumerate(zip(batch, outs)):
        if data is huge:
            assert got is None, i
        else:  # the forgeries decode in the reference too (zero bits past the end of the scan)
            _check(oracle, got, data, 3, f"image {i} of a batch with oversized neighbours")
    # staged form: what does not fit the budget of the one chunk fails alone with "outofmem", the rest decodes
    n = d.prepare(batch, 3)
    assert 6 <= n < len(batch) - 1, n
    d.run()
    d.sync()
    for i, data in enumerate(batch):
        if d.status(i):
            _check(oracle, d.fetch(i), data, 3, f"staged image {i}")
        else:
            assert data in (big, huge) and d.reason(i) == "outofmem", i
    d.close()


def test_blocks_no_scan_writes_are_zero_not_stale(oracle):
    """a baseline file that lost a restart marker (the scan ends early, jpeg.c:1184-1187) or a whole
    component leaves blocks that no scan writes.  The reference shows uninitialised heap there; here they must be
    ZERO coefficients -- never what an earlier batch left in the reused slab -- so the result is deterministic and
    equals the decode on a fresh object."""
    noise = [S.stream("420_rst16", 256, 128, 0), S.stream("444", 256, 128, 2)] * 4
    raw = bytearray(S.stream("420_rst1", 256, 128, 1))
    rst = [i for i in range(len(raw) - 1) if raw[i] == 0xFF and 0xD0 <= raw[i + 1] <= 0xD7]
    cut = rst[len(rst) // 2]
    eoi = raw.rindex(b"\xff\xd9")
    lost = bytes(raw[:cut]) + bytes(raw[eoi:])  # half the restart intervals, then EOI
    fresh = ic.BatchDecoder(device=0)
    want = fresh.decode([lost], 3)[0]
    fresh.close()
    assert want is not None
    d = ic.BatchDecoder(device=0)
    for _ in range(2):
        d.decode(noise, 3)  # fill the slab with other images' coefficients
        got = d.decode([lost], 3)[0]
        assert got is not None and np.array_equal(got, want)
    d.close()
    # rows the scan did reach equal the reference's; the rest is the IDCT of zero blocks (grey 128 luma, neutral chroma)
    ref, *_ = oracle.load_from_memory(lost, 3)
    assert ref is not None and np.array_equal(got[:48], ref[:48])
    assert np.all(got[-16:] == got[-1, -1])


def test_per_image_calls_are_reentrant(oracle):
    """The reference is re-entrant (fresh state per call, jpeg.c:2445-2450): 8 threads x 100 stbi_load_from_memory
    calls on different files must all be bit-exact, whatever interleaving the context pool gives them."""
    import threading

    cases = [("420", 233, 177), ("444", 129, 65), ("gray", 256, 128), ("420_rst16", 256, 128), ("420_prog", 129, 65),
             ("422", 64, 48), ("420_opt", 233, 177), ("cmyk", 64, 48)]
    datas = [S.stream(f, w, h, i % 3) for i, (f, w, h) in enumerate(cases)]
    wants = [oracle.load_from_memory(s_, 3)[0] for s_ in datas]
    errors = []

    def worker(t):
        try:
            for k in range(100):
                i = (t + k) % len(datas)
                got, x, y, c = ic.stbi_load_from_memory(datas[i], 3)
                if got is None or not np.array_equal(got, wants[i]):
                    errors.append((t, k, cases[i]))
                    return
            if ic.stbi_load_from_memory(b"junk", 3)[0] is not None or ic.failure_reason() != "unknown image type":
                errors.append((t, "failure reason is not thread-local"))
        except Exception as e:  # pragma: no cover
            errors.append((t, repr(e)))

    ths = [threading.Thread(target=worker, args=(t,)) for t in range(8)]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    assert not errors, errors[:3]


def test_pageable_destinations_go_through_the_shared_pinned_ring(oracle):
    """Downloads of 1 MB and more into malloc'ed (pageable) memory leave the device in 2 MB pieces through ONE ring of
    pinned slots per device, shared by every context on it (csrc/jpeg_cuda.cu, struct Bounce): the batched call with
    numpy buffers, and 6 threads of per-image loads (each borrows its own pooled context) at the same time.  Sizes with
    a ragged last piece; every byte against the reference."""
    import threading
    from tools import testenc as te

    datas = te.encode_many(3, 1920, 1080, kind=-1, base_seed=301, quality=90, subsampling="420", restart_interval=16)
    datas += te.encode_many(2, 1283, 727, kind=-1, base_seed=302, quality=90, subsampling="444")
    datas += te.encode_many(1, 801, 601, kind=-1, base_seed=303, quality=90, colorspace=te.CS_GRAY)
    wants = [oracle.load_from_memory(s_, 3)[0] for s_ in datas]
    errors = []

    def loads(t):
        try:
            for k in range(6):
                i = (t + k) % len(datas)
                got = ic.stbi_load_from_memory(datas[i], 3)[0]
                if got is None or not np.array_equal(got, wants[i]):
                    errors.append(("per-image", t, k, i))
        except Exception as e:  # pragma: no cover
            errors.append((t, repr(e)))

    def batches():
        try:
            d = ic.BatchDecoder(device=0, profile=False)
            for k in range(3):
                outs = d.decode(datas * 4, 3)  # numpy arrays: pageable
                for j, got in enumerate(outs):
                    if got is None or not np.array_equal(got, wants[j % len(datas)]):
                        errors.append(("batch", k, j))
            d.close()
        except Exception as e:  # pragma: no cover
            errors.append(("batch", repr(e)))

    ths = [threading.Thread(target=loads, args=(t,)) for t in range(6)] + [threading.Thread(target=batches)]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    assert not errors, errors[:3]


def test_multi_device_entry_point_in_c(oracle):
    """stbi_jpeg_decode_batch_multi: the split (contiguous, byte-balanced), one host thread + pipeline per device and
    the join all happen in C.  Two 'devices' are always driven -- the same GPU twice where the box has one, two
    different GPUs where it has more -- and every image must come back bit-exact and in place."""
    cases = S.small_matrix()[:48]
    datas = [S.stream(f, w, h, k) for f, w, h, k in cases]
    datas += [b"not a jpeg", datas[0][:50]]
    ngpu = ic.cuda_device_count()
    for devices in ([0, 0], [0, 1] if ngpu >= 2 else [0, 0, 0], None):
        m = ic.MultiDecoder(devices=devices, chunk_images=8)
        for rc in (3, 0):
            outs = m.decode(datas, rc)
            split = m.last_split()
            assert split[0] == 0 and split[-1] == len(datas) and split == sorted(split), split
            for i, (data, got) in enumerate(zip(datas, outs)):
                want, *_ = oracle.load_from_memory(data, rc)
                if want is None:
                    assert got is None, i
                else:
                    assert got is not None and np.array_equal(got, want), (devices, rc, i)
        if m.n_devices > 1:
            sizes = [sum(len(s_) for s_ in datas[a:b]) for a, b in zip(split, split[1:])]
            assert max(sizes) - min(sizes) <= 2 * max(len(s_) for s_ in datas), sizes
        assert m.decode([], 3) == []
        m.close()
