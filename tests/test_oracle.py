"""CPU tests: the oracle and the test-side stream generator against the committed golden fixtures
(tests/golden/, produced by the reference's own jpeg.c through tests/golden/make_golden.py)."""
import hashlib
import json
import os

import numpy as np
import pytest

from tests import streams as S

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def sha(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()[:24]


@pytest.fixture(scope="module")
def golden():
    return json.load(open(os.path.join(GOLD, "golden.json")))


def test_stream_generator_is_deterministic(golden):
    """The seeded generator must reproduce the exact bytes the golden hashes were taken from."""
    for ent in golden["matrix"]:
        data = S.stream(ent["fmt"], ent["w"], ent["h"], ent["kind"])
        assert len(data) == ent["stream_len"], ent["fmt"]
        assert sha(data) == ent["stream_sha"], (ent["fmt"], ent["w"], ent["h"])


def test_oracle_matches_golden_matrix(oracle, golden):
    """Whichever oracle is loaded (reference build or port) must reproduce the reference's bytes."""
    for ent in golden["matrix"]:
        data = S.stream(ent["fmt"], ent["w"], ent["h"], ent["kind"])
        for rc, want in ent["decoded"].items():
            px, x, y, c = oracle.load_from_memory(data, int(rc))
            assert px is not None, (ent["fmt"], rc, oracle.failure_reason())
            assert (x, y, c) == (want["x"], want["y"], want["comp"])
            assert sha(px.tobytes()) == want["sha"], (oracle.kind, ent["fmt"], ent["w"], ent["h"], rc)


def test_oracle_matches_golden_vectors(oracle):
    vec = np.load(os.path.join(GOLD, "vectors.npz"))
    names = sorted(k[: -len("_stream")] for k in vec.files if k.endswith("_stream"))
    assert names
    for name in names:
        data = vec[name + "_stream"].tobytes()
        for rc in (0, 3):
            px, _, _, _ = oracle.load_from_memory(data, rc)
            assert np.array_equal(px, vec[f"{name}_rc{rc}"]), (name, rc)


def test_oracle_kernels_match_golden(oracle):
    k = np.load(os.path.join(GOLD, "kernels.npz"))
    for c, want in zip(k["coef"], k["idct"]):
        assert np.array_equal(oracle.idct_block(c), want)
    ycc = k["ycc"]
    assert np.array_equal(oracle.ycbcr_row(ycc[:, 0].copy(), ycc[:, 1].copy(), ycc[:, 2].copy(), 3), k["rgb"])
    assert np.array_equal(oracle.resample_row(2, k["near"], k["far"], 2), k["hv2"])
    assert np.array_equal(oracle.resample_row(1, k["near"], k["far"], 2), k["h2"])
    assert np.array_equal(oracle.resample_row(0, k["near"], k["far"], 1), k["v2"])


def test_survey_probe_streams_have_the_recorded_lengths(oracle):
    """SURVEY.md 8(c) records four probe streams (jpeg_write.c on the LCG image) by length and FNV hash.  All four
    LENGTHS reproduce to the byte with the reference's own writer -- four different sizes and two sampling modes
    cannot agree by accident, so the streams are the survey's.  Its hash values do not reproduce under FNV-1a 64 nor
    FNV-1 64 (its probe program was not kept, so its hash routine cannot be inspected); the fixture therefore pins the
    hashes of THIS build (test below) and keeps the survey's values beside them for the record."""
    if not getattr(oracle, "has_writer", False):
        pytest.skip("this oracle has no jpeg_write.c")
    from tools import testenc as te

    for w, h, q, want in ((512, 512, 90, 104701), (1920, 1080, 90, 809327), (1920, 1080, 95, 2503615)):
        assert len(oracle.write_jpg(te.synth_rgb(w, h, 0, 12345), q)) == want, (w, h, q)


def test_probe_stream_decodes_to_golden_hash(oracle, golden):
    """512x512 q90 stream written by the reference's jpeg_write.c (BASELINE.json configs[0])."""
    from oracle import oracle_lib

    data = np.load(os.path.join(GOLD, "probe512.npz"))["stream"].tobytes()
    p = golden["probe"]
    assert len(data) == p["stream_len"] == p["survey_expected"]["stream_len"]
    px, x, y, c = oracle.load_from_memory(data, 3)
    assert (x, y, c) == (512, 512, 3)
    assert oracle_lib.fnv1a64(px.tobytes()) == p["decoded_fnv1a64"]


def test_same_coefficients_any_container(oracle):
    """SURVEY.md section 4 invariant: baseline / restart / optimised-Huffman / progressive / non-interleaved
    encodings of the same quantised coefficients decode to identical bytes."""
    ref = None
    for fmt in ("420", "420_rst16", "420_rst1", "420_opt", "420_prog", "420_prog_rst5", "420_nonint", "420_dri_late"):
        px, _, _, _ = oracle.load_from_memory(S.stream(fmt, 233, 177, 0), 3)
        assert px is not None, fmt
        if ref is None:
            ref = px
        assert np.array_equal(px, ref), fmt


def test_reference_build_agrees_with_its_sse2_build():
    """generic-C and SSE2 kernels agree on encoder-produced streams."""
    from oracle import oracle_lib

    if not (oracle_lib.available("reference") and oracle_lib.available("reference_sse2")):
        pytest.skip("oracle/_ref not built")
    a, b = oracle_lib.load("reference"), oracle_lib.load("reference_sse2")
    for fmt, w, h in [("420", 233, 177), ("444", 64, 48), ("gray", 129, 65), ("420_prog", 64, 48)]:
        d = S.stream(fmt, w, h, 1)
        assert np.array_equal(a.load_from_memory(d, 3)[0], b.load_from_memory(d, 3)[0])


# ---- the plain-C restatement (oracle/jpeg_port.c) is pinned the same way, and against the reference build ----
@pytest.fixture(scope="module")
def port():
    from oracle import oracle_lib

    try:
        return oracle_lib.load("port")
    except Exception as e:  # not built
        pytest.skip(f"oracle port not built: {e}")


def test_port_matches_golden_fixtures(port, golden):
    """oracle/jpeg_port.c against every committed fixture: the decoded-stream hashes of the whole format matrix,
    the stored pixel vectors, the kernel-level vectors and the jpeg_write.c probe stream."""
    test_oracle_matches_golden_matrix(port, golden)
    test_oracle_matches_golden_vectors(port)
    test_oracle_kernels_match_golden(port)
    test_probe_stream_decodes_to_golden_hash(port, golden)


def test_port_equals_reference_build(port):
    """Where the compiled reference is present (oracle/_ref), the port must agree with it byte for byte on the
    whole matrix, every req_comp, and on header queries."""
    from oracle import oracle_lib

    try:
        ref = oracle_lib.load("reference")
    except Exception:
        pytest.skip("oracle/_ref not present")
    for f, w, h, k in S.small_matrix():
        data = S.stream(f, w, h, k)
        assert ref.info_from_memory(data) == port.info_from_memory(data), (f, w, h)
        for rc in (0, 1, 2, 3, 4):
            a, b = ref.load_from_memory(data, rc), port.load_from_memory(data, rc)
            assert a[1:] == b[1:], (f, w, h, rc)
            assert np.array_equal(a[0], b[0]), (f, w, h, k, rc)


def test_port_follows_the_reference_bit_reader_on_damaged_entropy_data(port):
    """The port keeps the reference's 32-bit bit buffer and its refill points literally (jpeg.c:167-187, 205, 254, 271,
    283, 1185), because on damaged data they decide the verdict: the refill that meets a marker adds no bits, so a
    Huffman code that straddles the end of the data is "bad huffman code" (jpeg.c:218, 237) -- unless the fast-AC
    table took it (jpeg.c:335) -- and how far the reader stands ahead at a restart boundary decides whether the
    RSTn is seen (jpeg.c:1183).  Random damage inside the entropy-coded data (tables intact) of baseline files and
    of progressive files without restart markers -- where the reference never branches on memory it has not
    written: same verdict and same failure string, file by file."""
    from oracle import oracle_lib
    from tools import fuzz_probe as fp

    try:
        ref = oracle_lib.load("reference")
    except Exception:
        pytest.skip("oracle/_ref not present")
    checked = fails = 0
    for seed in (11, 12):
        for tag, data in fp.gen_cases(seed, 40):
            fmt = tag.split()[0]
            if "[entropy]" not in tag or ("prog" in fmt and "rst" in fmt):
                continue
            ref.load_from_memory(b"junk", 3)  # the reference keeps a stale string where it fails without naming a reason
            a = ref.load_from_memory(data, 3)[0]
            ra = ref.failure_reason() if a is None else None
            ra = None if ra == "unknown image type" else ra
            b = port.load_from_memory(data, 3)[0]
            rb = port.failure_reason() if b is None else None
            assert (a is None) == (b is None), (tag, ra, rb)
            if a is None:
                assert ra == rb, (tag, ra, rb)
                fails += 1
            checked += 1
    assert checked > 1500 and fails > 100, (checked, fails)
