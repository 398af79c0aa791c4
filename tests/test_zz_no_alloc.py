"""jcu_idct_blocks_noalloc (include/jpeg_cuda_abi.h): the allocation-free variant of the IDCT stage entry point
must give the bytes of the existing entry point (jcu_test_idct_blocks) -- bit-exact, like every other integer
test here -- on two calls in a row that share the library's static scratch array.

Collected last on purpose (file name): a failure here must not hide other tests.
Passes alone in a fresh process and after the rest of the suite (it creates and destroys its own context)."""
import ctypes as C

import numpy as np
import pytest

import image_codecs_b200 as ic

pytestmark = pytest.mark.gpu

MAX_BLOCKS = 32768  # JCU_NOALLOC_MAX_BLOCKS


def _blocks(rng, n):
    coef = np.zeros((n, 64), dtype=np.int16)
    for i in range(n):
        k = int(rng.integers(1, 64))
        idx = rng.choice(64, size=k, replace=False)
        amp = [8, 64, 512, 2048, 32767][i % 5]
        coef[i, idx] = rng.integers(-amp, amp + 1, size=k)
    return coef


def test_no_alloc_entry_point_matches_existing_twice_in_a_row():
    import torch

    if ic.cuda_device_count() < 1:
        pytest.skip("no CUDA device")
    lib = ic.lib()
    lib.jcu_idct_blocks_noalloc.restype = C.c_int
    lib.jcu_idct_blocks_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    ctx = C.c_void_p()
    assert lib.jcu_ctx_create(0, 0, 0, C.byref(ctx)) == 0
    try:
        rng = np.random.default_rng(20261105)
        dev = torch.device("cuda:0")
        # second call: fewer blocks, other values, other table -- stale scratch of the first call must not leak
        for call, n in enumerate((3000, 1111)):
            coef = _blocks(rng, n)
            qt = rng.integers(1, 64, size=64).astype(np.uint16)
            want = np.zeros((n, 64), dtype=np.uint8)
            assert lib.jcu_test_idct_blocks(ctx, coef.ctypes.data, qt.ctypes.data, n, want.ctypes.data) == 0
            # the test (not the entry point) owns the device buffers of inputs and outputs
            d_coef = torch.from_numpy(coef.view(np.int16)).to(dev)
            d_qt = torch.from_numpy(qt.view(np.int16)).to(dev)
            d_out = torch.full((n, 64), 0xEE, dtype=torch.uint8, device=dev)
            torch.cuda.synchronize()
            rc = lib.jcu_idct_blocks_noalloc(ctx, d_coef.data_ptr(), d_qt.data_ptr(), n, d_out.data_ptr(), None)
            assert rc == 0, lib.jcu_error_string(ctx)
            torch.cuda.synchronize()  # the entry point does not synchronise
            got = d_out.cpu().numpy()
            assert np.array_equal(got, want), f"call {call}: {int((got != want).sum())} differing samples"
        # beyond the stated maximum: rejected with a clear message, nothing launched
        rc = lib.jcu_idct_blocks_noalloc(ctx, d_coef.data_ptr(), d_qt.data_ptr(), MAX_BLOCKS + 1, d_out.data_ptr(), None)
        assert rc != 0
        msg = lib.jcu_error_string(ctx).decode()
        assert "JCU_NOALLOC_MAX_BLOCKS" in msg and str(MAX_BLOCKS) in msg, msg
        # exactly the maximum works
        n = MAX_BLOCKS
        coef = _blocks(rng, 64).repeat(n // 64, axis=0)
        qt = np.ones(64, dtype=np.uint16)
        want = np.zeros((n, 64), dtype=np.uint8)
        assert lib.jcu_test_idct_blocks(ctx, coef.ctypes.data, qt.ctypes.data, n, want.ctypes.data) == 0
        d_coef = torch.from_numpy(coef).to(dev)
        d_qt = torch.from_numpy(qt.view(np.int16)).to(dev)
        d_out = torch.zeros((n, 64), dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        assert lib.jcu_idct_blocks_noalloc(ctx, d_coef.data_ptr(), d_qt.data_ptr(), n, d_out.data_ptr(), None) == 0
        torch.cuda.synchronize()
        assert np.array_equal(d_out.cpu().numpy(), want)
    finally:
        lib.jcu_ctx_destroy(ctx)
