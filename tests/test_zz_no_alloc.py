"""gf_genbod_static (csrc/gf_gen.cu): the phase-space generator for callers that may not allocate. Same kernel as gf_genbod,
staging area in a file-scope __device__ array instead of per-call cudaMalloc. Called through the C ABI and compared with
gf_genbod on two calls in a row (the second smaller and with other arguments, so stale contents of the shared array would show).
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from goofit_b200 import capi

pytestmark = pytest.mark.gpu

MASSES4 = [1.8645, 0.13957018, 0.13957018, 0.493677, 0.13957018]
MASSES3 = [1.86484, 0.13957018, 0.13957018, 0.1349766]
MAX_EVENTS = 32768  # GF_GENBOD_STATIC_MAX_EVENTS in include/goofit_b200.h


def _signature(f):
    f.restype = C.c_int
    f.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int, C.c_size_t, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    return f


def _generate(f, ctx, masses, n, seed, offset, want_dim5):
    nd = len(masses) - 1
    m = (C.c_double * len(masses))(*masses)
    particles = np.full((nd, n, 4), np.nan)
    weights = np.full(n, np.nan)
    dim5 = np.full((5, n), np.nan) if want_dim5 else None
    rc = f(ctx, m, nd, n, seed, offset, particles.ctypes.data, weights.ctypes.data, dim5.ctypes.data if want_dim5 else None)
    assert rc == 0, capi.last_error()
    return particles, weights, dim5


def _free_bytes(lib, ctx):
    sm, free, total = C.c_int(), C.c_size_t(), C.c_size_t()
    lib.gf_device_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]
    assert lib.gf_device_info(ctx, C.byref(sm), C.byref(free), C.byref(total)) == 0, capi.last_error()
    return free.value


def test_genbod_static_matches_genbod_on_two_calls_in_a_row():
    lib = capi.lib()
    ref, static = _signature(lib.gf_genbod), _signature(lib.gf_genbod_static)
    ctx = lib.gf_create(0)
    assert ctx, capi.last_error()
    try:
        # first call: the largest size the static array takes, four daughters, all three outputs
        want = _generate(ref, ctx, MASSES4, MAX_EVENTS, 4242, 0, True)
        got = _generate(static, ctx, MASSES4, MAX_EVENTS, 4242, 0, True)
        for w, g in zip(want, got):
            np.testing.assert_array_equal(g, w)  # the same kernel on the same counters: bit-identical
        assert np.isfinite(got[1]).all() and (got[1] > 0).all() and (got[1] <= 1).all()
        # second call straight after: fewer events, three daughters, another seed and offset. The array still holds the first
        # call's values; nothing of them may come back. No device memory may be taken or returned by the call.
        free_before = _free_bytes(lib, ctx)
        got = _generate(static, ctx, MASSES3, 1001, 77, 123456, False)
        assert _free_bytes(lib, ctx) == free_before
        want = _generate(ref, ctx, MASSES3, 1001, 77, 123456, False)
        for w, g in zip(want[:2], got[:2]):
            np.testing.assert_array_equal(g, w)
        # event i of a seed is the same whatever the batch: the tail of the first range, regenerated through `offset`
        tail = _generate(static, ctx, MASSES4, 500, 4242, MAX_EVENTS - 500, True)
        first = _generate(ref, ctx, MASSES4, MAX_EVENTS, 4242, 0, True)
        np.testing.assert_array_equal(tail[0], first[0][:, -500:, :])
        np.testing.assert_array_equal(tail[1], first[1][-500:])
        np.testing.assert_array_equal(tail[2], first[2][:, -500:])
        # beyond the array: a clear error, nothing launched
        launches = lib.gf_launch_count(ctx)
        m = (C.c_double * 5)(*MASSES4)
        w = np.zeros(MAX_EVENTS + 1)
        assert static(ctx, m, 4, MAX_EVENTS + 1, 1, 0, None, w.ctypes.data, None) != 0
        assert "exceed the static scratch array" in capi.last_error() and str(MAX_EVENTS) in capi.last_error()
        assert lib.gf_launch_count(ctx) == launches
    finally:
        lib.gf_destroy(ctx)
