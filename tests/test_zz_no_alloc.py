"""rt_tonemap_noalloc — the no-allocation variant of rt_download's tonemap + read-back (scratch in a static __device__ array,
include/rt.h) — against the existing entry point, through the C ABI.  No counterpart in the reference.
In a file of its own that pytest collects last."""
import ctypes as C
import os

import numpy as np
import pytest

from rt_b200 import Renderer, capi

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _noalloc(lib, ptr, W, H, spp, gamma):
    out = np.full((H, W, 3), 0xAB, np.uint8)
    rc = lib.rt_tonemap_noalloc(0, None, C.c_void_p(ptr), W, H, spp, gamma, out.ctypes.data_as(C.POINTER(C.c_ubyte)))
    assert rc == 0, lib.rt_last_error(None)
    return out


def test_no_alloc_tonemap_equals_rt_download_on_two_calls_in_a_row():
    r = Renderer(0)
    lib = r.lib
    try:
        # first image: 320 x 180; two calls in a row (PNG tonemap, then the PPM writer's sqrt gamma) share the static scratch
        r.upload_json(os.path.join(G, "sample_scene.json"))
        r.override(width=320, height=180, spp=4)
        r.render(seed=3)
        _, want_png, _ = r.download(gamma=0)
        _, want_ppm, _ = r.download(gamma=1, want_sum=False)
        ptr, n = r.accum_device_ptr()
        assert n == 320 * 180 * 3
        got_png = _noalloc(lib, ptr, 320, 180, 4, 0)
        got_ppm = _noalloc(lib, ptr, 320, 180, 4, 1)
        assert np.array_equal(got_png, want_png) and np.array_equal(got_ppm, want_ppm)       # integer output: bit-exact
        assert not np.array_equal(got_png, got_ppm)
        # a smaller image next: the array still holds the previous call's bytes beyond the new image, which must not matter
        r.upload_json(os.path.join(G, "blue2.json"))
        r.override(width=96, height=54, spp=3)
        r.render(seed=4)
        _, want, _ = r.download(gamma=0)
        ptr, _ = r.accum_device_ptr()
        assert np.array_equal(_noalloc(lib, ptr, 96, 54, 3, 0), want)
        assert np.array_equal(_noalloc(lib, ptr, 96, 54, 3, 0), want)
        # sizes beyond the static scratch are rejected with a clear error, not truncated
        big_w = capi.RT_NOALLOC_MAX_BYTES // 3 + 1
        out = np.zeros(4, np.uint8)
        rc = lib.rt_tonemap_noalloc(0, None, C.c_void_p(ptr), big_w, 1, 1, 0, out.ctypes.data_as(C.POINTER(C.c_ubyte)))
        assert rc == capi.RT_ERR_INVALID and b"exceeds the static scratch" in lib.rt_last_error(None)
        assert lib.rt_tonemap_noalloc(0, None, None, 8, 8, 1, 0, out.ctypes.data_as(C.POINTER(C.c_ubyte))) == capi.RT_ERR_INVALID
    finally:
        r.close()
