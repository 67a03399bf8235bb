"""ptb_resolve_rgba8_noalloc (no counterpart in the reference): the resolve whose
device-side staging buffer is a static __device__ array instead of a per-call cudaMalloc.  Through the C-ABI, two calls in a row
(different frame sizes and tone maps, so that what the first call left in the shared array would show in the second), against the
existing entry point ptb_resolve_rgba8.  Tolerance: the one tests/test_gpu_output.py uses for the 8-bit picture — identical bytes.
Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C
import os

import numpy as np
import pytest

from opencl_path_tracer_b200 import capi, host
from tests import util

pytestmark = pytest.mark.gpu
MAX_PIXELS = 2 * 1024 * 1024            # PTB_NOALLOC_MAX_PIXELS of include/ptb200.h


def _rendered(ptb, spp):
    surf = host.load_ptbd(os.path.join(util.GOLDEN, "cornell_96x64.ptbd"))
    ptb.set_accel(capi.ACCEL_AUTO); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_AUTO)
    ptb.set_accumulate_sum(0); ptb.set_pixel_shard(0, 1, 1)
    s = host.scene_from_surface(surf, ptb)
    s.render_spp(spp)
    ptb.finish()
    return s


def _release(s):
    p = s.ptb
    p.finish()
    for h in (s.buffer_rays, s.buffer_rnds, s.cl_screen, s.buffer_colors, s.buffer_tris, s.buffer_mats, s.buffer_kd_tree, s.buffer_kd_tree_shift):
        if h:
            p.release(h)


def test_noalloc_resolve_equals_the_allocating_resolve_on_two_calls_in_a_row(ptb):
    assert "PTB_NOALLOC_MAX_PIXELS (2 * 1024 * 1024)" in open(os.path.join(util.ROOT, "include", "ptb200.h")).read()
    s = _rendered(ptb, 3)
    rng = np.random.default_rng(4)
    big = np.zeros((480, 640, 4), dtype=np.float32)                     # a second, larger frame of arbitrary radiance (with black pixels: L == 0)
    big[..., :3] = rng.random((480, 640, 3), dtype=np.float32) * 3.0
    big[100:140, 200:300, :3] = 0.0
    h_big = ptb.upload(big)
    try:
        # reference results from the existing entry point
        want_small = ptb.resolve_rgba8(s.buffer_colors, 96, 64, 0)
        want_big = ptb.resolve_rgba8(h_big, 640, 480, 1)
        want_big0 = ptb.resolve_rgba8(h_big, 640, 480, 0)
        # call 1: the big frame fills 1.2 MB of the shared scratch; call 2: the small frame must not see any of it; call 3: big again
        out_big = np.full((480, 640, 4), 7, dtype=np.uint8)
        out_small = np.full((64, 96, 4), 9, dtype=np.uint8)
        launches0 = ptb.launch_count()
        ptb.resolve_rgba8_noalloc(h_big, 640, 480, 1, out=out_big)
        ptb.resolve_rgba8_noalloc(s.buffer_colors, 96, 64, 0, out=out_small)
        assert ptb.launch_count() == launches0 + 2                      # one kernel per call
        assert (out_big == want_big).all()
        assert (out_small == want_small).all()
        assert (ptb.resolve_rgba8_noalloc(h_big, 640, 480, 0) == want_big0).all()
        assert (ptb.resolve_rgba8_noalloc(s.buffer_colors, 96, 64, 0) == want_small).all()
        assert want_small[..., :3].max() > 0 and (want_small[..., 3] == 255).all()
        # the largest frame the scratch holds, and one pixel more
        full = ptb.alloc(MAX_PIXELS * 16)
        edge = ptb.resolve_rgba8_noalloc(full, 2048, 1024, 0)
        assert (edge[..., :3] == 0).all() and (edge[..., 3] == 255).all()       # ptb_alloc zero-fills: black, opaque
        ptb.release(full)
        over = ptb.alloc((MAX_PIXELS + 2048) * 16)
        buf = np.zeros((1025, 2048, 4), dtype=np.uint8)
        rc = ptb.lib.ptb_resolve_rgba8_noalloc(over, 2048, 1025, 0, buf.ctypes.data)
        msg = ptb.last_error()
        ptb.clear_error()                                                # errors are sticky
        ptb.release(over)
        assert rc != 0 and "PTB_NOALLOC_MAX_PIXELS" in msg and "ptb_resolve_rgba8" in msg
        assert (buf == 0).all()                                          # rejected before anything ran
        # and the entry point still works after the rejected call
        assert (ptb.resolve_rgba8_noalloc(s.buffer_colors, 96, 64, 0) == want_small).all()
    finally:
        ptb.clear_error()
        ptb.release(h_big)
        _release(s)
