"""Parity at BASELINE.json's full sizes (C2 = Cornell 1920x1080, 1024 spp, depth 4; C3 = 104 k triangles at
1080p), where the CPU oracle cannot render the whole frame in test time.  Two kinds of checks:

* the ORACLE ITSELF on a slice of the full frame: the reference's kernels rendering every 16th band of 8 scanlines of
  the 1920x1080 frame (same per-pixel seeds as the whole frame, main.cpp:522-526), compared bit for bit with those
  pixels of the GPU's full-frame render;
* size-independent properties of the path: splitting the samples over launches, splitting the pixels over shards,
  swapping the closest-hit structure or the pipeline must not change a single bit in the parity flavour; sum-mode
  accumulation + resolve and the fast flavour must agree with it within the stated floating-point tolerance.
"""
import numpy as np
import pytest

from opencl_path_tracer_b200 import capi, host, scenes
from tests import util

pytestmark = pytest.mark.gpu

W, H, DEPTH = 1920, 1080, 4
CORNELL_EYE, CORNELL_FOV = (277.5, 277.5, -800), 39.3


def cornell(ptb, accel, math, pipeline=capi.PIPE_MEGAKERNEL, seed_skip=0):
    ptb.set_accel(accel); ptb.set_math(math); ptb.set_pipeline(pipeline)
    s = host.Scene(ptb)
    s.init_Scene(W, H, seed_skip)
    s.add_Obj(util.CORNELL, (555, 0, 0), (1, 1, 1), 0, 0)
    s.upload_Triangles(); s.upload_Materials()
    s.iterations = DEPTH
    cam = host.camera_at(W, H, CORNELL_FOV, 0, 0, CORNELL_EYE)
    s._camera = lambda: cam
    return s


def release(s):
    p = s.ptb
    p.finish()
    for h in (s.buffer_rays, s.buffer_rnds, s.cl_screen, s.buffer_colors, s.buffer_tris, s.buffer_mats, s.buffer_kd_tree, s.buffer_kd_tree_shift):
        if h:
            p.release(h)


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def c2_parity_64(ptb):
    """the full C2 frame, parity flavour, brute force, 64 spp in one launch: the image the other tests compare against"""
    s = cornell(ptb, capi.ACCEL_BRUTE, capi.MATH_PARITY)
    s.render_spp(64)
    out = s.download_colors().copy(), s.download_rnds().copy()
    release(s)
    return out


def test_c2_full_frame_slice_matches_the_reference_kernels(ptb, c2_parity_64):
    if not util.have_ref_cpu():
        pytest.skip("oracle/_ref/ref_cpu not built")
    n, group = 16, 8 * W                       # rank 0 of 16: the bands of 8 scanlines number 0, 16, ... 128 (9 of 135 bands)
    ref = util.run_ref_cpu(W, H, 64, DEPTH, obj=util.CORNELL, eye=CORNELL_EYE, fov=CORNELL_FOV, want=("colors", "rnds"),
                           extra=("--shard", "0", str(n), str(group)))
    own = ((np.arange(W * H) // group) % n == 0).reshape(H, W)
    assert own.sum() == 9 * 8 * W
    colors, rnds = c2_parity_64
    assert (bits(colors[own][:, :3]) == bits(ref["colors"][own][:, :3])).all()
    assert (rnds.reshape(H, W)[own] == ref["rnds"].reshape(H, W)[own]).all()


def test_c2_samples_split_over_launches_is_bitwise_the_same(ptb, c2_parity_64):
    s = cornell(ptb, capi.ACCEL_BRUTE, capi.MATH_PARITY)
    for n in (1, 7, 24, 32):                   # 64 samples as four launches of uneven size
        s.render_spp(n)
    colors, rnds = s.download_colors(), s.download_rnds()
    release(s)
    assert (bits(colors) == bits(c2_parity_64[0])).all() and (rnds == c2_parity_64[1]).all()


@pytest.mark.parametrize("accel,pipeline", [(capi.ACCEL_LBVH, capi.PIPE_MEGAKERNEL), (capi.ACCEL_LBVH, capi.PIPE_WAVEFRONT),
                                            (capi.ACCEL_LBVH, capi.PIPE_POOL), (capi.ACCEL_REFBVH, capi.PIPE_MEGAKERNEL)])
def test_c2_closest_hit_structure_and_pipeline_do_not_change_a_bit(ptb, c2_parity_64, accel, pipeline):
    s = cornell(ptb, accel, capi.MATH_PARITY, pipeline)
    s.render_spp(64)
    colors, rnds = s.download_colors(), s.download_rnds()
    release(s)
    assert (bits(colors) == bits(c2_parity_64[0])).all() and (rnds == c2_parity_64[1]).all()


def test_c2_parity_prefilter_does_not_change_a_bit(ptb, c2_parity_64):
    """the parity flavour's conservative pre-filter (on by default, so c2_parity_64 used it) against the exact test on
    every triangle: ~50 M rays, identical bits and RNG states"""
    try:
        ptb.set_parity_filter(0)
        s = cornell(ptb, capi.ACCEL_BRUTE, capi.MATH_PARITY)
        s.render_spp(64)
        colors, rnds = s.download_colors(), s.download_rnds()
        release(s)
    finally:
        ptb.set_parity_filter(1)
    assert (bits(colors) == bits(c2_parity_64[0])).all() and (rnds == c2_parity_64[1]).all()


def test_c2_pixel_shards_tile_the_frame_bitwise(ptb, c2_parity_64):
    n, group = 4, 8 * W
    total = np.zeros((H, W, 4), dtype=np.float32)
    seen = np.zeros((H, W), dtype=np.int32)
    try:
        for rank in range(n):
            ptb.set_pixel_shard(rank, n, group)
            s = cornell(ptb, capi.ACCEL_BRUTE, capi.MATH_PARITY)
            s.render_spp(64)
            c = s.download_colors()
            release(s)
            own = ((np.arange(W * H) // group) % n == rank).reshape(H, W)
            assert (c[~own] == 0).all()        # a shard never writes a pixel it does not own
            total[own] = c[own]
            seen += own
    finally:
        ptb.set_pixel_shard(0, 1, 1)
    assert (seen == 1).all()
    assert (bits(total) == bits(c2_parity_64[0])).all()


def test_c2_sum_mode_and_resolve_agree_with_the_running_mean(ptb, c2_parity_64):
    try:
        ptb.set_accumulate_sum(1)
        s = cornell(ptb, capi.ACCEL_BRUTE, capi.MATH_PARITY)
        s.render_spp(64)
        ptb.resolve_sum(s.buffer_colors, 64)
        colors = s.download_colors()
        release(s)
    finally:
        ptb.set_accumulate_sum(0)
    st = util.compare(colors, c2_parity_64[0])
    assert st["max_rel"] <= 2e-5, st           # sum/64 vs 64 running-mean steps: a few ulp per step


def test_c2_full_1024spp_fast_flavour_agrees_with_parity_flavour(ptb):
    """the bench configuration itself, both flavours, whole frame: converged images agree to Monte-Carlo-free tolerance
    (identical random streams; only rounding differs, and a path that flips at an edge is 1/1024 of its pixel)"""
    s = cornell(ptb, capi.ACCEL_AUTO, capi.MATH_FAST)
    s.render_spp(1024)
    fast = s.download_colors().copy()
    n_prims = ptb.scan_prims()
    release(s)
    s = cornell(ptb, capi.ACCEL_BRUTE, capi.MATH_PARITY)
    s.render_spp(1024)
    par = s.download_colors().copy()
    release(s)
    assert n_prims == 16
    st = util.compare(fast, par)
    assert st["rmse_over_mean"] <= 0.01 and st["frac_rel_le_1e-4"] >= 0.99, st
    assert abs(float(fast[..., :3].mean()) / float(par[..., :3].mean()) - 1.0) <= 1e-4
    assert np.isfinite(fast[..., :3]).all()


def test_c3_lbvh_pipelines_agree_bitwise_at_1080p(ptb):
    """C3 (104 k triangles): no CPU structure to compare a whole frame with in test time — the three LBVH kernels (plain
    megakernel loop is the tested one at small size against the oracle, tests/test_gpu_parity.py) must agree bitwise"""
    d = scenes.describe("spheres100k")
    out = []
    for pipeline in (capi.PIPE_MEGAKERNEL, capi.PIPE_WAVEFRONT, capi.PIPE_POOL):
        ptb.set_accel(capi.ACCEL_LBVH); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(pipeline)
        s = host.Scene(ptb)
        s.build_trees = False
        s.init_Scene(W, H)
        scenes.populate(s, d)
        s.upload_Triangles(); s.upload_Materials()
        s.iterations = DEPTH
        cam = host.camera_at(W, H, d["fov"], d.get("yaw", 0), d.get("pitch", 0), d["eye"])
        s._camera = lambda cam=cam: cam
        s.render_spp(4)
        out.append((s.download_colors().copy(), s.download_rnds().copy()))
        release(s)
    for o in out[1:]:
        assert (bits(out[0][0]) == bits(o[0])).all() and (out[0][1] == o[1]).all()
    assert out[0][0][..., :3].mean() > 0


def _render_lbvh(ptb, d, spp, w=640, h=360):
    s = host.Scene(ptb)
    s.build_trees = False
    s.init_Scene(w, h)
    scenes.populate(s, d)
    s.upload_Triangles(); s.upload_Materials()
    s.iterations = DEPTH
    cam = host.camera_at(w, h, d["fov"], d.get("yaw", 0), d.get("pitch", 0), d["eye"])
    s._camera = lambda cam=cam: cam
    s.render_spp(spp)
    out = (s.download_colors().copy(), s.download_rnds().copy())
    release(s)
    return out


def test_wavefront_image_does_not_depend_on_the_number_of_sub_wavefronts(ptb):
    """The wavefront deals the frame's paths to 1..4 independent queues on separate streams (ptb_set_wf_subs); paths never
    move between them and a pixel's samples stay sequential, so colours and RNG states are the same bits for every count,
    in both flavours, at a frame size that is not a multiple of the 256-pixel dealing unit."""
    d = scenes.describe("spheres100k")
    try:
        for math in (capi.MATH_PARITY, capi.MATH_FAST):
            ptb.set_accel(capi.ACCEL_LBVH); ptb.set_math(math); ptb.set_pipeline(capi.PIPE_WAVEFRONT)
            outs = []
            for subs in (1, 2, 3, 4):
                ptb.set_wf_subs(subs)
                outs.append(_render_lbvh(ptb, d, 3, w=611, h=173))
            for c, r in outs[1:]:
                assert (bits(c) == bits(outs[0][0])).all() and (r == outs[0][1]).all()
            assert outs[0][0][..., :3].mean() > 0
    finally:
        ptb.set_wf_subs(2)
        ptb.set_math(capi.MATH_FAST)


def test_lbvh_arena_survives_bigger_and_smaller_scenes(ptb):
    """The LBVH lives in one arena that is kept across rebuilds and regrown only for more triangles (lbvh.cu): rendering
    the 104 k-triangle scene, then the 1 M-triangle one (regrow), then the 104 k one again (reuse of the larger arena, no
    allocation) must give the first image back bit for bit."""
    small, big = scenes.describe("spheres100k"), scenes.describe("grid1m")
    ptb.set_accel(capi.ACCEL_LBVH); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_WAVEFRONT)
    try:
        a = _render_lbvh(ptb, small, 2, w=320, h=180)
        b = _render_lbvh(ptb, big, 1, w=320, h=180)
        c = _render_lbvh(ptb, small, 2, w=320, h=180)
        assert (bits(a[0]) == bits(c[0])).all() and (a[1] == c[1]).all()
        assert b[0][..., :3].mean() > 0 and np.isfinite(b[0][..., :3]).all()
        assert ptb.accel_stats()["tris"] == len(small["tri_mat"])          # the tree in place is the small scene's again
    finally:
        ptb.set_math(capi.MATH_FAST)
