"""GPU parity of what FOLLOWS the bounce loop in the reference and of BASELINE's configs as stated (SURVEY §8 rows a16, f1, f2;
VERDICT r1 items 1.i-1.vi):

  * the tone-mapped texture trace_ray writes (`write_imagef(tex, ..., reinhard_tone_map(colors[id]))`, prog.cl:247-269, 380)
    against the reference's own texture, bit patterns (NaN where L == 0);
  * ptb_resolve_rgba8 / ptb_resolve_frame against the 8-bit quantisation of that same texture;
  * filt_im (prog.cl:391-427) against the reference's kernel run by the oracle executor, ties included;
  * C1 exactly as BASELINE states it: Cornell 512x512, 16 spp, depth 4, full frame against the reference kernels;
  * C4/C5's scene (1 M triangles, 3840x2160, its own camera) against the reference's brute force on six scanlines
    (fixture tests/golden/grid1m_4k_rows_spp4_d4.*, generator next to it).

Tolerances: parity flavour = identical bits.  Fast flavour = >= 99.9 % of pixels within 1e-4 relative and RMSE <= 1 % of the mean.
"""
import json
import os

import numpy as np
import pytest

from opencl_path_tracer_b200 import capi, host, scenes
from tests import util

pytestmark = pytest.mark.gpu
G = util.GOLDEN


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def same_texels(a, b):
    """bit patterns equal; where the reference's texture holds a NaN (L == 0: 0 * 0 / 0, prog.cl:264-269) the other must hold a NaN
    too — WHICH NaN is not a property of the algorithm (x86 produces 0xffc00000 for 0/0, the GPU 0x7fffffff)"""
    a, b = np.ascontiguousarray(a, dtype=np.float32), np.ascontiguousarray(b, dtype=np.float32)
    return bool(((a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))).all())


def quantise(tex):
    """the 8-bit picture of a float RGBA texture as ptb200.h defines it: NaN and negatives -> 0, >= 1 -> 255,
    else (unsigned char)(v * 255 + 0.5) in float arithmetic; rows flipped (row 0 of the reference image is the bottom)"""
    t = np.asarray(tex, dtype=np.float32)[..., :3]
    v = (t * np.float32(255.0) + np.float32(0.5)).astype(np.float32)
    with np.errstate(invalid="ignore"):
        q = np.where(~(t > 0), 0, np.where(t >= 1, 255, np.floor(np.nan_to_num(v, nan=0.0, posinf=0.0, neginf=0.0)))).astype(np.uint8)
    out = np.full(t.shape[:2] + (4,), 255, dtype=np.uint8)
    out[..., :3] = q
    return out[::-1]


def release(s):
    p = s.ptb
    p.finish()
    for h in (s.buffer_rays, s.buffer_rnds, s.cl_screen, s.buffer_colors, s.buffer_tris, s.buffer_mats, s.buffer_kd_tree, s.buffer_kd_tree_shift):
        if h:
            p.release(h)


# ---------------------------------------------------------------- a16 / f1: the tone-mapped image
@pytest.fixture(scope="module")
def cornell_golden():
    surf = host.load_ptbd(os.path.join(G, "cornell_96x64.ptbd"))
    tex = np.fromfile(os.path.join(G, "cornell_96x64_spp3_d4.tex.f32"), dtype=np.float32).reshape(64, 96, 4)
    colors = np.fromfile(os.path.join(G, "cornell_96x64_spp3_d4.colors.f32"), dtype=np.float32).reshape(64, 96, 4)
    assert np.isnan(tex[..., 0]).sum() > 1000           # the fixture does hold the reference's L == 0 -> NaN pixels
    return surf, colors, tex


@pytest.mark.parametrize("accel,pipeline,mode", [
    (capi.ACCEL_BRUTE, capi.PIPE_MEGAKERNEL, "render_spp"), (capi.ACCEL_BRUTE, capi.PIPE_MEGAKERNEL, "launches"),
    (capi.ACCEL_REFBVH, capi.PIPE_MEGAKERNEL, "render_spp"), (capi.ACCEL_LBVH, capi.PIPE_MEGAKERNEL, "render_spp"),
    (capi.ACCEL_LBVH, capi.PIPE_WAVEFRONT, "render_spp"), (capi.ACCEL_LBVH, capi.PIPE_POOL, "render_spp"), (capi.ACCEL_LBVH, capi.PIPE_POOL, "launches")])
def test_texture_written_by_the_path_kernels_is_the_reference_texture(ptb, cornell_golden, accel, pipeline, mode):
    surf, colors, tex = cornell_golden
    ptb.set_accel(accel); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(pipeline)
    s = host.scene_from_surface(surf, ptb)
    if mode == "launches":
        for _ in range(3):
            s.render()                                  # gen_ray + trace_ray per sample: the texture is rewritten every sample (prog.cl:380)
    else:
        s.render_spp(3)
    ptb.finish()
    got = s.download_image()
    got_colors = s.download_colors()
    release(s)
    assert (bits(got_colors[..., :3]) == bits(colors[..., :3])).all()
    assert same_texels(got, tex)                        # bit patterns; NaN exactly where the reference has NaN (the L == 0 pixels)
    assert (np.isnan(got[..., 0]) == np.isnan(tex[..., 0])).all()


def test_resolve_rgba8_is_the_quantised_reference_texture(ptb, cornell_golden):
    surf, colors, tex = cornell_golden
    ptb.set_accel(capi.ACCEL_BRUTE); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_MEGAKERNEL)
    s = host.scene_from_surface(surf, ptb)
    s.render_spp(3)
    rgba = ptb.resolve_rgba8(s.buffer_colors, 96, 64, 0)
    release(s)
    want = quantise(tex)
    assert rgba.shape == want.shape and (rgba == want).all()
    assert (rgba[::-1][np.isnan(tex[..., 0])][:, :3] == 0).all()       # the reference's NaN pixels are black, alpha stays 255


def test_fused_root_resolve_sum_to_mean_to_texture_to_rgba8(ptb, cornell_golden):
    """ptb_resolve_frame = the spp-split root's epilogue in one pass: colors holds a SUM of samples on entry (sum-mode
    accumulation); afterwards colors is the mean, the image is reinhard_tone_map(mean) and the RGBA8 is its quantisation."""
    surf, _, _ = cornell_golden
    ptb.set_accel(capi.ACCEL_BRUTE); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_MEGAKERNEL)
    try:
        ptb.set_accumulate_sum(1)
        s = host.scene_from_surface(surf, ptb)
        s.render_spp(5)
        summed = s.download_colors().copy()
        rgba = ptb.resolve_frame(s.buffer_colors, 96, 64, divide_by=5, mode=0, image=s.cl_screen)
        mean, tex = s.download_colors().copy(), s.download_image().copy()
        filmic8 = ptb.resolve_frame(s.buffer_colors, 96, 64, divide_by=1, mode=1, image=0)
        release(s)
    finally:
        ptb.set_accumulate_sum(0)
    want_mean = (summed[..., :3] * np.float32(1.0 / 5.0)).astype(np.float32)
    assert (bits(mean[..., :3]) == bits(want_mean)).all()
    from oracle import pt_oracle_py as po
    fn = po.lib().pto_reinhard
    want_tex = np.zeros((64, 96, 4), dtype=np.float32)
    flat_mean, flat_tex = np.ascontiguousarray(mean[..., :3]).reshape(-1, 3), want_tex.reshape(-1, 4)
    for i in range(flat_mean.shape[0]):
        fn(flat_mean[i].ctypes.data, flat_tex[i].ctypes.data)
    assert same_texels(tex, want_tex)
    assert (rgba == quantise(want_tex)).all()
    c = np.maximum(np.float32(0), mean[..., :3] - np.float32(0.004))
    fil = (c * (c * np.float32(6.2) + np.float32(0.5))) / (c * (c * np.float32(6.2) + np.float32(1.7)) + np.float32(0.06))
    assert (filmic8 == quantise(np.concatenate([fil, np.ones((64, 96, 1), np.float32)], axis=-1))).all()


# ---------------------------------------------------------------- f2: filt_im
def _tie_image(w, h, seed=11):
    """random radiance with what makes the reference's unstable selection sort matter: runs of pixels with EQUAL grey
    value and different colours, flat black areas, repeated pixels"""
    rng = np.random.default_rng(seed)
    img = np.zeros((h, w, 4), dtype=np.float32)
    img[..., :3] = rng.random((h, w, 3), dtype=np.float32) * 2.0
    ties = np.array([(1, 2, 3), (3, 2, 1), (2, 2, 2), (0, 3, 3), (6, 0, 0), (2, 1, 3)], dtype=np.float32)
    img[4:14, 5:25, :3] = ties[rng.integers(0, len(ties), size=(10, 20))]
    img[16:24, 10:30, :3] = 0.0
    img[20:30, 28:44, :3] = ties[rng.integers(0, 3, size=(10, 16))]
    img[26:, :8, :3] = img[26, 0, :3]
    return img


def test_filt_im_golden_fixture(ptb):
    """the reference's own filt_im (executor: oracle/ref_backend.cpp) on a 48x32 image with constructed ties — committed fixture"""
    src = np.fromfile(os.path.join(G, "filt_im_48x32.in.f32"), dtype=np.float32).reshape(32, 48, 4)
    want = np.fromfile(os.path.join(G, "filt_im_48x32.out.f32"), dtype=np.float32).reshape(32, 48, 4)
    hc, hi = ptb.upload(src), ptb.alloc_image(48, 32)
    ptb.filter_median(hc, hi, 48, 32)
    got = ptb.read(hi, np.float32, 48 * 32 * 4).reshape(32, 48, 4)
    ptb.release(hc); ptb.release(hi)
    assert (bits(got) == bits(want)).all()
    # what the kernel writes and what it leaves: rows 1..H-2, columns 1..W-1, minus the last pixel of row H-2 (its
    # neighbourhood would run one element past the buffer: undefined in the reference, untouched here)
    assert (want[1:-1, 1:, 3].reshape(-1)[:-1] == 1).all() and (want[-2, -1] == 0).all() and (want[0] == 0).all() and (want[-1] == 0).all()


@pytest.mark.parametrize("w,h", [(48, 32), (33, 9), (130, 67), (3, 3), (2, 5), (1920, 1080)])
def test_filt_im_matches_the_oracle(ptb, w, h):
    """all pixels (written and untouched) against the C restatement of prog.cl:391-427, itself pinned to the reference's
    kernel by tests/test_oracle.py; sizes that are not multiples of the 32x8 tile, a degenerate 3x3 and a full 1080p frame"""
    from oracle import pt_oracle_py as po
    src = _tie_image(max(w, 48), max(h, 32))[:h, :w].copy() if w < 1000 else None
    if src is None:
        rng = np.random.default_rng(5)
        src = np.zeros((h, w, 4), dtype=np.float32)
        src[..., :3] = np.round(rng.random((h, w, 3), dtype=np.float32) * 8.0) / 4.0     # coarse values: thousands of grey ties
    want = po.filt_im(src, w, h)
    hc, hi = ptb.upload(src), ptb.alloc_image(w, h)
    ptb.filter_median(hc, hi, w, h)
    got = ptb.read(hi, np.float32, w * h * 4).reshape(h, w, 4)
    # the kernel id route of the reference's commented launch (main.cpp:665-668): filt_im(tex, colors) over (W, H)
    hi2 = ptb.alloc_image(w, h)
    ptb.set_arg_mem(capi.KERNEL_FILT_IM, 0, hi2); ptb.set_arg_mem(capi.KERNEL_FILT_IM, 1, hc)
    ptb.launch(capi.KERNEL_FILT_IM, w, h)
    got2 = ptb.read(hi2, np.float32, w * h * 4).reshape(h, w, 4)
    for m in (hc, hi, hi2):
        ptb.release(m)
    assert (bits(got) == bits(want)).all()
    assert (bits(got2) == bits(want)).all()


def test_filt_im_on_a_rendered_frame_matches_the_reference_kernel(ptb, tmp_path):
    if not util.have_ref_cpu():
        pytest.skip("oracle/_ref/ref_cpu not built")
    out = str(tmp_path / "filt.bin")
    ref = util.run_ref_cpu(128, 96, 4, 4, obj=util.CORNELL, eye=(277.5, 277.5, -800), fov=39.3, want=("colors", "dump"), extra=("--filt-out", out))
    want = np.fromfile(out, dtype=np.float32).reshape(96, 128, 4)
    ptb.set_accel(capi.ACCEL_AUTO); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_AUTO)
    s = host.scene_from_surface(host.load_ptbd(ref["files"]["dump"]), ptb)
    s.render_spp(4)
    ptb.write(s.cl_screen, np.zeros(128 * 96 * 4, dtype=np.float32))
    ptb.filter_median(s.buffer_colors, s.cl_screen, 128, 96)
    got = s.download_image()
    release(s)
    assert (bits(got) == bits(want)).all()


# ---------------------------------------------------------------- C1 as BASELINE states it
@pytest.fixture(scope="module")
def c1_ref():
    if not util.have_ref_cpu():
        pytest.skip("oracle/_ref/ref_cpu not built")
    return util.run_ref_cpu(512, 512, 16, 4, obj=util.CORNELL, eye=(277.5, 277.5, -800), fov=39.3, want=("colors", "rnds", "dump", "tex"))


@pytest.mark.parametrize("accel,pipeline", [(capi.ACCEL_AUTO, capi.PIPE_AUTO), (capi.ACCEL_LBVH, capi.PIPE_WAVEFRONT), (capi.ACCEL_LBVH, capi.PIPE_POOL),
                                            (capi.ACCEL_REFBVH, capi.PIPE_MEGAKERNEL)])
def test_c1_cornell_512_16spp_depth4_full_frame_is_bitwise_the_reference(ptb, c1_ref, accel, pipeline):
    ptb.set_accel(accel); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(pipeline)
    s = host.scene_from_surface(host.load_ptbd(c1_ref["files"]["dump"]), ptb)
    s.render_spp(16)
    colors, rnds, tex = s.download_colors(), s.download_rnds(), s.download_image()
    release(s)
    assert (bits(colors[..., :3]) == bits(c1_ref["colors"][..., :3])).all()
    assert (rnds == c1_ref["rnds"]).all()
    assert same_texels(tex, c1_ref["tex"])


def test_c1_fast_flavour_within_the_stated_tolerance(ptb, c1_ref):
    ptb.set_accel(capi.ACCEL_AUTO); ptb.set_math(capi.MATH_FAST); ptb.set_pipeline(capi.PIPE_AUTO)
    s = host.scene_from_surface(host.load_ptbd(c1_ref["files"]["dump"]), ptb)
    s.render_spp(16)
    colors = s.download_colors()
    release(s)
    st = util.compare(colors, c1_ref["colors"])
    assert st["frac_rel_le_1e-4"] >= 0.999 and st["rmse_over_mean"] <= 0.01, st


# ---------------------------------------------------------------- C4 / C5 scene against the reference's brute force
@pytest.fixture(scope="module")
def grid1m_rows():
    meta = json.load(open(os.path.join(G, "grid1m_4k_rows_spp4_d4.json")))
    rows = np.asarray(meta["rows"])
    W = meta["w"]
    colors = np.fromfile(os.path.join(G, "grid1m_4k_rows_spp4_d4.colors.f32"), dtype=np.float32).reshape(len(rows), W, 4)
    rnds = np.fromfile(os.path.join(G, "grid1m_4k_rows_spp4_d4.rnds.i32"), dtype=np.int32).reshape(len(rows), W)
    assert len(rows) * W >= 20000 and meta["spp"] >= 4 and meta["tris"] >= 1000000
    return meta, rows, colors, rnds


def _render_grid1m(ptb, meta, math, pipeline):
    d = scenes.describe("grid1m")
    assert [float(v) for v in d["eye"]] == meta["eye"] and d["fov"] == meta["fov"] and d["pitch"] == meta["pitch"]
    ptb.set_accel(capi.ACCEL_LBVH); ptb.set_math(math); ptb.set_pipeline(pipeline)
    s = host.Scene(ptb)
    s.build_trees = False
    W, H = meta["w"], meta["h"]
    s.init_Scene(W, H)
    scenes.populate(s, d)
    assert len(s.tris) + sum(len(p) for p in s._pending) == meta["tris"]
    s.upload_Triangles(); s.upload_Materials()
    s.iterations = meta["iterations"]
    cam = host.camera_at(W, H, d["fov"], d["yaw"], d["pitch"], d["eye"])
    s._camera = lambda cam=cam: cam
    s.render_spp(meta["spp"])
    colors, rnds = s.download_colors(), s.download_rnds().reshape(H, W)
    release(s)
    return colors, rnds


@pytest.mark.parametrize("pipeline", [capi.PIPE_MEGAKERNEL, capi.PIPE_WAVEFRONT, capi.PIPE_POOL])
def test_grid1m_4k_rows_are_bitwise_the_reference_brute_force(ptb, grid1m_rows, pipeline):
    """the full 3840x2160 frame of the 1 M-triangle scene, 4 spp, depth 4, parity flavour through the LBVH; its scanlines
    0, 360, ... 1800 (23 040 pixels) against the reference's own prog.cl testing EVERY triangle for every ray"""
    meta, rows, want_colors, want_rnds = grid1m_rows
    colors, rnds = _render_grid1m(ptb, meta, capi.MATH_PARITY, pipeline)
    assert (bits(colors[rows][..., :3]) == bits(want_colors[..., :3])).all()
    assert (rnds[rows] == want_rnds).all()
    assert np.isfinite(colors[..., :3]).all() and colors[..., :3].mean() > 0


def test_grid1m_4k_rows_fast_flavour_within_the_stated_tolerance(ptb, grid1m_rows):
    meta, rows, want_colors, want_rnds = grid1m_rows
    colors, rnds = _render_grid1m(ptb, meta, capi.MATH_FAST, capi.PIPE_AUTO)
    st = util.compare(colors[rows], want_colors)
    assert st["frac_rel_le_1e-4"] >= 0.999 and st["rmse_over_mean"] <= 0.01, st
    assert (rnds[rows] == want_rnds).mean() >= 0.999


# ---------------------------------------------------------------- checkpoint / resume
def test_checkpoint_resume_is_bitwise_an_uninterrupted_render(ptb, cornell_golden, tmp_path):
    """8 + 8 samples across ptb_save_state / a fresh set of buffers / ptb_load_state == 16 samples in one go (colours and RNG states)"""
    surf, _, _ = cornell_golden
    ptb.set_accel(capi.ACCEL_LBVH); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_AUTO)
    s = host.scene_from_surface(surf, ptb)
    s.render_spp(16)
    want_c, want_r = s.download_colors().copy(), s.download_rnds().copy()
    release(s)
    path = str(tmp_path / "state.ptbc")
    s = host.scene_from_surface(surf, ptb)
    s.render_spp(8)
    s._set_trace_args()                                  # arg 10 = 8: the index the next sample will have
    ptb.save_state(path)
    release(s)
    s = host.scene_from_surface(surf, ptb)               # fresh buffers: seeds back at the start, colours zero
    s.camera = s._camera()
    s._set_trace_args()
    s.current_sample = ptb.load_state(path)
    assert s.current_sample == 8
    s.render_spp(8)
    got_c, got_r = s.download_colors(), s.download_rnds()
    release(s)
    assert (bits(got_c) == bits(want_c)).all() and (got_r == want_r).all()
    with pytest.raises(capi.PtbError):
        ptb.load_state(str(tmp_path / "missing.ptbc"))
    ptb.clear_error()
