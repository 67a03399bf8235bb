"""Edge cases of the path, CUDA (through the C-ABI) against the C oracle (oracle/pt_oracle.c, brute-force closest hit
as prog.cl:318): empty and tiny scenes, ragged launch sizes, duplicate and degenerate triangles, zero bounces, every
closest-hit structure and both pipelines.  PARITY flavour: bit-identical colours and RNG states."""
import numpy as np
import pytest

from opencl_path_tracer_b200 import capi, host
from oracle import pt_oracle_py as po

pytestmark = pytest.mark.gpu

DIFFUSE = host.Material((0.7, 0.6, 0.5), (0, 0, 0), (0, 0, 0), (0, 0, 0), (0, 0, 0), 0.0, 0)
GLOSSY = host.Material((0.4, 0.4, 0.6), (0.5, 0.5, 0.5), (0, 0, 0), (0, 0, 0), (0, 0, 0), 20.0, 0)
LIGHT = host.Material((0, 0, 0), (0, 0, 0), (9, 8, 7), (0, 0, 0), (0, 0, 0), 0.0, 3)
MIRROR = host.Material((0, 0, 0), (0, 0, 0), (0, 0, 0), (0.2, 0.9, 1.1), (3.0, 2.4, 1.8), 0.0, 1)
GLASS = host.Material((0, 0, 0), (0, 0, 0), (0, 0, 0), (1.5, 1.5, 1.5), (0, 0, 0), 0.0, 2)
MATS = [DIFFUSE, GLOSSY, LIGHT, MIRROR, GLASS]


def quad(a, b, c, d, m):
    return [(a, b, c, m), (a, c, d, m)]


def open_box(extra=()):
    """floor, back wall, ceiling light and whatever `extra` adds, seen from z = -300"""
    t = []
    t += quad((-100, 0, -100), (100, 0, -100), (100, 0, 100), (-100, 0, 100), 0)
    t += quad((-100, 0, 100), (100, 0, 100), (100, 200, 100), (-100, 200, 100), 1)
    t += quad((-60, 199, -60), (60, 199, -60), (60, 199, 60), (-60, 199, 60), 2)
    t += list(extra)
    return t


def build(tris, W, H, depth):
    s = host.Scene(ptb=object())
    for m in MATS:
        s.add_Material(m)
    if tris:
        r1, r2, r3 = (np.array([t[k] for t in tris], dtype=np.float32) for k in range(3))
        s.add_Triangle(host.Triangles(r1, r2, r3, np.array([t[3] for t in tris])))
        s.end_Obj()
    s._flush()
    return {"W": W, "H": H, "iterations": depth, "camera": host.camera_at(W, H, 50.0, 0.0, 0.0, (0.0, 90.0, -300.0)),
            "tris": s.tris, "mats": np.stack(s.mats), "kd_tree": host.nodes_to_records(s.kd_tree),
            "kd_tree_shift": np.asarray(s.kd_tree_shift, dtype=np.int32)}


def both(ptb, surf, spp, accel, pipeline=capi.PIPE_MEGAKERNEL, math=capi.MATH_PARITY):
    # the reference-tree mode is checked against the oracle's own walk of those trees (prog.cl:144-184: on EXACT ties the
    # visit order decides, not the index); brute force and the LBVH against first_intersect over the whole array
    orc = po.OracleScene(surf, brute=accel != capi.ACCEL_REFBVH)
    want = orc.render(spp, threads=4).copy()
    ptb.set_accel(accel); ptb.set_math(math); ptb.set_pipeline(pipeline)
    s = host.scene_from_surface(surf, ptb)
    s.render_spp(spp)
    ptb.finish()
    got, rnds = s.download_colors(), s.download_rnds()
    p = s.ptb
    for h in (s.buffer_rays, s.buffer_rnds, s.cl_screen, s.buffer_colors, s.buffer_tris, s.buffer_mats, s.buffer_kd_tree, s.buffer_kd_tree_shift):
        if h:
            p.release(h)
    return got, rnds, want.reshape(got.shape), orc.rnds


def same_bits(a, b):
    a, b = np.ascontiguousarray(a[..., :3], dtype=np.float32), np.ascontiguousarray(b[..., :3], dtype=np.float32)
    return bool(((a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))).all())


ALL_ACCELS = [capi.ACCEL_BRUTE, capi.ACCEL_REFBVH, capi.ACCEL_LBVH]


@pytest.mark.parametrize("accel", ALL_ACCELS)
@pytest.mark.parametrize("pipeline", [capi.PIPE_MEGAKERNEL, capi.PIPE_WAVEFRONT, capi.PIPE_POOL])
def test_empty_scene_is_black_and_still_draws_the_camera_jitter(ptb, accel, pipeline):
    got, rnds, want, wrnds = both(ptb, build([], 37, 19, 4), 3, accel, pipeline)
    assert (got[..., :3] == 0).all() and same_bits(got, want)
    assert (rnds == wrnds).all() and (rnds != host.minstd_rand0_seeds(37 * 19)).all()     # two draws per sample (prog.cl:388)


@pytest.mark.parametrize("accel", ALL_ACCELS)
@pytest.mark.parametrize("n_tris", [1, 2, 3, 5])
def test_tiny_scenes_single_leaf_and_single_node_trees(ptb, accel, n_tris):
    tris = open_box()[:n_tris]
    got, rnds, want, wrnds = both(ptb, build(tris, 48, 40, 3), 4, accel)
    assert same_bits(got, want) and (rnds == wrnds).all()


@pytest.mark.parametrize("size", [(1, 1), (7, 3), (33, 17), (130, 9), (17, 129)])
@pytest.mark.parametrize("pipeline", [capi.PIPE_MEGAKERNEL, capi.PIPE_WAVEFRONT, capi.PIPE_POOL])
def test_ragged_launch_sizes(ptb, size, pipeline):
    """widths/heights that are no multiple of the 16x8 block or the 8x4 warp tile, down to one pixel"""
    W, H = size
    accel = capi.ACCEL_LBVH if pipeline != capi.PIPE_MEGAKERNEL else capi.ACCEL_BRUTE
    got, rnds, want, wrnds = both(ptb, build(open_box(), W, H, 4), 5, accel, pipeline)
    assert same_bits(got, want) and (rnds == wrnds).all()
    assert got[..., :3].sum() > 0 or W * H < 16


@pytest.mark.parametrize("accel", ALL_ACCELS)
def test_duplicate_triangles_tie_on_the_lower_upload_index(ptb, accel):
    """exactly coincident triangles with DIFFERENT materials: first_intersect keeps the first (strict '<', prog.cl:118),
    so the image must be that of the lower index — the tie rule every tree has to reproduce"""
    base = open_box()
    dup = [(t[0], t[1], t[2], 3) for t in base[:4]]          # mirror-material copies of floor and back wall, uploaded later
    got, rnds, want, wrnds = both(ptb, build(base + dup, 64, 48, 4), 4, accel)
    assert same_bits(got, want) and (rnds == wrnds).all()
    if accel != capi.ACCEL_REFBVH:
        got2, _, _, _ = both(ptb, build(base, 64, 48, 4), 4, accel)
        assert same_bits(got, got2)                            # the later duplicates are never seen


@pytest.mark.parametrize("accel", [capi.ACCEL_BRUTE, capi.ACCEL_LBVH])
def test_degenerate_triangles_do_not_poison_the_image(ptb, accel):
    """zero-area triangles carry a NaN normal (main.cpp:158-161 divides by a zero length): the reference test rejects them
    (every comparison with NaN is false) and so must the branch-free CUDA test and the LBVH built over them"""
    junk = [((5, 5, 5), (5, 5, 5), (5, 5, 5), 0), ((0, 50, 0), (10, 50, 0), (20, 50, 0), 1), ((-30, 20, 10), (-30, 20, 10), (40, 90, 10), 2)]
    got, rnds, want, wrnds = both(ptb, build(open_box(junk), 64, 48, 4), 4, accel)
    assert same_bits(got, want) and (rnds == wrnds).all()
    assert np.isfinite(got[..., :3]).all()


@pytest.mark.parametrize("pipeline", [capi.PIPE_MEGAKERNEL, capi.PIPE_WAVEFRONT, capi.PIPE_POOL])
@pytest.mark.parametrize("depth", [0, 1, 2, 7])
def test_bounce_count_extremes(ptb, depth, pipeline):
    """iterations = 0: the bounce loop never runs (black, only the jitter draws); 1: flat preview (prog.cl:323-325)"""
    got, rnds, want, wrnds = both(ptb, build(open_box(), 40, 30, depth), 3, capi.ACCEL_LBVH, pipeline)
    assert same_bits(got, want) and (rnds == wrnds).all()
    if depth == 0:
        assert (got[..., :3] == 0).all()


@pytest.mark.parametrize("pipeline", [capi.PIPE_MEGAKERNEL, capi.PIPE_WAVEFRONT, capi.PIPE_POOL])
def test_mirror_and_glass_inside_the_box(ptb, pipeline):
    extra = quad((-80, 0, 40), (-20, 0, 40), (-20, 90, 60), (-80, 90, 60), 3) + quad((20, 10, -20), (80, 10, -20), (80, 80, 0), (20, 80, 0), 4) \
        + quad((20, 10, 30), (80, 10, 30), (80, 80, 50), (20, 80, 50), 4)
    got, rnds, want, wrnds = both(ptb, build(open_box(extra), 72, 54, 6), 6, capi.ACCEL_LBVH, pipeline)
    assert same_bits(got, want) and (rnds == wrnds).all()


def test_fast_flavour_on_the_same_edge_scenes(ptb):
    """the fast flavour (merged quads, folded camera, FMA) on the scenes above: stated tolerance, no NaN/inf leaking in"""
    from tests import util
    for tris, depth in ((open_box(), 4), (open_box()[:1], 3), ([], 2)):
        got, rnds, want, wrnds = both(ptb, build(tris, 64, 48, depth), 8, capi.ACCEL_AUTO, math=capi.MATH_FAST)
        assert np.isfinite(got[..., :3]).all()
        if tris:
            st = util.compare(got, want)
            assert st["frac_rel_le_1e-4"] >= 0.995 and st["rmse_over_mean"] <= 0.02, st
        else:
            assert (got[..., :3] == 0).all() and (rnds == wrnds).all()


def soup(seed, n):
    """a hostile triangle soup around the open box: needles (aspect 1e4), specks, slivers through the camera position,
    triangles in planes that contain the eye (N.V ~ 0 for whole fans of rays) and far-away giants"""
    rng = np.random.default_rng(seed)
    t = open_box()
    eye = np.array([0.0, 90.0, -300.0])
    for k in range(n):
        kind = k % 6
        c = rng.uniform([-90, 10, -80], [90, 190, 90])
        if kind == 0:                                  # needle
            d = rng.normal(size=3); d /= np.linalg.norm(d)
            e = np.cross(d, rng.normal(size=3)); e /= np.linalg.norm(e)
            a, b, cc = c, c + 80 * d, c + 80 * d + 0.008 * e
        elif kind == 1:                                # speck
            a, b, cc = c, c + rng.normal(size=3) * 1e-3, c + rng.normal(size=3) * 1e-3
        elif kind == 2:                                # plane through the eye: every ray towards it grazes
            a, b, cc = eye + (c - eye) * 0.5, eye + (c - eye) * 1.5, eye + (c - eye) * 1.0 + rng.normal(size=3) * 20
        elif kind == 3:                                # giant, far away
            a, b, cc = c * 40 + 3000, c * 40 + 3000 + rng.normal(size=3) * 4000, c * 40 + 3000 + rng.normal(size=3) * 4000
        elif kind == 4:                                # ordinary
            a, b, cc = c, c + rng.normal(size=3) * 30, c + rng.normal(size=3) * 30
        else:                                          # sliver sharing an edge with the previous triangle
            a, b, cc = np.array(t[-1][0]), np.array(t[-1][1]), np.array(t[-1][1]) + rng.normal(size=3) * 1e-2
        t.append((tuple(map(float, a)), tuple(map(float, b)), tuple(map(float, cc)), int(rng.integers(0, 5))))
    return t


@pytest.mark.parametrize("seed,n", [(1, 40), (2, 120), (3, 300)])
def test_parity_prefilter_on_a_hostile_soup(ptb, seed, n):
    """the pre-filter of the parity brute-force scan must never drop a triangle the exact test accepts: oracle, filter on
    and filter off agree bit for bit on needles, specks, grazing planes and far giants"""
    surf = build(soup(seed, n), 96, 64, 5)
    try:
        ptb.set_parity_filter(1)
        on, rnds_on, want, wrnds = both(ptb, surf, 6, capi.ACCEL_BRUTE)
        ptb.set_parity_filter(0)
        off, rnds_off, _, _ = both(ptb, surf, 6, capi.ACCEL_BRUTE)
    finally:
        ptb.set_parity_filter(1)
    assert same_bits(on, off) and (rnds_on == rnds_off).all()
    assert same_bits(on, want) and (rnds_on == wrnds).all()


def test_material_index_past_the_buffer_is_an_error_not_a_stray_read(ptb):
    """prog.cl:322 indexes materials[] unchecked; the C-ABI refuses the launch and says why"""
    tris = open_box() + [((0, 0, 0), (10, 0, 0), (0, 10, 0), len(MATS) + 3)]
    surf = build(tris, 16, 8, 2)
    ptb.set_accel(capi.ACCEL_BRUTE); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_MEGAKERNEL)
    s = host.scene_from_surface(surf, ptb)
    try:
        with pytest.raises(RuntimeError, match="material"):
            s.render_spp(1)
    finally:
        ptb.clear_error()
    ptb.finish()


def test_read_write_range_checks_do_not_wrap(ptb):
    h = ptb.alloc(64)
    try:
        for call in (lambda: ptb.lib.ptb_write(h, 2**64 - 8, 16, np.zeros(16, np.uint8).ctypes.data),
                     lambda: ptb.lib.ptb_read(h, 2**64 - 8, 16, np.zeros(16, np.uint8).ctypes.data)):
            assert call() == -30                       # PTB_ERR_INVALID_VALUE
            ptb.clear_error()
    finally:
        ptb.clear_error()
        ptb.release(h)
