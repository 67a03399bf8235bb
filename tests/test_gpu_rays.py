"""Closest hit on arbitrary rays (ptb_closest_hit): every structure and pipeline must return exactly what brute force over
the uploaded array returns — the same t bits and the same triangle index — on millions of random rays, including rays
aimed at shared edges/vertices and at exactly duplicated triangles.  A sample of the rays is also
checked against the C oracle's first_intersect (oracle/pt_oracle.c)."""
import numpy as np
import pytest

from opencl_path_tracer_b200 import capi, host, scenes

pytestmark = pytest.mark.gpu


def bind(ptb, d_or_tris, W=64, H=64, dup=False, build_trees=True):
    s = host.Scene(ptb)
    s.build_trees = build_trees
    s.init_Scene(W, H)
    if isinstance(d_or_tris, dict):
        scenes.populate(s, d_or_tris)
    else:
        for m in d_or_tris[1]:
            s.add_Material(m)
        s.add_Triangle(d_or_tris[0])
        s.end_Obj()
    s.n_before_dup = None
    if dup:                                       # every 5th triangle once more, uploaded later, other material
        s._flush()
        s.n_before_dup = len(s.tris)
        extra = s.tris[::5].copy()
        extra["mati"] = (extra["mati"] + 1) % max(len(s.mats), 1)
        s.add_Triangle(extra)
        s.end_Obj()
    s.upload_Triangles(); s.upload_Materials()
    s.iterations = 1
    cam = host.camera_at(W, H, 60.0, 0, 0, (0.0, 0.0, -10.0))
    s._camera = lambda: cam
    s.camera = cam
    s._set_trace_args()                           # the scene is bound through trace_ray's arguments
    return s


def random_rays(rng, n, lo, hi, tris=None):
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    o = (lo + (hi - lo) * rng.random((n, 3), dtype=np.float32)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    if tris is not None:                          # half of the rays aimed at vertices, edge midpoints and centroids
        k = n // 2
        t = tris[rng.integers(0, len(tris), k)]
        a, b, c = t["r1"][:, :3], t["r2"][:, :3], t["r3"][:, :3]
        pick = rng.integers(0, 3, k)[:, None]
        target = np.where(pick == 0, a, np.where(pick == 1, (a + b) * np.float32(0.5), (a + b + c) / np.float32(3.0)))
        d[:k] = target - o[:k]
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    d[::97, 0] = 0.0                              # axis-parallel components: the slab test's 0 * inf corner
    d[::89, 1] = 0.0
    nz = np.linalg.norm(d, axis=1) > 0
    d[~nz] = (0, 0, 1)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    return o, d.astype(np.float32)


def same(a, b):
    return bool((a[0].view(np.uint32) == b[0].view(np.uint32)).all() and (a[1] == b[1]).all())


@pytest.mark.parametrize("math", [capi.MATH_PARITY, capi.MATH_FAST])
@pytest.mark.parametrize("scene,dup", [("small_mixed", False), ("small_mixed", True), ("spheres100k", False)])
def test_every_structure_equals_brute_force_on_random_rays(ptb, math, scene, dup):
    d = scenes.describe(scene)
    ptb.set_math(math)
    s = bind(ptb, d, dup=dup, build_trees=scene != "spheres100k")
    s._flush()
    v = np.concatenate([s.tris["r1"][:, :3], s.tris["r2"][:, :3], s.tris["r3"][:, :3]])
    lo, hi = v.min(axis=0), v.max(axis=0)
    rng = np.random.default_rng(20240 + len(s.tris))
    n = 2_000_000 if scene == "small_mixed" else 300_000       # brute force over 104 k triangles is the slow side
    o, dirs = random_rays(rng, n, lo, hi, s.tris)
    ptb.set_accel(capi.ACCEL_BRUTE); ptb.set_pipeline(capi.PIPE_MEGAKERNEL)
    want = ptb.closest_hit(o, dirs)
    assert (want[1] >= 0).mean() > 0.5            # closed rooms: most rays hit something
    try:
        ptb.set_merge_prims(0)                    # (fast flavour: plain triangles first — same test code as brute force, so exact)
        for accel, pipe in ((capi.ACCEL_LBVH, capi.PIPE_MEGAKERNEL), (capi.ACCEL_LBVH, capi.PIPE_WAVEFRONT), (capi.ACCEL_LBVH, capi.PIPE_POOL)):
            ptb.set_accel(accel); ptb.set_pipeline(pipe)
            for quant in (1, 0):                  # 64-byte quantised wide nodes forced on, then the 128-byte float nodes (the default picks by tree size)
                ptb.set_quant_nodes(quant)
                assert same(ptb.closest_hit(o, dirs), want), (accel, pipe, quant)
    finally:
        ptb.set_merge_prims(1)
        ptb.set_quant_nodes(2)
    if math == capi.MATH_FAST:
        # shipping configuration: the outlier list is scanned as merged quads, a hit on either half reports the pair's lower
        # index and t comes from the shared plane — same surface, same material, t within rounding
        ptb.set_accel(capi.ACCEL_LBVH); ptb.set_pipeline(capi.PIPE_WAVEFRONT)
        t, idx = ptb.closest_hit(o, dirs)
        hit = want[1] >= 0
        assert ((idx >= 0) == hit).mean() >= 1.0 - 1e-5
        both = hit & (idx >= 0)
        assert (np.abs(t[both] - want[0][both]) <= 1e-4 * np.maximum(want[0][both], 1.0)).mean() >= 1.0 - 1e-5
        assert (s.tris["mati"][idx[both]] == s.tris["mati"][want[1][both]]).mean() >= 1.0 - 1e-4
        nrm = np.abs(s.tris["N"][idx[both]][:, :3] - s.tris["N"][want[1][both]][:, :3]).max(axis=1)
        assert (nrm <= 1e-5).mean() >= 1.0 - 1e-4
    if dup:                                       # exact ties: the later copies never win (first_intersect, prog.cl:118)
        n_orig = s.n_before_dup
        assert (want[1] < n_orig).all()


def test_reference_trees_equal_brute_force_where_no_exact_tie(ptb):
    d = scenes.describe("small_mixed")
    ptb.set_math(capi.MATH_PARITY)
    s = bind(ptb, d)
    s._flush()
    v = np.concatenate([s.tris["r1"][:, :3], s.tris["r2"][:, :3], s.tris["r3"][:, :3]])
    rng = np.random.default_rng(7)
    o, dirs = random_rays(rng, 500_000, v.min(axis=0), v.max(axis=0))
    ptb.set_accel(capi.ACCEL_BRUTE); ptb.set_pipeline(capi.PIPE_MEGAKERNEL)
    want = ptb.closest_hit(o, dirs)
    ptb.set_accel(capi.ACCEL_REFBVH)
    got = ptb.closest_hit(o, dirs)
    assert (got[0].view(np.uint32) == want[0].view(np.uint32)).all()       # same t everywhere
    assert (got[1] != want[1]).mean() <= 1e-4                              # the index may differ only on exact ties (visit order)


def test_sample_of_the_rays_against_the_c_oracle(ptb):
    from oracle import pt_oracle_py as po
    d = scenes.describe("small_mixed")
    ptb.set_math(capi.MATH_PARITY); ptb.set_accel(capi.ACCEL_LBVH); ptb.set_pipeline(capi.PIPE_WAVEFRONT)
    s = bind(ptb, d)
    s._flush()
    v = np.concatenate([s.tris["r1"][:, :3], s.tris["r2"][:, :3], s.tris["r3"][:, :3]])
    rng = np.random.default_rng(11)
    o, dirs = random_rays(rng, 4000, v.min(axis=0), v.max(axis=0), s.tris)
    got = ptb.closest_hit(o, dirs)
    ptb.set_pipeline(capi.PIPE_POOL)
    assert same(ptb.closest_hit(o, dirs), got)
    surf = {"W": 64, "H": 64, "iterations": 1, "camera": host.camera_at(64, 64, 60.0, 0, 0, (0.0, 0.0, -10.0)), "tris": s.tris,
            "mats": np.stack(s.mats), "kd_tree": host.nodes_to_records(s.kd_tree), "kd_tree_shift": np.asarray(s.kd_tree_shift, dtype=np.int32)}
    orc = po.OracleScene(surf, brute=True)
    t, idx = orc.closest_hit(o, dirs)
    assert (got[0].view(np.uint32) == t.view(np.uint32)).all()
    hit = t > 0
    assert (got[1][hit] == idx[hit]).all() and (got[1][~hit] == -1).all()
