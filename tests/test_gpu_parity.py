"""GPU parity: the CUDA path, driven through the C-ABI exactly as the reference host drives OpenCL,
against the reference's own prog.cl executed on the CPU (oracle/_ref/ref_cpu) on the same scene,
camera and per-pixel seeds.

Tolerances (written here):
  * PARITY arithmetic: colors and RNG states are expected BIT-IDENTICAL to the oracle; the test
    allows <= 1e-5 of the pixels to differ (double-function rounding boundary cases, see oclshim.h).
  * FAST arithmetic (FMA, rsqrt, sincospi): >= 99.9 % of pixels within 1e-4 relative and image RMSE
    <= 1 % of the mean radiance at fixed spp.
"""
import numpy as np
import pytest

from opencl_path_tracer_b200 import capi, host
from tests import util

pytestmark = pytest.mark.gpu

W, H, SPP, DEPTH = 160, 120, 6, 4


@pytest.fixture(scope="module")
def cornell_ref():
    if not util.have_ref_cpu():
        pytest.skip("oracle/_ref/ref_cpu not built")
    return util.run_ref_cpu(W, H, SPP, DEPTH, obj=util.CORNELL, eye=(277.5, 277.5, -800), fov=39.3)


def render(ptb, ref, accel, math, mode, spp=SPP, pipeline=capi.PIPE_MEGAKERNEL):
    surf = host.load_ptbd(ref["files"]["dump"])
    ptb.set_accel(accel)
    ptb.set_math(math)
    if pipeline is not None:
        ptb.set_pipeline(pipeline)
    s = host.scene_from_surface(surf, ptb)
    if mode == "two_kernel":          # literal gen_ray + trace_ray launches per sample, unfused
        ptb.set_fuse_gen(0)
        for _ in range(spp):
            s.render()
    elif mode == "fused_launches":    # same calls, gen_ray folded into trace_ray
        ptb.set_fuse_gen(1)
        for _ in range(spp):
            s.render()
    else:                             # in-kernel sample loop
        s.render_spp(spp)
    ptb.finish()
    ptb.set_fuse_gen(1)
    return s.download_colors(), s.download_rnds()


@pytest.mark.parametrize("accel", [capi.ACCEL_BRUTE, capi.ACCEL_REFBVH, capi.ACCEL_LBVH])
@pytest.mark.parametrize("mode", ["two_kernel", "fused_launches", "render_spp"])
def test_parity_math_is_bit_identical(ptb, cornell_ref, accel, mode):
    colors, rnds = render(ptb, cornell_ref, accel, capi.MATH_PARITY, mode)
    st = util.compare(colors, cornell_ref["colors"])
    assert st["bit_equal_frac"] >= 1.0 - 1e-5, st
    assert (rnds == cornell_ref["rnds"]).mean() >= 1.0 - 1e-5


@pytest.mark.parametrize("accel", [capi.ACCEL_BRUTE, capi.ACCEL_LBVH])
def test_fast_math_within_tolerance(ptb, cornell_ref, accel):
    colors, _ = render(ptb, cornell_ref, accel, capi.MATH_FAST, "render_spp")
    st = util.compare(colors, cornell_ref["colors"])
    assert st["frac_rel_le_1e-4"] >= 0.999, st
    assert st["rmse_over_mean"] <= 0.01, st


def test_merged_quads_match_triangle_scan(ptb, cornell_ref):
    """Fast brute-force scan over merged primitives (prims.cu) vs. the same scan over the plain triangles:
    the Cornell box's fan-triangulated quads must merge, and the image must stay within the fast tolerance."""
    try:
        ptb.set_merge_prims(0)
        plain, _ = render(ptb, cornell_ref, capi.ACCEL_BRUTE, capi.MATH_FAST, "render_spp")
        n_plain = ptb.scan_prims()
        ptb.set_merge_prims(1)
        merged, _ = render(ptb, cornell_ref, capi.ACCEL_BRUTE, capi.MATH_FAST, "render_spp")
        n_merged = ptb.scan_prims()
    finally:
        ptb.set_merge_prims(1)
    assert n_plain == cornell_ref["info"]["n_tris"] if "n_tris" in cornell_ref["info"] else n_plain > 0
    assert n_merged < n_plain and 2 * n_merged >= n_plain, (n_merged, n_plain)
    st = util.compare(merged, plain)
    assert st["frac_rel_le_1e-4"] >= 0.999, st
    st = util.compare(merged, cornell_ref["colors"])
    assert st["frac_rel_le_1e-4"] >= 0.999 and st["rmse_over_mean"] <= 0.01, st


def test_depth1_preview_is_exact(ptb):
    """iterations == 1: flat kd+emission preview (prog.cl:323-325) — a sharp test of ray generation + closest hit."""
    if not util.have_ref_cpu():
        pytest.skip("oracle/_ref/ref_cpu not built")
    ref = util.run_ref_cpu(W, H, 2, 1, obj=util.CORNELL, eye=(277.5, 277.5, -800), fov=39.3)
    colors, rnds = render(ptb, ref, capi.ACCEL_LBVH, capi.MATH_PARITY, "render_spp", spp=2)
    st = util.compare(colors, ref["colors"])
    assert st["bit_equal_frac"] == 1.0, st
    assert (rnds == ref["rnds"]).all()


@pytest.mark.parametrize("accel", [capi.ACCEL_REFBVH, capi.ACCEL_BRUTE])
def test_work_counters_match_reference(ptb, accel):
    """rays / box tests / triangle tests of the REFBVH traversal and of the brute-force scan (prog.cl:318)
    equal the reference's own counts (idle lanes of the converged scan are not work)."""
    if not util.have_ref_cpu():
        pytest.skip("oracle/_ref/ref_cpu not built")
    ref = util.run_ref_cpu(96, 64, 2, 4, obj=util.CORNELL, eye=(277.5, 277.5, -800), fov=39.3,
                           count=accel != capi.ACCEL_BRUTE, brute=accel == capi.ACCEL_BRUTE)
    surf = host.load_ptbd(ref["files"]["dump"])
    ptb.set_accel(accel)
    ptb.set_math(capi.MATH_PARITY)
    ptb.set_count_work(1)
    ptb.reset_counters()
    s = host.scene_from_surface(surf, ptb)
    s.render_spp(2)
    c = ptb.counters()
    ptb.set_count_work(0)
    if accel == capi.ACCEL_BRUTE:     # the reference's brute-force build carries no counters: every ray tests every triangle
        assert c["rays"] == ref["info"]["rays"], (c, ref["info"])
        assert c["tri_tests"] == ref["info"]["tris"] * c["rays"] and c["box_tests"] == 0, c
        return
    for k in ("rays", "box_tests", "tri_tests", "tri_stage1", "tri_stage2", "tri_stage3"):
        assert c[k] == ref["info"][k], (k, c, ref["info"])


# ---------------------------------------------------------------- every material type, many objects
@pytest.fixture(scope="module")
def mixed_ref(tmp_path_factory):
    if not util.have_ref_cpu():
        pytest.skip("oracle/_ref/ref_cpu not built")
    from opencl_path_tracer_b200 import scenes
    d = scenes.describe("small_mixed")
    path = str(tmp_path_factory.mktemp("scenes") / "small_mixed.ptbs")
    scenes.write_ptbs(d, path)
    return util.run_ref_cpu(W, H, 8, 6, tribin=path, eye=d["eye"], fov=d["fov"])


@pytest.mark.parametrize("accel", [capi.ACCEL_BRUTE, capi.ACCEL_REFBVH, capi.ACCEL_LBVH])
def test_mixed_materials_parity(ptb, mixed_ref, accel):
    """mirror (type 1), glass (type 2), glossy lobe (ks != 0) and emitter paths, depth 6, 770 triangles in 4 objects."""
    colors, rnds = render(ptb, mixed_ref, accel, capi.MATH_PARITY, "render_spp", spp=8)
    st = util.compare(colors, mixed_ref["colors"])
    assert st["bit_equal_frac"] >= 1.0 - 1e-4, st
    assert (rnds == mixed_ref["rnds"]).mean() >= 1.0 - 1e-4


def test_mixed_materials_fast(ptb, mixed_ref):
    colors, _ = render(ptb, mixed_ref, capi.ACCEL_LBVH, capi.MATH_FAST, "render_spp", spp=8)
    st = util.compare(colors, mixed_ref["colors"])
    print("mixed fast vs reference:", st)
    assert st["frac_rel_le_1e-4"] >= 0.999, st
    assert st["rmse_over_mean"] <= 0.01, st


# ---------------------------------------------------------------- 100 k triangles: LBVH == reference trees == oracle
@pytest.fixture(scope="module")
def spheres_ref(tmp_path_factory):
    if not util.have_ref_cpu():
        pytest.skip("oracle/_ref/ref_cpu not built")
    from opencl_path_tracer_b200 import scenes
    d = scenes.describe("spheres100k")
    path = str(tmp_path_factory.mktemp("scenes") / "spheres100k.ptbs")
    scenes.write_ptbs(d, path)
    return util.run_ref_cpu(192, 108, 4, 4, tribin=path, eye=d["eye"], fov=d["fov"])


@pytest.mark.parametrize("accel", [capi.ACCEL_REFBVH, capi.ACCEL_LBVH])
def test_spheres100k_parity(ptb, spheres_ref, accel):
    colors, rnds = render(ptb, spheres_ref, accel, capi.MATH_PARITY, "render_spp", spp=4)
    st = util.compare(colors, spheres_ref["colors"])
    assert st["bit_equal_frac"] >= 1.0 - 1e-4, st
    assert (rnds == spheres_ref["rnds"]).mean() >= 1.0 - 1e-4


# ---------------------------------------------------------------- the reference's own host drives the CUDA library
def test_reference_host_over_cuda_shim(tmp_path):
    """oracle/_ref/ref_host_b200 = /root/reference/main.cpp compiled UNMODIFIED (but for the resolution constants)
    behind shim/CL/cl.hpp, linked against libptb200.so: same Scene methods, same launches, CUDA underneath."""
    import json, os, subprocess
    exe = os.path.join(util.ROOT, "oracle", "_ref", "ref_host_b200")
    if not (os.access(exe, os.X_OK) and util.have_ref_cpu()):
        pytest.skip("oracle/_ref binaries not built")
    ref = util.run_ref_cpu(128, 96, 5, 4, obj=util.CORNELL, want=("colors", "rnds"))
    out = str(tmp_path / "cuda.bin")
    r = subprocess.run([exe, "--w", "128", "--h", "96", "--spp", "5", "--iterations", "4", "--obj", util.CORNELL, "--pos", "555", "0", "0",
                        "--math", "0", "--out", out], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    info = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert info["backend"] == "cuda-sm100a" and info["error"] == ""
    colors = np.fromfile(out, dtype=np.float32).reshape(96, 128, 4)
    st = util.compare(colors, ref["colors"])
    assert st["bit_equal_frac"] >= 1.0 - 1e-4, st
    assert info["rays"] == ref["info"]["rays"]


# ---------------------------------------------------------------- wavefront pipeline == megakernel == oracle
@pytest.mark.parametrize("pipeline", [capi.PIPE_WAVEFRONT, capi.PIPE_POOL])
@pytest.mark.parametrize("which", ["mixed", "spheres"])
def test_wavefront_parity(ptb, mixed_ref, spheres_ref, which, pipeline):
    ref, spp = (mixed_ref, 8) if which == "mixed" else (spheres_ref, 4)
    colors, rnds = render(ptb, ref, capi.ACCEL_LBVH, capi.MATH_PARITY, "render_spp", spp=spp, pipeline=pipeline)
    st = util.compare(colors, ref["colors"])
    assert st["bit_equal_frac"] >= 1.0 - 1e-4, st
    assert (rnds == ref["rnds"]).mean() >= 1.0 - 1e-4


@pytest.mark.parametrize("pipeline", [capi.PIPE_MEGAKERNEL, capi.PIPE_WAVEFRONT, capi.PIPE_POOL])
def test_spheres100k_fast_flavour_against_the_reference(ptb, spheres_ref, pipeline):
    """the SHIPPING configuration of the LBVH scenes (fast arithmetic, merged outlier quads scanned from shared memory, `rf`
    triangle records, rcp.approx t) against the reference's own kernels on the 104 k-triangle scene, stated tolerance"""
    colors, rnds = render(ptb, spheres_ref, capi.ACCEL_LBVH, capi.MATH_FAST, "render_spp", spp=4, pipeline=pipeline)
    st = util.compare(colors, spheres_ref["colors"])
    assert st["frac_rel_le_1e-4"] >= 0.999 and st["rmse_over_mean"] <= 0.01, st
    assert (rnds == spheres_ref["rnds"]).mean() >= 0.999


@pytest.mark.parametrize("pipeline", [capi.PIPE_WAVEFRONT, capi.PIPE_POOL])
def test_quantised_wide_nodes_do_not_change_a_bit(ptb, spheres_ref, mixed_ref, pipeline):
    """the 64-byte quantised 4-ary nodes forced on for trees the default would walk as float nodes: same image, same RNG states"""
    try:
        ptb.set_quant_nodes(1)
        for ref, spp in ((spheres_ref, 4), (mixed_ref, 8)):
            colors, rnds = render(ptb, ref, capi.ACCEL_LBVH, capi.MATH_PARITY, "render_spp", spp=spp, pipeline=pipeline)
            st = util.compare(colors, ref["colors"])
            assert st["bit_equal_frac"] >= 1.0 - 1e-4, st
            assert (rnds == ref["rnds"]).mean() >= 1.0 - 1e-4
    finally:
        ptb.set_quant_nodes(2)


def test_fast_flavour_pipelines_agree_to_rounding_parity_flavour_to_the_bit(ptb, spheres_ref):
    """pool kernel vs wavefront on the same scene: the parity flavour (no FMA contraction) is the same bits through both; the fast
    flavour is the same SOURCE compiled into two kernels whose FMA contraction may differ, so it is held to the fast tolerance only"""
    out = {}
    for math in (capi.MATH_PARITY, capi.MATH_FAST):
        for pipe in (capi.PIPE_WAVEFRONT, capi.PIPE_POOL):
            out[math, pipe] = render(ptb, spheres_ref, capi.ACCEL_LBVH, math, "render_spp", spp=4, pipeline=pipe)
    a, b = out[capi.MATH_PARITY, capi.PIPE_WAVEFRONT], out[capi.MATH_PARITY, capi.PIPE_POOL]
    assert util.compare(a[0], b[0])["bit_equal_frac"] == 1.0 and (a[1] == b[1]).all()
    a, b = out[capi.MATH_FAST, capi.PIPE_WAVEFRONT], out[capi.MATH_FAST, capi.PIPE_POOL]
    st = util.compare(a[0], b[0])
    print("fast flavour, pool vs wavefront:", st)
    assert st["frac_rel_le_1e-4"] >= 0.999 and st["rmse_over_mean"] <= 0.01, st


@pytest.mark.parametrize("mode", ["fused_launches", "render_spp"])
def test_pool_kernel_on_cornell_is_bit_identical(ptb, cornell_ref, mode):
    """the persistent pool kernel (PTB_PIPE_POOL) forced onto the 32-triangle scene through the LBVH: per-sample launches
    (gen_ray folded in, one kernel launch per sample) and the in-kernel sample loop"""
    colors, rnds = render(ptb, cornell_ref, capi.ACCEL_LBVH, capi.MATH_PARITY, mode, pipeline=capi.PIPE_POOL)
    st = util.compare(colors, cornell_ref["colors"])
    assert st["bit_equal_frac"] == 1.0, st
    assert (rnds == cornell_ref["rnds"]).all()


# ---------------------------------------------------------------- scene ingest on the device
def _records_equal(a, b):
    """bit patterns equal; a NaN (degenerate face: 0/0 normal) matches any NaN"""
    for k in ("r1", "r2", "r3", "N"):
        x, y = a[k][:, :3], b[k][:, :3]
        same = (x.view(np.uint32) == y.view(np.uint32)) | (np.isnan(x) & np.isnan(y))
        if not same.all():
            bad = np.argwhere(~same)
            print(k, len(bad), "mismatches; first:", bad[0], x[bad[0][0]], y[bad[0][0]])
            return False
    return np.array_equal(a["mati"], b["mati"])


@pytest.mark.parametrize("pose", [((555, 0, 0), (1, 1, 1), 0.0, 0.0), ((12.5, -3.0, 40.0), (0.5, 2.0, 1.25), 33.0, -71.5)])
def test_device_ingest_matches_host_records(ptb, pose):
    """ptb_ingest_mesh writes the same 80-byte Triangle records (vertices and unit normal, bit for bit) as the host
    mirror of Scene::add_Obj + the Triangle constructor (main.cpp:589-612, 144-166), itself checked against the
    reference host's dumps in tests/test_host.py."""
    pos, scale, pitch, yaw = pose
    ref = host.Scene(ptb)
    ref.build_trees = False
    ref.add_Obj(util.CORNELL, pos, scale, pitch, yaw)
    ref._flush()
    dev = host.Scene(ptb)
    dev.add_Obj_device(util.CORNELL, pos, scale, pitch, yaw)
    dev.upload_Triangles()
    got = ptb.read(dev.buffer_tris, host.TRIANGLE, dev.tris_size)
    assert dev.tris_size == len(ref.tris) == 32
    assert _records_equal(got, ref.tris)
    assert len(dev.mats) == len(ref.mats)


def test_device_ingest_large_random_mesh(ptb):
    rng = np.random.default_rng(7)
    verts = (rng.random((5000, 3), dtype=np.float32) - 0.5) * 800.0
    faces = rng.integers(0, len(verts), size=(20000, 3), dtype=np.int32)
    mats = rng.integers(0, 9, size=len(faces), dtype=np.int32)
    pos, scale, pitch, yaw = (100.0, 50.0, -25.0), (1.5, 0.75, 1.0), -12.0, 205.0
    v = verts.copy()
    v[:, 0] = -v[:, 0]
    v = host.rotate_y(host.rotate_x(v, pitch), yaw)
    v = v * np.asarray(scale, dtype=np.float32) + np.asarray(pos, dtype=np.float32)
    want = host.Triangles(v[faces[:, 0]], v[faces[:, 1]], v[faces[:, 2]], 3 + mats)
    h = ptb.alloc(len(faces) * host.TRIANGLE.itemsize)
    ptb.ingest_mesh(h, 0, verts, faces, mats, pos, scale, pitch, yaw, 3)
    got = ptb.read(h, host.TRIANGLE, len(faces))
    ptb.release(h)
    # degenerate faces (repeated vertex) give NaN normals on both sides: compare bit patterns
    assert _records_equal(got, want)
    with pytest.raises(capi.PtbError):
        ptb.ingest_mesh(h, 0, verts, faces, mats, pos, scale, pitch, yaw, 3)      # released handle
    ptb.clear_error()                                                             # errors are sticky
    h2 = ptb.alloc(10 * host.TRIANGLE.itemsize)
    with pytest.raises(capi.PtbError):
        ptb.ingest_mesh(h2, 0, verts, faces, mats, pos, scale, pitch, yaw, 3)     # buffer too small
    ptb.clear_error()
    bad = faces.copy(); bad[5, 1] = len(verts)
    h3 = ptb.alloc(len(faces) * host.TRIANGLE.itemsize)
    with pytest.raises(capi.PtbError):
        ptb.ingest_mesh(h3, 0, verts, bad, mats, pos, scale, pitch, yaw, 3)       # index out of range
    ptb.clear_error()
    ptb.release(h2); ptb.release(h3)


def test_device_ingested_scene_renders_like_host_scene(ptb, cornell_ref):
    """Cornell ingested on the device (face order, LBVH / brute force) renders the reference image: closest hits do not
    depend on the triangle order except through exact-t ties, which the Cornell box does not produce at this size."""
    surf = host.load_ptbd(cornell_ref["files"]["dump"])
    ptb.set_accel(capi.ACCEL_AUTO)
    ptb.set_math(capi.MATH_PARITY)
    ptb.set_pipeline(capi.PIPE_MEGAKERNEL)
    s = host.Scene(ptb)
    s.init_Scene(W, H)
    s.add_Obj_device(util.CORNELL, (555, 0, 0), (1, 1, 1), 0.0, 0.0)
    s.upload_Triangles()
    s.upload_Materials()
    s.iterations = DEPTH
    cam = np.array(surf["camera"])
    s._camera = lambda: cam
    s.render_spp(SPP)
    ptb.finish()
    st = util.compare(s.download_colors(), cornell_ref["colors"])
    assert st["bit_equal_frac"] >= 1.0 - 1e-4, st


# ---------------------------------------------------------------- RNG states outside [0, 2^31)
@pytest.mark.parametrize("math", [capi.MATH_PARITY, capi.MATH_FAST])
def test_negative_rng_states_follow_the_reference_conversion(ptb, cornell_ref, math):
    """`(ulong)seed` of a NEGATIVE int sign-extends (prog.cl:73): such states never come from the reference host, but the
    buffer is host-writable, so the kernels must follow the conversion.  The fast flavour's branch-free generator
    (pt_rand_nonneg) may only be selected after the C-ABI has scanned the uploaded states — this is the test of that scan."""
    from oracle import pt_oracle_py as po
    surf = host.load_ptbd(cornell_ref["files"]["dump"])
    seeds = host.minstd_rand0_seeds(W * H).astype(np.int32)
    seeds[::7] = -seeds[::7]
    seeds[3] = -2147483648
    orc = po.OracleScene(surf, brute=True, seeds=seeds.copy())
    want = orc.render(SPP).copy()
    ptb.set_accel(capi.ACCEL_BRUTE); ptb.set_math(math); ptb.set_pipeline(capi.PIPE_MEGAKERNEL)
    s = host.scene_from_surface(surf, ptb)
    ptb.write(s.buffer_rnds, seeds)
    s.render_spp(SPP)
    ptb.finish()
    got, rnds = s.download_colors(), s.download_rnds()
    st = util.compare(got, want)
    if math == capi.MATH_PARITY:
        assert st["bit_equal_frac"] >= 1.0 - 1e-5, st
        assert (rnds == orc.rnds).mean() >= 1.0 - 1e-5
    else:
        assert st["frac_rel_le_1e-4"] >= 0.999 and st["rmse_over_mean"] <= 0.01, st
        assert (rnds == orc.rnds).mean() >= 0.999
    # and the same scene object, re-seeded with valid states, is back on the fast generator with the right image
    ptb.write(s.buffer_rnds, host.minstd_rand0_seeds(W * H))
    s.current_sample = 0
    s.render_spp(SPP)
    ptb.finish()
    st = util.compare(s.download_colors(), cornell_ref["colors"])
    assert st["frac_rel_le_1e-4"] >= 0.999, st
