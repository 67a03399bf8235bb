"""The oracle (oracle/pt_oracle.c, a plain-C restatement of prog.cl's hot path) against
  * the survey's known-answer vectors for the RNG, the seeds and the struct layout,
  * golden fixtures produced by the REFERENCE's own prog.cl + main.cpp (tests/golden/make_golden.py),
  * the reference itself run live (oracle/_ref/ref_cpu) when it is built in this container.
All comparisons are bit-exact: the oracle and the reference share the canonical arithmetic (oracle/oclshim.h)."""
import json
import os

import numpy as np
import pytest

from opencl_path_tracer_b200 import host
from oracle import pt_oracle_py as po
from tests import util

G = util.GOLDEN


def test_struct_sizes_are_the_kernel_abi():
    # prog.cl:1-35 ↔ main.cpp:92-348 (SURVEY §4 KAT)
    assert po.struct_sizes() == {"Material": 80, "Ray": 32, "Triangle": 80, "Node": 48, "Camera": 80, "float3": 16}


def test_rand_known_answers():
    # prog.cl:72-77 from state 16807: 0.377786666, 0.140372753, state afterwards 301448195
    vals, state = po.rand_stream(16807, 2)
    assert vals.tolist() == [np.float32(0.377786666), np.float32(0.140372753)]
    assert state == 301448195


def test_rand_can_return_exactly_one():
    # states >= 2147483584 round to 2^31 in float and 2147483647.0f == 2^31: range is (0, 1]
    # find the predecessor state s with s*48271 mod M == M-1 : s = (M-1) * inverse(48271)
    M = 2147483647
    s = ((M - 1) * pow(48271, -1, M)) % M
    vals, state = po.rand_stream(s, 1)
    assert state == M - 1 and vals[0] == np.float32(1.0)


def test_host_seeds_are_minstd_rand0():
    # main.cpp:45,522-525: RNDS[0..2] = 16807, 282475249, 1622650073
    assert po.minstd_rand0(3).tolist() == [16807, 282475249, 1622650073]
    assert (po.minstd_rand0(5000, skip=1234) == host.minstd_rand0_seeds(5000, skip=1234)).all()


@pytest.mark.parametrize("name,surface,spp", [
    ("cornell_96x64_spp3_d4", "cornell_96x64.ptbd", 3),
    ("cornell_96x64_spp2_d1", "cornell_96x64.ptbd", 2),
    ("small_mixed_80x60_spp4_d6", "small_mixed_80x60.ptbd", 4),
])
def test_port_reproduces_reference_golden(name, surface, spp):
    info = json.load(open(os.path.join(G, name + ".json")))
    surf = host.load_ptbd(os.path.join(G, surface))
    surf["iterations"] = info["iterations"]
    gold = np.fromfile(os.path.join(G, name + ".colors.f32"), dtype=np.float32).reshape(info["h"], info["w"], 4)
    gold_rnds = np.fromfile(os.path.join(G, name + ".rnds.i32"), dtype=np.int32)
    o = po.OracleScene(surf, want_tex=True)
    colors = o.render(spp, threads=4)
    assert colors[..., :3].tobytes() == gold[..., :3].tobytes()
    assert (o.rnds == gold_rnds).all()
    for k in ("rays", "box_tests", "tri_tests", "tri_stage1", "tri_stage2", "tri_stage3"):
        assert o.counters[k] == info[k], k
    tex_path = os.path.join(G, name + ".tex.f32")
    if os.path.exists(tex_path):     # tone-mapped image (prog.cl:380); NaN where L == 0, compare bit patterns
        tex = np.fromfile(tex_path, dtype=np.float32)
        assert o.tex.view(np.uint32).tolist() == tex.view(np.uint32).tolist()


def test_brute_force_equals_reference_trees():
    # the north_star contract (prog.cl:318 vs :319): SURVEY [probe] found them bit-identical on Cornell
    surf = host.load_ptbd(os.path.join(G, "small_mixed_80x60.ptbd"))
    a = po.OracleScene(surf)
    b = po.OracleScene(surf, brute=True)
    assert a.render(2, 4).tobytes() == b.render(2, 4).tobytes()
    assert (a.rnds == b.rnds).all()


@pytest.mark.skipif(not util.have_ref_cpu(), reason="oracle/_ref/ref_cpu not built (no /root/reference here)")
def test_port_equals_reference_live():
    ref = util.run_ref_cpu(72, 40, 5, 4, obj=util.CORNELL, eye=(200.0, 300.0, -700.0), fov=50.0, yaw=8.0, pitch=-4.0, count=True)
    surf = host.load_ptbd(ref["files"]["dump"])
    o = po.OracleScene(surf)
    colors = o.render(5, threads=3)
    assert colors[..., :3].tobytes() == ref["colors"][..., :3].tobytes()
    assert (o.rnds == ref["rnds"]).all()
    assert o.counters["rays"] == ref["info"]["rays"] and o.counters["tri_tests"] == ref["info"]["tri_tests"]


def test_samples_are_sequential_per_pixel():
    # the RNG state persists across samples: 2 + 3 samples == 5 samples
    surf = host.load_ptbd(os.path.join(G, "cornell_96x64.ptbd"))
    a = po.OracleScene(surf)
    a.render(2, 2)
    a.render(3, 4)
    b = po.OracleScene(surf)
    b.render(5, 1)
    assert a.colors.tobytes() == b.colors.tobytes() and (a.rnds == b.rnds).all()


def test_filt_im_port_reproduces_the_reference_kernel_fixture():
    # tests/golden/filt_im_48x32.*: the reference's own filt_im (prog.cl:391-427) on an image with constructed grey ties
    src = np.fromfile(os.path.join(G, "filt_im_48x32.in.f32"), dtype=np.float32).reshape(32, 48, 4)
    want = np.fromfile(os.path.join(G, "filt_im_48x32.out.f32"), dtype=np.float32).reshape(32, 48, 4)
    got = po.filt_im(src, 48, 32)
    assert got.view(np.uint32).tolist() == want.view(np.uint32).tolist()
    assert int((want[..., 3] == 1).sum()) == 30 * 47 - 1       # rows 1..30 x columns 1..47, minus the pixel whose reads leave the buffer


@pytest.mark.skipif(not util.have_ref_cpu(), reason="oracle/_ref/ref_cpu not built (no /root/reference here)")
def test_filt_im_port_equals_reference_live(tmp_path):
    rng = np.random.default_rng(3)
    w, h = 37, 21
    src = np.zeros((h, w, 4), dtype=np.float32)
    src[..., :3] = np.round(rng.random((h, w, 3), dtype=np.float32) * 6.0) / 3.0     # coarse values: many equal greys with different colours
    cin, cout = str(tmp_path / "in.f32"), str(tmp_path / "out.f32")
    src.tofile(cin)
    util.run_ref_cpu(w, h, 0, 4, obj=util.CORNELL, want=(), extra=("--colors-in", cin, "--filt-out", cout))
    want = np.fromfile(cout, dtype=np.float32).reshape(h, w, 4)
    assert po.filt_im(src, w, h).view(np.uint32).tolist() == want.view(np.uint32).tolist()
