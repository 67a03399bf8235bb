"""Host-side mirror of the reference's Scene interface (opencl_path_tracer_b200/host.py) against dumps of the
kernel-argument surface written by the reference's own host code (main.cpp compiled unmodified)."""
import os

import numpy as np
import pytest

from opencl_path_tracer_b200 import host, scenes
from tests import util


class _NoDevice:
    """host.Scene's record building needs no device."""


def fields_equal(a, b, fields):
    for f in fields:
        x, y = a[f], b[f]
        if x.ndim >= 1 and x.shape[-1] == 4:      # float3: lane 3 is padding (uninitialised in the reference)
            x, y = x[..., :3], y[..., :3]
        assert x.tobytes() == y.tobytes(), f


def test_cornell_records_match_reference_host():
    d = host.load_ptbd(os.path.join(util.GOLDEN, "cornell_96x64.ptbd"))
    s = host.Scene(ptb=_NoDevice())
    s.add_Obj(util.CORNELL, (555, 0, 0), (1, 1, 1), 0, 0)       # main.cpp:552-617
    s._flush()
    fields_equal(s.tris, d["tris"], ["r1", "r2", "r3", "N", "mati"])
    fields_equal(np.stack(s.mats), d["mats"], ["kd", "ks", "emission", "F0", "n", "shininess", "type"])
    fields_equal(host.nodes_to_records(s.kd_tree), d["kd_tree"], ["trii", "bl", "tr"])   # main.cpp:195-304
    assert s.kd_tree_shift == d["kd_tree_shift"].tolist() == [0, 3, 6]                    # main.cpp:537-540
    cam = host.camera_at(96, 64, 39.3, 0, 0, (277.5, 277.5, -800))                       # main.cpp:311-347
    fields_equal(cam.reshape(1), d["camera"].reshape(1), ["eye", "lookat", "up", "right", "XM", "YM"])


def test_procedural_scene_records_match_reference_host():
    d = host.load_ptbd(os.path.join(util.GOLDEN, "small_mixed_80x60.ptbd"))
    s = host.Scene(ptb=_NoDevice())
    desc = scenes.describe("small_mixed")
    scenes.populate(s, desc)                                      # add_Material / add_Triangle / end_Obj
    s._flush()
    fields_equal(s.tris, d["tris"], ["r1", "r2", "r3", "N", "mati"])
    fields_equal(np.stack(s.mats), d["mats"], ["kd", "ks", "emission", "F0", "n", "shininess", "type"])
    fields_equal(host.nodes_to_records(s.kd_tree), d["kd_tree"], ["trii", "bl", "tr"])
    assert s.kd_tree_shift == d["kd_tree_shift"].tolist()


@pytest.mark.skipif(not util.have_ref_cpu(), reason="oracle/_ref/ref_cpu not built")
def test_rotated_camera_matches_reference_host():
    ref = util.run_ref_cpu(40, 30, 0, 4, obj=util.CORNELL, eye=(100.0, 350.0, -600.0), fov=61.5, yaw=-13.8, pitch=15.6, want=("dump",))
    d = host.load_ptbd(ref["files"]["dump"])
    cam = host.camera_at(40, 30, 61.5, -13.8, 15.6, (100.0, 350.0, -600.0))
    fields_equal(cam.reshape(1), d["camera"].reshape(1), ["eye", "lookat", "up", "right", "XM", "YM"])


def test_seed_jump_ahead_equals_sequential():
    x, ref = 1, []
    for _ in range(20000):
        x = (x * 16807) % 2147483647
        ref.append(x)
    assert (host.minstd_rand0_seeds(20000) == np.asarray(ref, dtype=np.int32)).all()
    assert (host.minstd_rand0_seeds(777, skip=9000) == np.asarray(ref[9000:9777], dtype=np.int32)).all()
    assert host.minstd_rand0_seeds(0).size == 0


def test_missing_custom_mtl_key_raises(tmp_path):
    # main.cpp:568-571: unknown_parameter.at("Kn") throws when the custom keys are absent
    (tmp_path / "a.mtl").write_text("newmtl m\nKd 1 1 1\nKs 0 0 0\nKe 0 0 0\nNs 1\n")
    (tmp_path / "a.obj").write_text("mtllib a.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nusemtl m\nf 1 2 3\n")
    with pytest.raises(KeyError):
        host.load_obj(str(tmp_path / "a.obj"))


def test_coincident_centroids_are_rejected():
    # the reference's builder loops forever on > 6 triangles with identical centroids (main.cpp:246-257)
    t = host.Triangles(np.zeros((8, 3)), np.tile([1, 0, 0], (8, 1)), np.tile([0, 1, 0], (8, 1)), np.zeros(8))
    with pytest.raises(ValueError):
        host.build_object_tree(t, 0, [], 0)


def test_scene_generators_are_deterministic_and_sized():
    a, b = scenes.describe("spheres100k"), scenes.describe("spheres100k")
    assert a["verts"].tobytes() == b["verts"].tobytes()
    assert 90_000 <= len(a["verts"]) <= 120_000
    g = scenes.describe("small_mixed")
    assert len(g["verts"]) == g["obj_end"][-1] == len(g["tri_mat"])


def test_image_writers(tmp_path):
    rgba = np.arange(2 * 3 * 4, dtype=np.uint8).reshape(2, 3, 4)
    host.write_ppm(tmp_path / "a.ppm", rgba)
    raw = (tmp_path / "a.ppm").read_bytes()
    assert raw.startswith(b"P6\n3 2\n255\n") and raw[11:] == rgba[..., :3].tobytes()
    rgb = np.linspace(0, 5, 2 * 3 * 3, dtype=np.float32).reshape(2, 3, 3)
    host.write_pfm(tmp_path / "a.pfm", rgb)
    raw = (tmp_path / "a.pfm").read_bytes()
    assert raw.startswith(b"PF\n3 2\n-1.0\n") and np.frombuffer(raw[len(b"PF\n3 2\n-1.0\n"):], dtype="<f4").tolist() == rgb.ravel().tolist()


def _read_exr(path):
    """minimal reader of what host.write_exr writes (scanline, uncompressed, FLOAT channels): returns {name: (H, W)}"""
    import struct
    raw = open(path, "rb").read()
    assert struct.unpack_from("<ii", raw, 0) == (20000630, 2)
    off, attrs = 8, {}
    while raw[off] != 0:
        e = raw.index(b"\0", off); name = raw[off:e].decode(); off = e + 1
        e = raw.index(b"\0", off); typ = raw[off:e].decode(); off = e + 1
        (size,) = struct.unpack_from("<i", raw, off); off += 4
        attrs[name] = (typ, raw[off:off + size]); off += size
    off += 1
    assert attrs["compression"] == ("compression", b"\0") and attrs["lineOrder"] == ("lineOrder", b"\0")
    x0, y0, x1, y1 = struct.unpack("<iiii", attrs["dataWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    names, c = [], attrs["channels"][1]
    i = 0
    while c[i] != 0:
        e = c.index(b"\0", i); names.append(c[i:e].decode())
        ptype, plin, xs, ys = struct.unpack_from("<iBxxxii", c, e + 1)
        assert (ptype, xs, ys) == (2, 1, 1)
        i = e + 1 + 16
    assert names == sorted(names)
    table = np.frombuffer(raw, dtype="<u8", count=h, offset=off)
    out = {n: np.zeros((h, w), np.float32) for n in names}
    for y in range(h):
        o = int(table[y])
        yy, size = struct.unpack_from("<ii", raw, o)
        assert yy == y0 + y and size == len(names) * w * 4
        for k, n in enumerate(names):
            out[n][y] = np.frombuffer(raw, dtype="<f4", count=w, offset=o + 8 + k * w * 4)
    assert int(table[-1]) + 8 + len(names) * w * 4 == len(raw)
    return out


def test_exr_writer_round_trips(tmp_path):
    rng = np.random.default_rng(5)
    rgb = rng.random((7, 5, 3), dtype=np.float32) * 30.0          # linear radiance, row 0 = bottom
    rgb[0, 0] = (0.0, 17.0, np.float32(1e-30))
    host.write_exr(tmp_path / "a.exr", rgb)
    ch = _read_exr(tmp_path / "a.exr")
    assert sorted(ch) == ["B", "G", "R"]
    for k, n in enumerate("RGB"):
        assert ch[n].tobytes() == np.ascontiguousarray(rgb[::-1, :, k]).tobytes()   # EXR rows run top to bottom
