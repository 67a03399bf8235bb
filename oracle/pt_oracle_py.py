"""TEST INFRASTRUCTURE — ctypes front end of oracle/libpt_oracle.so (pt_oracle.c, the plain-C restatement of
prog.cl's hot path).  May be imported only by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline leg;
the product path never touches it."""
import ctypes as C
import os
import subprocess
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class F3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float), ("w", C.c_float)]


class Camera(C.Structure):
    _fields_ = [("eye", F3), ("lookat", F3), ("up", F3), ("right", F3), ("XM", C.c_float), ("YM", C.c_float), ("pad", C.c_float * 2)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays", "box_tests", "tri_tests", "tri_stage1", "tri_stage2", "tri_stage3")]


class Scene(C.Structure):
    _fields_ = [("tris", C.c_void_p), ("tris_size", C.c_int32), ("materials", C.c_void_p), ("kd_tree", C.c_void_p),
                ("kd_tree_shift", C.c_void_p), ("kd_tree_shift_size", C.c_int32), ("rays", C.c_void_p), ("rnds", C.c_void_p),
                ("colors", C.c_void_p), ("tex", C.c_void_p), ("cam", Camera), ("width", C.c_int32), ("height", C.c_int32),
                ("iterations", C.c_int32), ("brute", C.c_int32)]


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libpt_oracle.so")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(_HERE, "pt_oracle.c")):
            subprocess.run(["make", "-C", _HERE, "port"], check=True, capture_output=True)
        _LIB = C.CDLL(path)
        _LIB.pto_rand.restype = C.c_float
        _LIB.pto_rand.argtypes = [C.POINTER(C.c_int32)]
        _LIB.pto_render.argtypes = [C.POINTER(Scene), C.c_int32, C.c_int32, C.c_int32, C.POINTER(Counters)]
        _LIB.pto_closest_hit.restype = C.c_float
        _LIB.pto_closest_hit.argtypes = [C.POINTER(Scene), C.c_void_p, C.c_void_p, C.POINTER(C.c_int32)]
        _LIB.pto_minstd_rand0.argtypes = [C.c_void_p, C.c_int64, C.c_int64]
        _LIB.pto_reinhard.restype = C.c_float
        _LIB.pto_reinhard.argtypes = [C.c_void_p, C.c_void_p]
        _LIB.pto_struct_sizes.argtypes = [C.c_void_p]
    return _LIB


def rand_stream(seed, count):
    """(values, final state) of `count` draws of prog.cl:72-77 from `seed`."""
    s = C.c_int32(seed)
    vals = [lib().pto_rand(C.byref(s)) for _ in range(count)]
    return np.asarray(vals, dtype=np.float32), s.value


def minstd_rand0(count, skip=0):
    out = np.empty(count, dtype=np.int32)
    lib().pto_minstd_rand0(out.ctypes.data, count, skip)
    return out


def struct_sizes():
    o = np.zeros(6, dtype=np.int32)
    lib().pto_struct_sizes(o.ctypes.data)
    return dict(zip(["Material", "Ray", "Triangle", "Node", "Camera", "float3"], o.tolist()))


def filt_im(colors, width, height):
    """prog.cl:391-427 over a (H, W, 4) float32 accumulation buffer: returns the (H, W, 4) RGBA float image, zero where the
    kernel writes nothing."""
    c = np.ascontiguousarray(colors, dtype=np.float32).reshape(height * width * 4)
    tex = np.zeros(height * width * 4, dtype=np.float32)
    fn = lib().pto_filt_im
    fn.restype, fn.argtypes = None, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32]
    fn(c.ctypes.data, tex.ctypes.data, width, height)
    return tex.reshape(height, width, 4)


class OracleScene:
    """The kernel-argument surface (a dict as returned by host.load_ptbd / built by the host mirror) bound to the C oracle."""

    def __init__(self, surface, width=None, height=None, brute=False, seeds=None, want_tex=False):
        self.W, self.H = width or surface["W"], height or surface["H"]
        n = self.W * self.H
        self.tris = np.ascontiguousarray(surface["tris"])
        self.mats = np.ascontiguousarray(surface["mats"])
        self.nodes = np.ascontiguousarray(surface["kd_tree"])
        self.shift = np.ascontiguousarray(surface["kd_tree_shift"], dtype=np.int32)
        self.rays = np.zeros(n * 8, dtype=np.float32)
        self.rnds = np.ascontiguousarray(seeds if seeds is not None else minstd_rand0(n), dtype=np.int32)
        self.colors = np.zeros(n * 4, dtype=np.float32)
        self.tex = np.zeros(n * 4, dtype=np.float32) if want_tex else None
        s = Scene()
        s.tris, s.tris_size = self.tris.ctypes.data, len(self.tris)
        s.materials = self.mats.ctypes.data
        s.kd_tree, s.kd_tree_shift, s.kd_tree_shift_size = self.nodes.ctypes.data, self.shift.ctypes.data, len(self.shift)
        s.rays, s.rnds, s.colors = self.rays.ctypes.data, self.rnds.ctypes.data, self.colors.ctypes.data
        s.tex = self.tex.ctypes.data if want_tex else None
        C.memmove(C.byref(s.cam), np.ascontiguousarray(surface["camera"]).tobytes(), 80)
        s.width, s.height, s.iterations, s.brute = self.W, self.H, int(surface["iterations"]), int(brute or len(self.shift) == 0)
        self.c = s
        self.sample = 0

    def render(self, spp, threads=0):
        cnt = Counters()
        threads = threads or (os.cpu_count() or 1)
        t0 = time.time()
        lib().pto_render(C.byref(self.c), self.sample, spp, threads, C.byref(cnt))
        self.seconds = time.time() - t0
        self.sample += spp
        self.counters = {n: int(getattr(cnt, n)) for n, _ in Counters._fields_}
        return self.colors.reshape(self.H, self.W, 4)

    def closest_hit(self, origins, dirs):
        origins = np.ascontiguousarray(origins, dtype=np.float32)
        dirs = np.ascontiguousarray(dirs, dtype=np.float32)
        t = np.empty(len(origins), dtype=np.float32)
        idx = np.empty(len(origins), dtype=np.int32)
        k = C.c_int32()
        for i in range(len(origins)):
            t[i] = lib().pto_closest_hit(C.byref(self.c), origins[i].ctypes.data, dirs[i].ctypes.data, C.byref(k))
            idx[i] = k.value
        return t, idx


def time_render(scene_name, W, H, spp, depth, threads):
    """bench.py's CPU leg when oracle/_ref/ref_cpu cannot serve the scene: the port on a surface built by the
    host mirror (brute force is hopeless beyond tiny scenes, so the reference-style trees are built)."""
    import sys
    sys.path.insert(0, os.path.dirname(_HERE))
    from opencl_path_tracer_b200 import host, scenes

    class _NoDev:
        pass
    s = host.Scene(ptb=_NoDev())
    if scene_name == "cornell":
        s.add_Obj(os.path.join(os.path.dirname(_HERE), "scenes", "cornell.obj"), (555, 0, 0), (1, 1, 1), 0, 0)
        fov, yaw, pitch, eye = 39.3, 0.0, 0.0, (277.5, 277.5, -800.0)
    else:
        d = scenes.describe(scene_name)
        scenes.populate(s, d)
        fov, yaw, pitch, eye = d["fov"], d["yaw"], d["pitch"], d["eye"]
    s._flush()
    surf = {"W": W, "H": H, "iterations": depth, "camera": host.camera_at(W, H, fov, yaw, pitch, eye), "tris": s.tris,
            "mats": np.stack(s.mats), "kd_tree": host.nodes_to_records(s.kd_tree), "kd_tree_shift": np.asarray(s.kd_tree_shift, dtype=np.int32)}
    o = OracleScene(surf)
    o.render(spp, threads)
    return {"kind": "port", "seconds": o.seconds, "rays": o.counters["rays"], "threads": threads}
