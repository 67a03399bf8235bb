"""ctypes bindings of include/ptb200.h (the C-ABI that replaces the reference's cl:: calls,
main.cpp:389-528, 618-682).  One thin method per entry point; errors raise :class:`PtbError`."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

KERNEL_GEN_RAY, KERNEL_TRACE_RAY, KERNEL_FILT_IM = 0, 1, 2
ACCEL_AUTO, ACCEL_BRUTE, ACCEL_REFBVH, ACCEL_LBVH = 0, 1, 2, 3
MATH_PARITY, MATH_FAST = 0, 1
PIPE_MEGAKERNEL, PIPE_WAVEFRONT, PIPE_AUTO, PIPE_POOL = 0, 1, 2, 3
SPLIT_TILE, SPLIT_SPP = 0, 1

# every symbol include/ptb200.h declares (checked by tests/test_abi.py against the header text)
SYMBOLS = [
    "ptb_init", "ptb_shutdown", "ptb_device_string", "ptb_device_compute_units", "ptb_backend_name",
    "ptb_alloc", "ptb_alloc_image", "ptb_retain", "ptb_release", "ptb_write", "ptb_read", "ptb_mem_size", "ptb_ingest_mesh",
    "ptb_wrap_device", "ptb_device_ptr", "ptb_ipc_export", "ptb_ipc_import", "ptb_kernel_id", "ptb_set_arg", "ptb_set_arg_mem", "ptb_launch",
    "ptb_finish", "ptb_last_error", "ptb_clear_error", "ptb_set_accel", "ptb_set_math", "ptb_set_pipeline", "ptb_set_wf_subs",
    "ptb_set_fuse_gen", "ptb_set_lbvh_leaf", "ptb_set_merge_prims", "ptb_set_parity_filter", "ptb_scan_prims", "ptb_get_accel", "ptb_build_accel", "ptb_render",
    "ptb_set_pixel_shard", "ptb_set_accumulate_sum", "ptb_resolve_sum", "ptb_set_count_work", "ptb_get_counters",
    "ptb_reset_counters", "ptb_resolve_rgba8", "ptb_filter_median", "ptb_set_stream", "ptb_get_stream",
    "ptb_accel_stats", "ptb_launch_count", "ptb_closest_hit", "ptb_invalidate", "ptb_resolve_frame", "ptb_save_state", "ptb_load_state",
    "ptb_comm_unique_id", "ptb_comm_init", "ptb_comm_ranks", "ptb_comm_rank", "ptb_comm_reduce", "ptb_comm_broadcast", "ptb_comm_barrier", "ptb_comm_destroy",
    "ptb_set_gpus", "ptb_get_gpus", "ptb_set_quant_nodes", "ptb_resolve_rgba8_noalloc",
]


class PtbError(RuntimeError):
    pass


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays", "box_tests", "tri_tests", "tri_stage1", "tri_stage2", "tri_stage3", "pixel_samples")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


def lib_path():
    """libptb200.so next to this file; PTB200_LIB names another build of the same library (A/B experiments)."""
    return os.environ.get("PTB200_LIB") or os.path.join(_HERE, "libptb200.so")


def load_library(path=None):
    """dlopen libptb200.so and declare the prototypes.  Raises if the library has not been built —
    there is deliberately no fallback implementation."""
    path = path or lib_path()
    if not os.path.exists(path):
        raise PtbError(f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(make -C opencl_path_tracer_b200/csrc); there is no CPU fallback")
    lib = C.CDLL(path)
    u64, sz, i, vp, cp = C.c_uint64, C.c_size_t, C.c_int, C.c_void_p, C.c_char_p
    proto = {
        "ptb_init": (i, [i]), "ptb_shutdown": (i, []), "ptb_device_string": (i, [i, cp, sz]),
        "ptb_device_compute_units": (i, []), "ptb_backend_name": (cp, []),
        "ptb_alloc": (u64, [sz]), "ptb_alloc_image": (u64, [i, i]), "ptb_retain": (i, [u64]), "ptb_release": (i, [u64]),
        "ptb_write": (i, [u64, sz, sz, vp]), "ptb_read": (i, [u64, sz, sz, vp]), "ptb_mem_size": (sz, [u64]), "ptb_closest_hit": (i, [u64, sz, u64]),
        "ptb_ingest_mesh": (i, [u64, sz, vp, sz, vp, vp, sz, vp, vp, C.c_float, C.c_float, i]),
        "ptb_wrap_device": (u64, [vp, sz]), "ptb_device_ptr": (vp, [u64]),
        "ptb_ipc_export": (i, [u64, vp]), "ptb_ipc_import": (u64, [vp, sz]),
        "ptb_kernel_id": (i, [cp]), "ptb_set_arg": (i, [i, i, sz, vp]), "ptb_set_arg_mem": (i, [i, i, u64]),
        "ptb_launch": (i, [i, sz, sz]), "ptb_finish": (i, []), "ptb_last_error": (cp, []), "ptb_clear_error": (None, []),
        "ptb_set_accel": (i, [i]), "ptb_set_math": (i, [i]), "ptb_set_pipeline": (i, [i]), "ptb_set_wf_subs": (i, [i]), "ptb_set_fuse_gen": (i, [i]),
        "ptb_set_lbvh_leaf": (i, [i]), "ptb_set_merge_prims": (i, [i]), "ptb_set_parity_filter": (i, [i]), "ptb_scan_prims": (i, []), "ptb_get_accel": (i, []), "ptb_build_accel": (i, []), "ptb_render": (i, [i]),
        "ptb_set_pixel_shard": (i, [i, i, i]), "ptb_set_accumulate_sum": (i, [i]), "ptb_resolve_sum": (i, [u64, i]),
        "ptb_set_count_work": (i, [i]), "ptb_get_counters": (i, [C.POINTER(Counters)]), "ptb_reset_counters": (i, []),
        "ptb_resolve_rgba8": (i, [u64, i, i, i, vp]), "ptb_filter_median": (i, [u64, u64, i, i]),
        "ptb_set_stream": (i, [vp]), "ptb_get_stream": (vp, []),
        "ptb_accel_stats": (i, [C.POINTER(i), C.POINTER(i), C.POINTER(i), C.POINTER(C.c_float)]),
        "ptb_launch_count": (u64, []),
        "ptb_invalidate": (i, [u64]), "ptb_resolve_frame": (i, [u64, i, i, i, i, u64, vp]),
        "ptb_save_state": (i, [cp]), "ptb_load_state": (i, [cp, C.POINTER(i)]),
        "ptb_comm_unique_id": (i, [vp]), "ptb_comm_init": (i, [i, i, vp]), "ptb_comm_ranks": (i, []), "ptb_comm_rank": (i, []),
        "ptb_resolve_rgba8_noalloc": (i, [u64, i, i, i, vp]), "ptb_set_quant_nodes": (i, [i]), "ptb_set_gpus": (i, [i, i]), "ptb_get_gpus": (i, []),
        "ptb_comm_reduce": (i, [u64, i]), "ptb_comm_broadcast": (i, [u64, i]), "ptb_comm_barrier": (i, []), "ptb_comm_destroy": (i, []),
    }
    for name, (res, args) in proto.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


class Ptb:
    """Object façade over the C-ABI.  Method names follow the C names without the ``ptb_`` prefix."""

    def __init__(self, device=-1, path=None):
        self.lib = load_library(path)
        self._ck(self.lib.ptb_init(device), "ptb_init")

    # -- error handling: the reference ignores return codes; this host does not
    def _ck(self, rc, what):
        if rc != 0:
            msg = self.lib.ptb_last_error().decode() or f"error {rc}"
            raise PtbError(f"{what}: {msg}")
        return rc

    def last_error(self):
        return self.lib.ptb_last_error().decode()

    def clear_error(self):
        self.lib.ptb_clear_error()

    def backend_name(self):
        return self.lib.ptb_backend_name().decode()

    def device_string(self, what=0):
        buf = C.create_string_buffer(512)
        self._ck(self.lib.ptb_device_string(what, buf, 512), "ptb_device_string")
        return buf.value.decode()

    def compute_units(self):
        return self.lib.ptb_device_compute_units()

    # -- memory
    def alloc(self, nbytes):
        h = self.lib.ptb_alloc(nbytes)
        if not h:
            raise PtbError("ptb_alloc: " + self.last_error())
        return h

    def alloc_image(self, w, h):
        m = self.lib.ptb_alloc_image(w, h)
        if not m:
            raise PtbError("ptb_alloc_image: " + self.last_error())
        return m

    def release(self, h):
        self.lib.ptb_release(h)

    def write(self, h, array, offset=0):
        a = np.ascontiguousarray(array)
        self._ck(self.lib.ptb_write(h, offset, a.nbytes, a.ctypes.data), "ptb_write")

    def write_ptr(self, h, ptr, nbytes, offset=0):
        self._ck(self.lib.ptb_write(h, offset, nbytes, ptr), "ptb_write")

    def read(self, h, dtype, count, offset=0):
        out = np.empty(count, dtype=dtype)
        self._ck(self.lib.ptb_read(h, offset, out.nbytes, out.ctypes.data), "ptb_read")
        return out

    def read_ptr(self, h, ptr, nbytes, offset=0):
        self._ck(self.lib.ptb_read(h, offset, nbytes, ptr), "ptb_read")

    def ingest_mesh(self, h, first_tri, vertices, faces, mat_ids, pos, scale, pitch, yaw, mat_offset):
        """ptb_ingest_mesh: transform + normals + Triangle records on the device (main.cpp:589-612, 144-166)."""
        v = np.ascontiguousarray(vertices, dtype=np.float32).reshape(-1, 3)
        f = np.ascontiguousarray(faces, dtype=np.int32).reshape(-1, 3)
        m = np.ascontiguousarray(mat_ids, dtype=np.int32).reshape(-1)
        if len(m) != len(f):
            raise ValueError("one material id per face")
        ps = np.ascontiguousarray(pos, dtype=np.float32).reshape(3)
        sc = np.ascontiguousarray(scale, dtype=np.float32).reshape(3)
        self._ck(self.lib.ptb_ingest_mesh(h, first_tri, v.ctypes.data, len(v), f.ctypes.data, m.ctypes.data, len(f),
                                          ps.ctypes.data, sc.ctypes.data, float(pitch), float(yaw), int(mat_offset)),
                 "ptb_ingest_mesh")

    def closest_hit(self, origins, dirs):
        """ptb_closest_hit on (n,3) origins / directions: returns (t float32[n], triangle int32[n]); (-1, -1) = miss."""
        o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        n = len(o)
        rays = np.zeros((n, 8), dtype=np.float32)
        rays[:, 0:3], rays[:, 4:7] = o, d
        h_rays, h_out = self.upload(rays), self.alloc(max(n, 1) * 8)
        try:
            self._ck(self.lib.ptb_closest_hit(h_rays, n, h_out), "ptb_closest_hit")
            raw = self.read(h_out, np.float32, 2 * n).reshape(n, 2)
        finally:
            self.release(h_rays); self.release(h_out)
        return raw[:, 0].copy(), raw[:, 1].copy().view(np.int32)

    def upload(self, array):
        a = np.ascontiguousarray(array)
        h = self.alloc(max(a.nbytes, 1))
        if a.nbytes:
            self.write(h, a)
        return h

    def wrap_device(self, ptr, nbytes):
        h = self.lib.ptb_wrap_device(ptr, nbytes)
        if not h:
            raise PtbError("ptb_wrap_device failed")
        return h

    def device_ptr(self, h):
        return self.lib.ptb_device_ptr(h)

    def invalidate(self, h):
        """tell the library that wrapped memory was rewritten outside of it (a torch op, NCCL, a peer GPU)"""
        self._ck(self.lib.ptb_invalidate(h), "ptb_invalidate")

    def ipc_export(self, h):
        """64-byte CUDA IPC handle of a ptb.alloc'ed buffer (bytes), for a process on another GPU of the box."""
        buf = C.create_string_buffer(64)
        self._ck(self.lib.ptb_ipc_export(h, buf), "ptb_ipc_export")
        return buf.raw

    def ipc_import(self, handle_bytes, nbytes):
        """Map another process's exported buffer; the returned handle binds as a kernel argument like a local one."""
        buf = C.create_string_buffer(bytes(handle_bytes), 64)
        h = self.lib.ptb_ipc_import(buf, nbytes)
        if not h:
            raise PtbError("ptb_ipc_import: " + self.last_error())
        return h

    # -- kernels
    def set_arg(self, kernel, index, value):
        """value: int → 4-byte int, bytes/ndarray → by-value blob (Camera)."""
        if isinstance(value, (int, np.integer)):
            v = C.c_int(int(value))
            self._ck(self.lib.ptb_set_arg(kernel, index, 4, C.byref(v)), "ptb_set_arg")
        else:
            b = bytes(value) if not isinstance(value, np.ndarray) else value.tobytes()
            buf = C.create_string_buffer(b, len(b))
            self._ck(self.lib.ptb_set_arg(kernel, index, len(b), buf), "ptb_set_arg")

    def set_arg_mem(self, kernel, index, h):
        self._ck(self.lib.ptb_set_arg_mem(kernel, index, h), "ptb_set_arg_mem")

    def launch(self, kernel, gx, gy=1):
        self._ck(self.lib.ptb_launch(kernel, gx, gy), "ptb_launch")

    def finish(self):
        self._ck(self.lib.ptb_finish(), "ptb_finish")

    def render(self, spp):
        self._ck(self.lib.ptb_render(spp), "ptb_render")

    # -- knobs
    def set_accel(self, a): self._ck(self.lib.ptb_set_accel(a), "ptb_set_accel")
    def set_math(self, m): self._ck(self.lib.ptb_set_math(m), "ptb_set_math")
    def set_pipeline(self, p): self._ck(self.lib.ptb_set_pipeline(p), "ptb_set_pipeline")
    def set_quant_nodes(self, on): self._ck(self.lib.ptb_set_quant_nodes(int(on)), "ptb_set_quant_nodes")
    def set_wf_subs(self, n): self._ck(self.lib.ptb_set_wf_subs(n), "ptb_set_wf_subs")
    def set_fuse_gen(self, on): self._ck(self.lib.ptb_set_fuse_gen(int(on)), "ptb_set_fuse_gen")
    def set_lbvh_leaf(self, n): self._ck(self.lib.ptb_set_lbvh_leaf(n), "ptb_set_lbvh_leaf")
    def set_merge_prims(self, on): self._ck(self.lib.ptb_set_merge_prims(int(on)), "ptb_set_merge_prims")
    def set_parity_filter(self, on): self._ck(self.lib.ptb_set_parity_filter(int(on)), "ptb_set_parity_filter")
    def scan_prims(self): return self.lib.ptb_scan_prims()
    def get_accel(self): return self.lib.ptb_get_accel()
    def build_accel(self): self._ck(self.lib.ptb_build_accel(), "ptb_build_accel")
    def set_pixel_shard(self, rank, n, group): self._ck(self.lib.ptb_set_pixel_shard(rank, n, group), "ptb_set_pixel_shard")
    def set_accumulate_sum(self, on): self._ck(self.lib.ptb_set_accumulate_sum(int(on)), "ptb_set_accumulate_sum")
    def resolve_sum(self, h, spp): self._ck(self.lib.ptb_resolve_sum(h, spp), "ptb_resolve_sum")
    def set_count_work(self, on): self._ck(self.lib.ptb_set_count_work(int(on)), "ptb_set_count_work")
    def reset_counters(self): self._ck(self.lib.ptb_reset_counters(), "ptb_reset_counters")
    def set_stream(self, s): self._ck(self.lib.ptb_set_stream(s), "ptb_set_stream")
    def get_stream(self): return self.lib.ptb_get_stream()
    def launch_count(self): return int(self.lib.ptb_launch_count())

    def counters(self):
        c = Counters()
        self._ck(self.lib.ptb_get_counters(C.byref(c)), "ptb_get_counters")
        return c.as_dict()

    def accel_stats(self):
        a, b, c, d = C.c_int(), C.c_int(), C.c_int(), C.c_float()
        self.lib.ptb_accel_stats(C.byref(a), C.byref(b), C.byref(c), C.byref(d))
        return {"nodes": a.value, "tris": b.value, "max_depth": c.value, "build_ms": d.value}

    def resolve_rgba8(self, colors, w, h, mode=0):
        out = np.empty((h, w, 4), dtype=np.uint8)
        self._ck(self.lib.ptb_resolve_rgba8(colors, w, h, mode, out.ctypes.data), "ptb_resolve_rgba8")
        return out

    def resolve_rgba8_noalloc(self, colors, w, h, mode=0, out=None):
        """ptb_resolve_rgba8_noalloc: the same picture through the static device scratch (no allocation on the library side);
        `out` lets the caller bring its own (h, w, 4) uint8 array so that nothing is allocated on this side either"""
        if out is None:
            out = np.empty((h, w, 4), dtype=np.uint8)
        self._ck(self.lib.ptb_resolve_rgba8_noalloc(colors, w, h, mode, out.ctypes.data), "ptb_resolve_rgba8_noalloc")
        return out

    def resolve_frame(self, colors, w, h, divide_by=1, mode=0, image=0, want_rgba8=True):
        """ptb_resolve_frame: sum -> mean (in place) -> tone-mapped float image -> RGBA8 in one pass; returns the RGBA8 array or None"""
        out = np.empty((h, w, 4), dtype=np.uint8) if want_rgba8 else None
        self._ck(self.lib.ptb_resolve_frame(colors, w, h, divide_by, mode, image, out.ctypes.data if want_rgba8 else None), "ptb_resolve_frame")
        return out

    # -- several GPUs inside this process
    def set_gpus(self, n, mode=0): self._ck(self.lib.ptb_set_gpus(n, mode), "ptb_set_gpus")
    def get_gpus(self): return self.lib.ptb_get_gpus()

    # -- one process per GPU: NCCL behind the boundary
    def comm_unique_id(self):
        buf = C.create_string_buffer(128)
        self._ck(self.lib.ptb_comm_unique_id(buf), "ptb_comm_unique_id")
        return buf.raw

    def comm_init(self, rank, nranks, unique_id):
        buf = C.create_string_buffer(bytes(unique_id), 128)
        self._ck(self.lib.ptb_comm_init(rank, nranks, buf), "ptb_comm_init")

    def comm_ranks(self): return self.lib.ptb_comm_ranks()
    def comm_rank(self): return self.lib.ptb_comm_rank()
    def comm_reduce(self, h, root=0): self._ck(self.lib.ptb_comm_reduce(h, root), "ptb_comm_reduce")
    def comm_broadcast(self, h, root=0): self._ck(self.lib.ptb_comm_broadcast(h, root), "ptb_comm_broadcast")
    def comm_barrier(self): self._ck(self.lib.ptb_comm_barrier(), "ptb_comm_barrier")
    def comm_destroy(self): self.lib.ptb_comm_destroy()

    def save_state(self, path):
        self._ck(self.lib.ptb_save_state(os.fsencode(path)), "ptb_save_state")

    def load_state(self, path):
        s = C.c_int(0)
        self._ck(self.lib.ptb_load_state(os.fsencode(path), C.byref(s)), "ptb_load_state")
        return s.value

    def filter_median(self, colors, image, w, h):
        self._ck(self.lib.ptb_filter_median(colors, image, w, h), "ptb_filter_median")

    def shutdown(self):
        self.lib.ptb_shutdown()
