"""Host-side mirror of the reference's ``Scene`` interface (main.cpp:363-742) over the C-ABI.

Same method names, argument meaning and call order as the reference host: ``init_Scene``,
``add_Material``, ``add_Triangle``, ``end_Obj``, ``add_Obj``, ``upload_Triangles``,
``upload_Materials``, ``generate_rays``, ``trace_rays``, ``render`` — so a test written against the
reference's Scene reads the same here.  The records built here are the kernel ABI structs byte for
byte (prog.cl:1-35): tests compare them with dumps written by the reference's own host code.

All float arithmetic follows main.cpp's types: ``float`` expressions in float32, the unqualified
``cos/sin/tan/sqrt`` calls in double (``<cmath>`` only puts the double versions in the global
namespace) narrowed to float on assignment.
"""
import os

import numpy as np

from . import capi

# ---------------------------------------------------------------- kernel ABI records (prog.cl:1-35)
MATERIAL = np.dtype([("kd", "<f4", 4), ("ks", "<f4", 4), ("emission", "<f4", 4), ("F0", "<f4", 4),
                     ("n", "<f4"), ("shininess", "<f4"), ("type", "<i4"), ("pad", "<i4")])
TRIANGLE = np.dtype([("r1", "<f4", 4), ("r2", "<f4", 4), ("r3", "<f4", 4), ("N", "<f4", 4),
                     ("mati", "<u2"), ("pad", "<u2", 7)])
NODE = np.dtype([("trii", "<i4", 2), ("pad", "<i4", 2), ("bl", "<f4", 4), ("tr", "<f4", 4)])
CAMERA = np.dtype([("eye", "<f4", 4), ("lookat", "<f4", 4), ("up", "<f4", 4), ("right", "<f4", 4),
                   ("XM", "<f4"), ("YM", "<f4"), ("pad", "<f4", 2)])
RAY = np.dtype([("P", "<f4", 4), ("D", "<f4", 4)])
assert MATERIAL.itemsize == 80 and TRIANGLE.itemsize == 80 and NODE.itemsize == 48 and CAMERA.itemsize == 80 and RAY.itemsize == 32

f32 = np.float32
M31 = 2147483647


def minstd_rand0_seeds(count, skip=0):
    """Outputs ``skip .. skip+count-1`` of a default-constructed ``std::minstd_rand0`` (multiplier
    16807, modulus 2^31-1, seed 1) — the per-pixel RNG seeds of main.cpp:45,522-526 (pixel i gets
    the i-th output).  ``skip`` continues the sequence for further ranks of an spp-split.
    Computed by jump-ahead: x_k = 16807^(k+1) mod M."""
    if count <= 0:
        return np.zeros(0, dtype=np.int32)
    block = 4096
    tab = np.empty(block, dtype=np.uint64)
    x = 1
    for i in range(block):
        tab[i] = x
        x = (x * 16807) % M31
    a_block = x                                   # 16807^block
    n_rows = (count + block - 1) // block
    row = pow(16807, skip + 1, M31)               # first wanted output
    rows = np.empty(n_rows, dtype=np.uint64)
    for j in range(n_rows):
        rows[j] = row
        row = (row * a_block) % M31
    out = (rows[:, None] * tab[None, :]) % np.uint64(M31)   # products < 2^62
    return out.reshape(-1)[:count].astype(np.int32)


def _rot(v, deg, axis):
    """rotate_x / rotate_y of main.cpp:56-70 on an (n,3) float32 array: angle in float, cos/sin in double,
    the products in double, narrowed to float."""
    a = f32(f32(f32(deg) / f32(180.0)) * f32(3.141593))
    c, s = np.cos(np.float64(a)), np.sin(np.float64(a))
    v = np.asarray(v, dtype=np.float32).reshape(-1, 3)
    x, y, z = (v[:, i].astype(np.float64) for i in range(3))
    if axis == "x":
        r = np.stack([v[:, 0], (y * c - z * s).astype(np.float32), (y * s + z * c).astype(np.float32)], axis=1)
    else:
        r = np.stack([(x * c + z * s).astype(np.float32), v[:, 1], (-x * s + z * c).astype(np.float32)], axis=1)
    return r.astype(np.float32)


def rotate_x(v, deg):
    return _rot(v, deg, "x")


def rotate_y(v, deg):
    return _rot(v, deg, "y")


def Material(kd, ks, emission, N, K, shininess, type):
    """main.cpp:101-111 — one 80-byte Material record; N = refractive index, K = extinction per channel."""
    m = np.zeros((), dtype=MATERIAL)
    N = np.asarray(N, dtype=np.float32)
    K = np.asarray(K, dtype=np.float32)
    m["kd"][:3] = kd
    m["ks"][:3] = ks
    m["emission"][:3] = emission
    m["n"] = f32(f32(f32(N[0] + N[1]) + N[2]) / f32(3.0))
    one = f32(1.0)
    a = (N - one) * (N - one)
    b = (N + one) * (N + one)
    with np.errstate(divide="ignore", invalid="ignore"):
        m["F0"][:3] = (K * K + a) / (K * K + b)
    m["shininess"] = shininess
    m["type"] = type
    return m


def Triangles(r1, r2, r3, mati):
    """main.cpp:144-166 vectorised — (n,) Triangle records with the host-computed unit normal."""
    r1, r2, r3 = (np.asarray(a, dtype=np.float32).reshape(-1, 3) for a in (r1, r2, r3))
    n = r1.shape[0]
    t = np.zeros(n, dtype=TRIANGLE)
    t["r1"][:, :3], t["r2"][:, :3], t["r3"][:, :3] = r1, r2, r3
    v1, v2 = r2 - r1, r3 - r1
    nx = v1[:, 1] * v2[:, 2] - v1[:, 2] * v2[:, 1]
    ny = v1[:, 2] * v2[:, 0] - v1[:, 0] * v2[:, 2]
    nz = v1[:, 0] * v2[:, 1] - v1[:, 1] * v2[:, 0]
    with np.errstate(divide="ignore", invalid="ignore"):
        length = np.sqrt(((nx * nx + ny * ny) + nz * nz).astype(np.float64)).astype(np.float32)
        t["N"][:, 0], t["N"][:, 1], t["N"][:, 2] = nx / length, ny / length, nz / length
    t["mati"] = np.asarray(mati).astype(np.int64).astype(np.uint16)     # cl_ushort conversion wraps
    return t


def Triangle(r1, r2, r3, mati):
    return Triangles([r1], [r2], [r3], [mati])[0]


def camera_step(width, height, fov, yaw, pitch, shift, move=(0.0, 0.0, 0.0)):
    """main.cpp:311-347 — the Camera constructor INCLUDING its side effect: it integrates the interactive motion
    (global_forward, global_rightward, global_upward) into global_shift along the rotated unit basis before the basis
    is scaled.  Returns (Camera record, new global_shift).  All arithmetic in float, cos/sin/tan in double."""
    c = np.zeros((), dtype=CAMERA)
    XM, YM = f32(width), f32(height)
    up_len, right_len = f32(YM / f32(2.0)), f32(XM / f32(2.0))
    ang = f32(f32(f32(f32(fov) / f32(2.0)) / f32(180.0)) * f32(3.141593))
    ahead_len = f32(np.float64(right_len) / np.tan(np.float64(ang)))
    basis = np.array([[0, 1, 0], [1, 0, 0], [0, 0, 1]], dtype=np.float32)    # up, right, ahead
    basis = rotate_y(rotate_x(basis, pitch), yaw)
    fwd, rgt, upw = (f32(m) for m in move)
    shift = np.asarray(shift, dtype=np.float32).copy()
    # global_shift[i] = global_shift[i] + ahead[i]*forward + right[i]*rightward + up[i]*upward  (left to right, float)
    shift = ((shift + basis[2] * fwd) + basis[1] * rgt) + basis[0] * upw
    up, right, ahead = basis[0] * up_len, basis[1] * right_len, basis[2] * ahead_len
    eye = np.array([f32(500.0) + shift[0], f32(500.0) + shift[1], f32(-1299.037842) + shift[2]], dtype=np.float32)
    c["eye"][:3], c["up"][:3], c["right"][:3], c["lookat"][:3] = eye, up, right, eye + ahead
    c["XM"], c["YM"] = XM, YM
    return c, shift.astype(np.float32)


def Camera(width, height, fov, yaw, pitch, shift):
    """main.cpp:311-347 with the interactive globals at rest (global_forward/rightward/upward = 0)."""
    return camera_step(width, height, fov, yaw, pitch, shift)[0]


def camera_at(width, height, fov, yaw, pitch, eye):
    """Camera with a given eye position (the reference expresses it as a shift from (500,500,-1299.037842))."""
    eye = np.asarray(eye, dtype=np.float32)
    shift = np.array([eye[0] - f32(500.0), eye[1] - f32(500.0), eye[2] + f32(1299.037842)], dtype=np.float32)
    return Camera(width, height, fov, yaw, pitch, shift)


# ---------------------------------------------------------------- host BVH (main.cpp:195-304, 536-551)
def _bbox(t):
    v = np.stack([t["r1"][:, :3], t["r2"][:, :3], t["r3"][:, :3]])
    return v.min(axis=0), v.max(axis=0)


def _midpoints(t):
    return ((t["r1"][:, :3] + t["r2"][:, :3]) + t["r3"][:, :3]) / f32(3.0)


def build_object_tree(tris, tri_shift, kd_tree, kd_shift):
    """NodeOnHost::build + convert for one object: median-of-centroid-mean split on axis depth%3
    (next axis while a side is empty), leaves of <= 6 triangles, BFS flatten into the implicit heap
    at ``kd_tree[k + kd_shift]``; returns the triangles in leaf order.  ``kd_tree`` is a python list of
    (from, to, bl, tr) tuples that is extended in place."""
    lo_all, hi_all = _bbox(tris)
    mids = _midpoints(tris)

    def build(idx, depth):
        lo, hi = lo_all[idx].min(axis=0), hi_all[idx].max(axis=0)
        if len(idx) <= 6:
            return (idx, None, None, lo, hi)
        # midpoint accumulates in float, one triangle at a time (main.cpp:224-234)
        acc = np.add.accumulate(mids[idx], axis=0, dtype=np.float32)[-1]
        mean = acc / f32(len(idx))
        axis = depth % 3
        for _ in range(4):
            right = mean[axis] >= mids[idx, axis]
            if right.any() and (~right).any():
                break
            axis = (axis + 1) % 3
        else:
            raise ValueError("coincident centroids: the reference builder does not terminate on this input (main.cpp:246-257)")
        return (None, build(idx[~right], depth + 1), build(idx[right], depth + 1), lo, hi)

    root = build(np.arange(len(tris)), 0)
    order = []
    frm = to = tri_shift
    queue = [(root, 1)]
    qi = 0
    while qi < len(queue):
        node, ptr = queue[qi]
        qi += 1
        idx, left, right, lo, hi = node
        while len(kd_tree) <= ptr + kd_shift:
            kd_tree.append((-1, -1, lo, hi))
        if idx is not None:
            to += len(idx)
            kd_tree[ptr + kd_shift] = (frm, to, lo, hi)
            order.extend(idx.tolist())
            frm += len(idx)
        else:
            kd_tree[ptr + kd_shift] = (-1, -1, lo, hi)
            queue.append((left, ptr * 2))
            queue.append((right, ptr * 2 + 1))
    return tris[np.asarray(order, dtype=np.int64)]


def nodes_to_records(kd_tree):
    n = np.zeros(len(kd_tree), dtype=NODE)
    for i, (a, b, lo, hi) in enumerate(kd_tree):
        n["trii"][i] = (a, b)
        n["bl"][i, :3] = lo
        n["tr"][i, :3] = hi
    return n


# ---------------------------------------------------------------- OBJ / MTL (main.cpp:552-617 contract)
def load_mtl(path):
    """The subset of MTL the reference consumes: Kd Ks Ke Ns + its custom keys Kn Kk Tp (main.cpp:564-571)."""
    mats, cur = [], None
    with open(path) as f:
        for line in f:
            p = line.split()
            if not p or p[0].startswith("#"):
                continue
            if p[0] == "newmtl":
                cur = {"name": p[1], "Kd": (0, 0, 0), "Ks": (0, 0, 0), "Ke": (0, 0, 0), "Ns": 1.0}
                mats.append(cur)
            elif cur is not None and p[0] in ("Kd", "Ks", "Ke", "Kn", "Kk"):
                cur[p[0]] = tuple(float(v) for v in p[1:4])
            elif cur is not None and p[0] == "Ns":
                cur["Ns"] = float(p[1])
            elif cur is not None and p[0] == "Tp":
                cur["Tp"] = int(p[1])
    for m in mats:
        for k in ("Kn", "Kk", "Tp"):
            if k not in m:
                raise KeyError(f"material {m['name']}: mandatory key {k} missing (main.cpp:568-571 throws here)")
    return mats


def load_obj(path):
    """Vertices, and per shape (``g``/``o`` group) the fan-triangulated faces with material ids."""
    verts, shapes, cur, mat_names, mtl, cur_mat = [], [], None, {}, [], -1
    base = os.path.dirname(path)
    with open(path) as f:
        for line in f:
            p = line.split()
            if not p or p[0].startswith("#"):
                continue
            if p[0] == "v":
                verts.append([float(p[1]), float(p[2]), float(p[3])])
            elif p[0] == "mtllib":
                mtl = load_mtl(os.path.join(base, p[1]))
                mat_names = {m["name"]: i for i, m in enumerate(mtl)}
            elif p[0] == "usemtl":
                cur_mat = mat_names.get(p[1], -1)
            elif p[0] in ("g", "o"):
                cur = {"faces": [], "mats": []}
                shapes.append(cur)
            elif p[0] == "f":
                if cur is None:
                    cur = {"faces": [], "mats": []}
                    shapes.append(cur)
                idx = []
                for tok in p[1:]:
                    i = int(tok.split("/")[0])
                    idx.append(i - 1 if i > 0 else len(verts) + i)
                for k in range(1, len(idx) - 1):
                    cur["faces"].append((idx[0], idx[k], idx[k + 1]))
                    cur["mats"].append(cur_mat)
    shapes = [s for s in shapes if s["faces"]]
    return np.asarray(verts, dtype=np.float32), shapes, mtl


# ---------------------------------------------------------------- the Scene mirror
class Scene:
    """Mirror of ``class Scene`` (main.cpp:363-742).  ``screen_width/height``, ``iterations`` and the camera
    globals of main.cpp:20-43 are attributes instead of compile-time / global state."""

    def __init__(self, ptb=None, device=-1):
        self.ptb = ptb or capi.Ptb(device)
        self.tris = np.zeros(0, dtype=TRIANGLE)
        self._pending = []
        self._device_meshes = []
        self.tri_shift = 0
        self.kd_tree = []
        self.kd_tree_shift = []
        self.mats = []
        self.iterations = 1                 # main.cpp:27
        self.current_sample = 0             # main.cpp:28
        self.global_fov, self.global_yaw, self.global_pitch = 75.0, -13.800002 - 50, 5.599997 + 10     # main.cpp:30-32
        self.global_shift = np.array([265.055481, 162.305969, 360.414001], dtype=np.float32)            # main.cpp:39
        self.global_forward = self.global_rightward = self.global_upward = 0.0                          # main.cpp:33-35
        self.build_trees = True             # False: skip the host trees (they are optional for the CUDA path)
        self.buffer_tris = self.buffer_mats = self.buffer_kd_tree = self.buffer_kd_tree_shift = 0
        self.buffer_rays = self.buffer_rnds = self.buffer_colors = self.cl_screen = 0

    # main.cpp:456-527
    def init_Scene(self, screen_width, screen_height, seed_skip=0):
        p = self.ptb
        self.screen_width, self.screen_height = int(screen_width), int(screen_height)
        self.rays_size = self.screen_width * self.screen_height
        self.buffer_rays = p.alloc(RAY.itemsize * self.rays_size)
        self.buffer_rnds = p.alloc(4 * self.rays_size)
        self.cl_screen = p.alloc_image(self.screen_width, self.screen_height)
        self.buffer_colors = p.alloc(16 * self.rays_size)
        p.write(self.buffer_rnds, minstd_rand0_seeds(self.rays_size, seed_skip))

    def add_Triangle(self, tri):
        self._pending.append(np.asarray(tri, dtype=TRIANGLE).reshape(-1))

    def add_Material(self, mat):
        self.mats.append(np.asarray(mat, dtype=MATERIAL).reshape(()))
        return len(self.mats) - 1

    def _flush(self):
        if self._pending:
            self.tris = np.concatenate([self.tris] + self._pending)
            self._pending = []

    # main.cpp:536-551
    def end_Obj(self):
        self._flush()
        if self.build_trees:
            shift = 0 if not self.kd_tree_shift else len(self.kd_tree) - 1
            self.kd_tree_shift.append(shift)
            obj = self.tris[self.tri_shift:]
            self.tris[self.tri_shift:] = build_object_tree(obj, self.tri_shift, self.kd_tree, shift)
        self.tri_shift = len(self.tris)

    # main.cpp:552-617
    def add_Obj(self, file, pos, scale, pitch, yaw):
        verts, shapes, mtl = load_obj(file)
        mat_offset = len(self.mats)
        for m in mtl:
            self.add_Material(Material(m["Kd"], m["Ks"], m["Ke"], m["Kn"], m["Kk"], m["Ns"], m["Tp"]))
        v = verts.copy()
        v[:, 0] = -v[:, 0]                                            # main.cpp:598
        v = rotate_y(rotate_x(v, pitch), yaw)
        v = v * np.asarray(scale, dtype=np.float32) + np.asarray(pos, dtype=np.float32)
        for s in shapes:
            f = np.asarray(s["faces"], dtype=np.int64)
            self.add_Triangle(Triangles(v[f[:, 0]], v[f[:, 1]], v[f[:, 2]], mat_offset + np.asarray(s["mats"])))
            self.end_Obj()

    def add_Obj_device(self, file, pos, scale, pitch, yaw):
        """add_Obj with the vertex transform, the normals and the Triangle records computed ON THE DEVICE
        (ptb_ingest_mesh) at upload_Triangles(): only the OBJ/MTL text parse stays on the host.  The records are
        bit-identical to add_Obj's, in face order; no reference tree is built for them (the LBVH takes them)."""
        verts, shapes, mtl = load_obj(file)
        mat_offset = len(self.mats)
        for m in mtl:
            self.add_Material(Material(m["Kd"], m["Ks"], m["Ke"], m["Kn"], m["Kk"], m["Ns"], m["Tp"]))
        for s in shapes:
            self.add_Mesh_device(verts, s["faces"], s["mats"], pos, scale, pitch, yaw, mat_offset)

    def add_Mesh_device(self, verts, faces, mat_ids, pos=(0, 0, 0), scale=(1, 1, 1), pitch=0.0, yaw=0.0, mat_offset=0):
        self._device_meshes.append((np.asarray(verts, dtype=np.float32), np.asarray(faces, dtype=np.int32),
                                    np.asarray(mat_ids, dtype=np.int32), pos, scale, pitch, yaw, mat_offset))

    # main.cpp:618-630
    def upload_Triangles(self):
        self._flush()
        p = self.ptb
        if self._device_meshes:
            return self._upload_with_device_ingest()
        for h in (self.buffer_kd_tree, self.buffer_kd_tree_shift, self.buffer_tris):
            if h:
                p.release(h)
        nodes = nodes_to_records(self.kd_tree)
        self.buffer_kd_tree = p.upload(nodes)
        self.kd_tree_shift_size = len(self.kd_tree_shift)
        self.buffer_kd_tree_shift = p.upload(np.asarray(self.kd_tree_shift, dtype=np.int32))
        self.tris_size = len(self.tris)
        self.buffer_tris = p.upload(self.tris)

    def _upload_with_device_ingest(self):
        """host-built records first (their reference trees are dropped: indices into a longer array would still be
        valid, but a scene that mixes both is traversed by the LBVH as one), then every device mesh in call order"""
        p = self.ptb
        for h in (self.buffer_kd_tree, self.buffer_kd_tree_shift, self.buffer_tris):
            if h:
                p.release(h)
        n_dev = sum(len(m[1]) for m in self._device_meshes)
        self.tris_size = len(self.tris) + n_dev
        self.buffer_tris = p.alloc(max(self.tris_size, 1) * TRIANGLE.itemsize)
        if len(self.tris):
            p.write(self.buffer_tris, self.tris)
        first = len(self.tris)
        for verts, faces, mats, pos, scale, pitch, yaw, mat_offset in self._device_meshes:
            p.ingest_mesh(self.buffer_tris, first, verts, faces, mats, pos, scale, pitch, yaw, mat_offset)
            first += len(faces)
        self.buffer_kd_tree = p.upload(np.zeros(1, dtype=NODE))
        self.kd_tree_shift_size = 0
        self.buffer_kd_tree_shift = p.upload(np.zeros(1, dtype=np.int32))

    # main.cpp:631-634
    def upload_Materials(self):
        if self.buffer_mats:
            self.ptb.release(self.buffer_mats)
        self.buffer_mats = self.ptb.upload(np.stack(self.mats) if self.mats else np.zeros(0, dtype=MATERIAL))

    def _camera(self):
        """`Camera cam;` of main.cpp:636: constructing it moves global_shift by the pending interactive motion"""
        cam, self.global_shift = camera_step(self.screen_width, self.screen_height, self.global_fov, self.global_yaw,
                                             self.global_pitch, self.global_shift,
                                             (self.global_forward, self.global_rightward, self.global_upward))
        return cam

    # main.cpp:635-643
    def generate_rays(self):
        p, K = self.ptb, capi.KERNEL_GEN_RAY
        self.camera = self._camera()
        p.set_arg_mem(K, 0, self.buffer_rays)
        p.set_arg(K, 1, self.camera.tobytes())
        p.set_arg_mem(K, 2, self.buffer_rnds)
        p.launch(K, self.rays_size)

    def _set_trace_args(self):
        p, K = self.ptb, capi.KERNEL_TRACE_RAY
        p.set_arg_mem(K, 0, self.cl_screen)
        p.set_arg_mem(K, 1, self.buffer_tris)
        p.set_arg(K, 2, self.tris_size)
        p.set_arg_mem(K, 3, self.buffer_mats)
        p.set_arg_mem(K, 4, self.buffer_kd_tree)
        p.set_arg_mem(K, 5, self.buffer_kd_tree_shift)
        p.set_arg(K, 6, self.kd_tree_shift_size)
        p.set_arg_mem(K, 7, self.buffer_rays)
        p.set_arg_mem(K, 8, self.buffer_rnds)
        p.set_arg(K, 9, self.iterations)
        p.set_arg(K, 10, self.current_sample)
        p.set_arg(K, 11, self.camera.tobytes())
        p.set_arg_mem(K, 12, self.buffer_colors)

    # main.cpp:645-682
    def trace_rays(self):
        self._set_trace_args()
        self.ptb.launch(capi.KERNEL_TRACE_RAY, self.screen_width, self.screen_height)
        if self.current_sample % 3 == 0:
            self.ptb.finish()

    # main.cpp:683-687
    def render(self):
        self.generate_rays()
        self.trace_rays()
        self.current_sample += 1

    # ---- extension: the in-kernel sample loop (ptb_render); same result as `spp` calls of render()
    def render_spp(self, spp):
        self.camera = self._camera()
        self._set_trace_args()
        self.ptb.render(spp)
        self.current_sample += spp

    # ---- headless read-back (dead code in the reference: main.cpp:727-741)
    def download_colors(self):
        c = self.ptb.read(self.buffer_colors, np.float32, self.rays_size * 4)
        return c.reshape(self.screen_height, self.screen_width, 4)

    def download_rnds(self):
        return self.ptb.read(self.buffer_rnds, np.int32, self.rays_size)

    def download_image(self):
        c = self.ptb.read(self.cl_screen, np.float32, self.rays_size * 4)
        return c.reshape(self.screen_height, self.screen_width, 4)

    # ---- headless stand-in for the window: tone-mapped 8-bit PPM and linear PFM of the accumulation buffer
    def save_ppm(self, path, mode=0):
        write_ppm(path, self.ptb.resolve_rgba8(self.buffer_colors, self.screen_width, self.screen_height, mode))

    def save_pfm(self, path):
        write_pfm(path, self.download_colors()[..., :3])

    def save_exr(self, path):
        write_exr(path, self.download_colors()[..., :3])


def write_ppm(path, rgba8):
    """binary PPM (P6) from an (H, W, 3|4) uint8 image whose first row is the TOP of the picture"""
    img = np.ascontiguousarray(np.asarray(rgba8, dtype=np.uint8)[..., :3])
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (img.shape[1], img.shape[0]))
        f.write(img.tobytes())


def write_pfm(path, rgb):
    """little-endian colour PFM from an (H, W, 3) float32 image whose first row is the BOTTOM of the picture — which
    is both PFM's convention and the reference's (row 0 = bottom scanline, prog.cl:86-88)"""
    img = np.ascontiguousarray(np.asarray(rgb, dtype="<f4"))
    with open(path, "wb") as f:
        f.write(b"PF\n%d %d\n-1.0\n" % (img.shape[1], img.shape[0]))
        f.write(img.tobytes())


def write_exr(path, rgb):
    """OpenEXR 2 scanline file, three 32-bit float channels, no compression, from an (H, W, 3) float32 image whose first
    row is the BOTTOM of the picture (the reference's convention, prog.cl:86-88; EXR's y = 0 is the top, so rows are
    flipped).  Linear radiance as accumulated — what the reference could only show tone-mapped in its window."""
    import struct
    img = np.asarray(rgb, dtype="<f4")[::-1, :, :3]
    h, w = img.shape[:2]

    def attr(name, typ, payload):
        return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<i", len(payload)) + payload

    chlist = b"".join(c + b"\0" + struct.pack("<iBxxxii", 2, 0, 1, 1) for c in (b"B", b"G", b"R")) + b"\0"   # FLOAT, sampling 1 x 1
    box = struct.pack("<iiii", 0, 0, w - 1, h - 1)
    header = (struct.pack("<ii", 20000630, 2)
              + attr("channels", "chlist", chlist) + attr("compression", "compression", b"\0")
              + attr("dataWindow", "box2i", box) + attr("displayWindow", "box2i", box)
              + attr("lineOrder", "lineOrder", b"\0") + attr("pixelAspectRatio", "float", struct.pack("<f", 1.0))
              + attr("screenWindowCenter", "v2f", struct.pack("<ff", 0.0, 0.0))
              + attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) + b"\0")
    row_bytes = 3 * w * 4
    first = len(header) + 8 * h
    offsets = (first + (8 + row_bytes) * np.arange(h, dtype=np.uint64)).astype("<u8")
    planes = np.ascontiguousarray(img[:, :, ::-1].transpose(0, 2, 1))        # per scanline: all B, all G, all R
    with open(path, "wb") as f:
        f.write(header)
        f.write(offsets.tobytes())
        for y in range(h):
            f.write(struct.pack("<ii", y, row_bytes))
            f.write(planes[y].tobytes())


# ---------------------------------------------------------------- dumps of the kernel-argument surface
def load_ptbd(path):
    """'PTBD' file written by oracle/ref_driver.cpp --dump: the buffers exactly as the reference host uploads them."""
    raw = open(path, "rb").read()
    if raw[:4] != b"PTBD":
        raise ValueError(f"{path}: not a PTBD file")
    hdr = np.frombuffer(raw, dtype="<i4", count=8, offset=4)
    ver, W, H, iters, n_tris, n_mats, n_nodes, n_obj = (int(v) for v in hdr)
    off = 36
    cam = np.frombuffer(raw, dtype=CAMERA, count=1, offset=off)[0]; off += 80
    tris = np.frombuffer(raw, dtype=TRIANGLE, count=n_tris, offset=off); off += 80 * n_tris
    mats = np.frombuffer(raw, dtype=MATERIAL, count=n_mats, offset=off); off += 80 * n_mats
    nodes = np.frombuffer(raw, dtype=NODE, count=n_nodes, offset=off); off += 48 * n_nodes
    shift = np.frombuffer(raw, dtype="<i4", count=n_obj, offset=off)
    return {"W": W, "H": H, "iterations": iters, "camera": cam, "tris": tris, "mats": mats, "kd_tree": nodes, "kd_tree_shift": shift}


def scene_from_surface(surface, ptb=None, width=None, height=None, seed_skip=0):
    """A Scene whose device buffers hold a dumped kernel-argument surface verbatim (no host recomputation)."""
    s = Scene(ptb)
    W, H = width or surface["W"], height or surface["H"]
    s.init_Scene(W, H, seed_skip)
    s.tris = np.array(surface["tris"])
    s.mats = [m for m in surface["mats"]]
    s.kd_tree_records = np.array(surface["kd_tree"])
    s.kd_tree_shift = [int(v) for v in surface["kd_tree_shift"]]
    p = s.ptb
    s.buffer_kd_tree = p.upload(s.kd_tree_records)
    s.kd_tree_shift_size = len(s.kd_tree_shift)
    s.buffer_kd_tree_shift = p.upload(np.asarray(s.kd_tree_shift, dtype=np.int32))
    s.tris_size = len(s.tris)
    s.buffer_tris = p.upload(s.tris)
    s.upload_Materials()
    s.iterations = surface["iterations"]
    cam = np.array(surface["camera"])
    if (W, H) != (surface["W"], surface["H"]):
        raise ValueError("a dumped camera is tied to its resolution")
    s._camera = lambda: cam
    return s
