"""Procedural scenes of BASELINE.json configs 3-5 (the reference ships no models: main.cpp:1002-1010
loads ../models/*.obj that are not in its repository).  Fixed generator seeds; coordinates at the
reference's scene scale (~10^3 units, the self-intersection offset is a fixed 0.001: prog.cl:217).

A scene description is a dict: mats (list of (kd, ks, emission, N, K, shininess, type)), verts (n,9)
float32, tri_mat (n,) int32, obj_end (triangle count at which each object ends: one tree per object in
the reference, main.cpp:536-551), camera (eye, fov, yaw, pitch).  ``populate`` feeds it through the
reference's Scene interface; ``write_ptbs`` writes the soup file oracle/ref_driver.cpp --tribin reads,
so the reference's own host/kernels render exactly the same triangles.
"""
import struct

import numpy as np

WHITE = ((0.73, 0.73, 0.73), (0, 0, 0), (0, 0, 0), (1, 1, 1), (0, 0, 0), 50.0, 0)
RED = ((0.65, 0.05, 0.05), (0, 0, 0), (0, 0, 0), (1, 1, 1), (0, 0, 0), 50.0, 0)
GREEN = ((0.12, 0.45, 0.15), (0, 0, 0), (0, 0, 0), (1, 1, 1), (0, 0, 0), 50.0, 0)
BLUE = ((0.15, 0.25, 0.65), (0.2, 0.2, 0.2), (0, 0, 0), (1, 1, 1), (0, 0, 0), 30.0, 0)
LIGHT = ((0, 0, 0), (0, 0, 0), (17, 12, 4), (1, 1, 1), (0, 0, 0), 0.0, 3)
MIRROR = ((0, 0, 0), (0, 0, 0), (0, 0, 0), (0.17, 0.35, 1.5), (3.1, 2.7, 1.9), 0.0, 1)
GLASS = ((0, 0, 0), (0, 0, 0), (0, 0, 0), (1.5, 1.5, 1.5), (0, 0, 0), 0.0, 2)


def quad(a, b, c, d):
    a, b, c, d = (np.asarray(v, dtype=np.float32) for v in (a, b, c, d))
    return np.stack([np.concatenate([a, b, c]), np.concatenate([a, c, d])])


def room(lo, hi, front=True):
    """6 quads = 12 triangles of an axis-aligned box (front=False: 10 triangles, open towards -z like a Cornell box)."""
    x0, y0, z0 = lo
    x1, y1, z1 = hi
    q = [quad((x0, y0, z0), (x1, y0, z0), (x1, y0, z1), (x0, y0, z1)),      # floor
         quad((x0, y1, z0), (x0, y1, z1), (x1, y1, z1), (x1, y1, z0)),      # ceiling
         quad((x0, y0, z1), (x1, y0, z1), (x1, y1, z1), (x0, y1, z1)),      # back
         quad((x0, y0, z0), (x0, y0, z1), (x0, y1, z1), (x0, y1, z0)),      # left
         quad((x1, y0, z0), (x1, y1, z0), (x1, y1, z1), (x1, y0, z1)),      # right
         quad((x0, y0, z0), (x0, y1, z0), (x1, y1, z0), (x1, y0, z0))]      # front
    return np.concatenate(q if front else q[:5])


def grid_mesh(P):
    """(nu, nv, 3) vertex grid → (2*(nu-1)*(nv-1), 9) triangles."""
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    t1 = np.concatenate([a, b, c], axis=-1).reshape(-1, 9)
    t2 = np.concatenate([a, c, d], axis=-1).reshape(-1, 9)
    return np.concatenate([t1, t2]).astype(np.float32)


def uv_sphere(center, radius, nu, nv):
    u = np.linspace(0, 2 * np.pi, nu + 1)[:, None]
    v = np.linspace(0, np.pi, nv + 1)[None, :]
    P = np.stack([np.cos(u) * np.sin(v), np.ones_like(u) * np.cos(v), np.sin(u) * np.sin(v)], axis=-1)
    t = grid_mesh(P * radius + np.asarray(center))
    # drop the degenerate pole triangles (zero area → NaN normal in the reference's Triangle ctor)
    a, b, c = t[:, 0:3], t[:, 3:6], t[:, 6:9]
    area = np.linalg.norm(np.cross(b - a, c - a), axis=1)
    return t[area > 1e-6 * radius * radius]


def torus(center, R, r, nu, nv, tilt):
    u = np.linspace(0, 2 * np.pi, nu + 1)[:, None]
    v = np.linspace(0, 2 * np.pi, nv + 1)[None, :]
    P = np.stack([(R + r * np.cos(v)) * np.cos(u), r * np.sin(v) * np.ones_like(u), (R + r * np.cos(v)) * np.sin(u)], axis=-1)
    ca, sa = np.cos(tilt), np.sin(tilt)
    rot = np.array([[1, 0, 0], [0, ca, -sa], [0, sa, ca]])
    return grid_mesh(P @ rot.T + np.asarray(center))


def spheres100k(seed=12345):
    """~100 k triangles: tessellated spheres and tori inside a closed diffuse room with a ceiling emitter."""
    rng = np.random.RandomState(seed)
    mats = [WHITE, RED, GREEN, BLUE, LIGHT, MIRROR, GLASS]
    parts, pm, obj_end = [], [], []
    r = room((0, 0, 0), (1000, 1000, 1000))
    parts.append(r); pm.append(np.array([0, 0, 0, 0, 0, 0, 1, 1, 2, 2, 0, 0]))
    parts.append(quad((200, 999.5, 200), (800, 999.5, 200), (800, 999.5, 800), (200, 999.5, 800))); pm.append(np.array([4, 4]))
    obj_end.append(14)
    n = 14
    for i in range(24):
        c = rng.uniform([120, 100, 150], [880, 700, 880])
        rad = rng.uniform(45, 95)
        t = uv_sphere(c, rad, 48, 32)
        parts.append(t); pm.append(np.full(len(t), [0, 1, 2, 3, 3, 0, 5, 6][i % 8]))
        n += len(t); obj_end.append(n)
    for i in range(8):
        c = rng.uniform([150, 120, 150], [850, 600, 850])
        t = torus(c, rng.uniform(70, 110), rng.uniform(18, 32), 64, 32, rng.uniform(0, np.pi))
        parts.append(t); pm.append(np.full(len(t), [3, 0, 1, 2][i % 4]))
        n += len(t); obj_end.append(n)
    return {"mats": mats, "verts": np.concatenate(parts).astype(np.float32), "tri_mat": np.concatenate(pm).astype(np.int32),
            "obj_end": obj_end, "eye": (500.0, 500.0, 5.0), "fov": 70.0, "yaw": 0.0, "pitch": 0.0}


def grid1m(seed=2468, nx=1000, nz=500):
    """1 000 000 triangles: a displaced 1000x500-cell height field under a large emitter, side walls keep rays in."""
    rng = np.random.RandomState(seed)
    x = np.linspace(0, 2000, nx + 1)[:, None]
    z = np.linspace(0, 1000, nz + 1)[None, :]
    h = np.zeros((nx + 1, nz + 1))
    for _ in range(6):
        fx, fz, ph, amp = rng.uniform(0.004, 0.05), rng.uniform(0.004, 0.05), rng.uniform(0, 6.28), rng.uniform(4, 40)
        h += amp * np.sin(fx * x + ph) * np.cos(fz * z + 0.7 * ph)
    h += rng.uniform(-0.4, 0.4, size=h.shape)        # breaks exact coplanarity / coincident centroids
    P = np.stack([x * np.ones_like(z), 150 + h, z * np.ones_like(x)], axis=-1)
    g = grid_mesh(P)
    mats = [WHITE, RED, GREEN, BLUE, LIGHT]
    walls = room((0, 0, 0), (2000, 900, 1000))
    light = quad((300, 899.5, 200), (1700, 899.5, 200), (1700, 899.5, 800), (300, 899.5, 800))
    gm = (np.arange(len(g)) // 4096 % 3 == 0).astype(np.int32) * 3      # stripes of the glossy material
    verts = np.concatenate([walls, light, g]).astype(np.float32)
    tri_mat = np.concatenate([np.array([0, 0, 0, 0, 0, 0, 1, 1, 2, 2, 0, 0]), np.array([4, 4]), gm]).astype(np.int32)
    return {"mats": mats, "verts": verts, "tri_mat": tri_mat, "obj_end": [14, 14 + len(g)],
            "eye": (1000.0, 620.0, 6.0), "fov": 75.0, "yaw": 0.0, "pitch": 25.0}


def small_mixed(seed=7, n_spheres=3):
    """A few hundred triangles with every material type (mirror, glass, glossy): parity coverage of prog.cl:341-357."""
    rng = np.random.RandomState(seed)
    mats = [WHITE, RED, GREEN, BLUE, LIGHT, MIRROR, GLASS]
    parts = [room((0, 0, 0), (555, 555, 555), front=False), quad((213, 554.5, 227), (343, 554.5, 227), (343, 554.5, 332), (213, 554.5, 332))]
    pm = [np.array([0, 0, 0, 0, 3, 3, 1, 1, 2, 2]), np.array([4, 4])]
    obj_end = [12]
    n = 12
    for i in range(n_spheres):
        c = rng.uniform([120, 90, 150], [430, 300, 430])
        t = uv_sphere(c, rng.uniform(55, 85), 14, 10)
        parts.append(t); pm.append(np.full(len(t), [5, 6, 3][i % 3]))
        n += len(t); obj_end.append(n)
    return {"mats": mats, "verts": np.concatenate(parts).astype(np.float32), "tri_mat": np.concatenate(pm).astype(np.int32),
            "obj_end": obj_end, "eye": (277.5, 277.5, -800.0), "fov": 39.3, "yaw": 0.0, "pitch": 0.0}


GENERATORS = {"spheres100k": spheres100k, "grid1m": grid1m, "small_mixed": small_mixed}


def describe(name):
    return GENERATORS[name]()


def populate(scene, name_or_desc):
    """Feed a description through the reference's Scene interface; returns the eye position."""
    from . import host
    d = describe(name_or_desc) if isinstance(name_or_desc, str) else name_or_desc
    off = len(scene.mats)
    for m in d["mats"]:
        scene.add_Material(host.Material(*m))
    v = d["verts"]
    start = 0
    for end in d["obj_end"]:
        scene.add_Triangle(host.Triangles(v[start:end, 0:3], v[start:end, 3:6], v[start:end, 6:9], off + d["tri_mat"][start:end]))
        scene.end_Obj()
        start = end
    scene.global_fov, scene.global_yaw, scene.global_pitch = d["fov"], d["yaw"], d["pitch"]
    return np.asarray(d["eye"], dtype=np.float32)


def write_ptbs(d, path):
    """The soup file oracle/ref_driver.cpp --tribin loads through the reference's add_Material/add_Triangle/end_Obj."""
    mats = d["mats"]
    with open(path, "wb") as f:
        f.write(b"PTBS")
        f.write(struct.pack("<4i", 1, len(mats), len(d["verts"]), len(d["obj_end"])))
        m = np.zeros((len(mats), 16), dtype=np.float32)
        for i, (kd, ks, ke, N, K, sh, _tp) in enumerate(mats):
            m[i, 0:3], m[i, 3:6], m[i, 6:9], m[i, 9:12], m[i, 12:15], m[i, 15] = kd, ks, ke, N, K, sh
        f.write(m.tobytes())
        f.write(np.asarray([mt[6] for mt in mats], dtype=np.int32).tobytes())
        f.write(np.ascontiguousarray(d["verts"], dtype=np.float32).tobytes())
        f.write(np.ascontiguousarray(d["tri_mat"], dtype=np.int32).tobytes())
        f.write(np.asarray(d["obj_end"], dtype=np.int32).tobytes())
