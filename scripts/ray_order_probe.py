#!/usr/bin/env python
"""How much would ray REORDERING buy k_wf_trace?  Upper bound, measured through ptb_closest_hit (device-resident ray list, the
wavefront trace kernel): the same 2 M bounce-like rays (origins on the scene's surfaces, directions uniform on the sphere) in
four orders — A: by Morton code of the origin (what the queue's pixel order gives a bounce), B: by direction octant, then origin
Morton, C: by coarse origin cell (4096 cells), then octant, then fine Morton, D: shuffled.  Host clock around blocking calls.
  python scripts/ray_order_probe.py [scene ...]"""
import os, sys, time
import numpy as np
sys.path.insert(0, ".")
from opencl_path_tracer_b200 import capi, host, scenes


def morton3(q):                       # q: (n,3) uint32 in [0, 1024)
    def spread(v):
        v = v.astype(np.uint64) & 0x3ff
        v = (v | (v << 16)) & 0x30000ff
        v = (v | (v << 8)) & 0x300f00f
        v = (v | (v << 4)) & 0x30c30c3
        v = (v | (v << 2)) & 0x9249249
        return v
    return (spread(q[:, 0]) << 2) | (spread(q[:, 1]) << 1) | spread(q[:, 2])


def main():
    names = sys.argv[1:] or ["spheres100k", "grid1m"]
    ptb = capi.Ptb(0)
    ptb.set_math(capi.MATH_FAST); ptb.set_accel(capi.ACCEL_LBVH)
    rng = np.random.default_rng(5)
    n = 2_000_000
    for name in names:
        d = scenes.describe(name)
        s = host.Scene(ptb); s.build_trees = False
        s.init_Scene(64, 64); scenes.populate(s, d); s._flush()
        s.upload_Triangles(); s.upload_Materials(); s.iterations = 1
        cam = host.camera_at(64, 64, 60.0, 0, 0, (0.0, 0.0, -10.0)); s._camera = lambda: cam; s.camera = cam; s._set_trace_args()
        t = s.tris[rng.integers(0, len(s.tris), n)]
        u, v = rng.random(n, dtype=np.float32), rng.random(n, dtype=np.float32)
        f = u + v > 1; u[f], v[f] = 1 - u[f], 1 - v[f]
        a, b, c = t["r1"][:, :3], t["r2"][:, :3], t["r3"][:, :3]
        o = (a + (b - a) * u[:, None] + (c - a) * v[:, None]).astype(np.float32)
        dirs = rng.normal(size=(n, 3)).astype(np.float32); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
        o += dirs * np.float32(1e-3)
        lo, hi = o.min(axis=0), o.max(axis=0)
        q = np.minimum(((o - lo) / (hi - lo + 1e-20) * 1024).astype(np.uint32), 1023)
        m = morton3(q)
        octant = ((dirs[:, 0] < 0).astype(np.uint64) | ((dirs[:, 1] < 0).astype(np.uint64) << 1) | ((dirs[:, 2] < 0).astype(np.uint64) << 2))
        orders = {"A origin Morton (pixel-order proxy)": np.argsort(m, kind="stable"),
                  "B octant, origin Morton": np.argsort((octant << 30) | m, kind="stable"),
                  "C 4096 origin cells, octant, fine Morton": np.argsort(((m >> 18) << 33) | (octant << 30) | (m & ((1 << 18) - 1)), kind="stable"),
                  "D shuffled": rng.permutation(n)}
        for pipe, pname in ((capi.PIPE_WAVEFRONT, "wavefront"), (capi.PIPE_POOL, "pool")):
            ptb.set_pipeline(pipe)
            for label, idx in orders.items():
                rays = np.zeros((n, 8), dtype=np.float32); rays[:, 0:3], rays[:, 4:7] = o[idx], dirs[idx]
                h_rays, h_out = ptb.upload(rays), ptb.alloc(n * 8)
                for _ in range(int(os.environ.get('PROBE_WARM', 3))):
                    ptb.lib.ptb_closest_hit(h_rays, n, h_out)
                ptb.finish(); t0 = time.perf_counter()
                reps = int(os.environ.get('PROBE_REPS', 10))
                for _ in range(reps):
                    ptb.lib.ptb_closest_hit(h_rays, n, h_out)
                ptb.finish(); dt = (time.perf_counter() - t0) / reps
                hits = (ptb.read(h_out, np.float32, 2 * n).reshape(n, 2)[:, 0] > 0).mean()
                print(f"{name:12s} {pname:9s} {label:42s} {dt * 1e3:7.3f} ms  {n / dt / 1e6:7.0f} Mrays/s  hit {hits:.3f}", flush=True)
                ptb.release(h_rays); ptb.release(h_out)
    ptb.shutdown()


if __name__ == "__main__":
    main()
