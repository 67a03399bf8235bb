#!/usr/bin/env python
"""tests/golden/grid1m_4k_rows_spp4_d4.* — the 1 M-triangle scene of BASELINE configs[3]/[4] at its own camera and
resolution (3840x2160, depth 4), rendered by the REFERENCE's own prog.cl (oracle/_ref/ref_cpu, built from
/root/reference by oracle/Makefile) in its brute-force mode (`first_intersect` over all triangles, prog.cl:318 — the
reference's host BVH builder does not survive 1 M triangles) on a pixel subset: the six scanlines
y = 0, 360, ... 1800 (23 040 pixels, pixel-shard rank 0 of 360 with groups of one scanline), 4 samples each, with the
reference's per-pixel seeds (main.cpp:522-526).  About 3e11 ray-triangle tests: minutes on 8 cores, which is why the
result is committed as a fixture (colors + RNG states of those rows) instead of being recomputed by every test run.

  python tests/golden/make_golden_grid1m.py        # only where /root/reference is mounted (ref_cpu built)
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.path.join(ROOT, "oracle", "_ref", "ref_cpu")

W, H, SPP, DEPTH = 3840, 2160, 4, 4
SHARD_N, SHARD_GROUP = 360, W                      # rank 0: scanlines 0, 360, 720, 1080, 1440, 1800
NAME = "grid1m_4k_rows_spp4_d4"


def main():
    from opencl_path_tracer_b200 import scenes
    d = scenes.describe("grid1m")
    tmp = tempfile.mkdtemp(prefix="ptb_grid1m_")
    ptbs = os.path.join(tmp, "grid1m.ptbs")
    scenes.write_ptbs(d, ptbs)
    colors, rnds = os.path.join(tmp, "colors.bin"), os.path.join(tmp, "rnds.bin")
    cmd = [REF, "--w", str(W), "--h", str(H), "--spp", str(SPP), "--iterations", str(DEPTH), "--tribin", ptbs, "--no-bvh", "--brute",
           "--eye"] + [repr(float(v)) for v in d["eye"]] + ["--fov", repr(float(d["fov"])), "--cam-yaw", repr(float(d["yaw"])),
           "--cam-pitch", repr(float(d["pitch"])), "--shard", "0", str(SHARD_N), str(SHARD_GROUP), "--out", colors, "--rnds-out", rnds]
    if len(sys.argv) > 1:
        cmd += ["--threads", sys.argv[1]]
    out = subprocess.run(cmd, capture_output=True, text=True, check=True).stdout
    info = json.loads([ln for ln in out.splitlines() if ln.startswith("{")][-1])
    rows = np.arange(0, H, SHARD_N)
    c = np.fromfile(colors, dtype=np.float32).reshape(H, W, 4)[rows]
    r = np.fromfile(rnds, dtype=np.int32).reshape(H, W)[rows]
    base = os.path.join(HERE, NAME)
    np.ascontiguousarray(c).tofile(base + ".colors.f32")
    np.ascontiguousarray(r).tofile(base + ".rnds.i32")
    meta = {k: info[k] for k in ("w", "h", "spp", "iterations", "tris", "rays")}
    meta.update({"rows": rows.tolist(), "shard": [0, SHARD_N, SHARD_GROUP], "scene": "grid1m", "mode": "brute (prog.cl:318)",
                 "eye": list(map(float, d["eye"])), "fov": d["fov"], "yaw": d["yaw"], "pitch": d["pitch"]})
    json.dump(meta, open(base + ".json", "w"), indent=1, sort_keys=True)
    print(NAME, meta, "seconds:", info["seconds"])


if __name__ == "__main__":
    main()
