#!/usr/bin/env python
"""Regenerates tests/golden/* by running the REFERENCE's own prog.cl + main.cpp (oracle/_ref/ref_cpu, built
from /root/reference by oracle/Makefile) — only possible where /root/reference is mounted.  The fixtures pin
the oracle port (oracle/pt_oracle.c) and, through it, the CUDA path on boxes that have no reference tree.

  python tests/golden/make_golden.py
"""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.path.join(ROOT, "oracle", "_ref", "ref_cpu")


def run(name, w, h, spp, depth, scene_args, extra=(), tex=False):
    base = os.path.join(HERE, name)
    cmd = [REF, "--w", str(w), "--h", str(h), "--spp", str(spp), "--iterations", str(depth), "--threads", "4", "--count",
           "--out", base + ".colors.f32", "--rnds-out", base + ".rnds.i32"] + scene_args + list(extra)
    if tex:
        cmd += ["--tex-out", base + ".tex.f32"]
    out = subprocess.run(cmd, capture_output=True, text=True, check=True).stdout
    info = json.loads([ln for ln in out.splitlines() if ln.startswith("{")][-1])
    info.pop("seconds", None); info.pop("threads", None)
    json.dump(info, open(base + ".json", "w"), indent=1, sort_keys=True)
    print(name, info)


def main():
    from opencl_path_tracer_b200 import scenes
    cornell = ["--obj", os.path.join(ROOT, "scenes", "cornell.obj"), "--pos", "555", "0", "0"]
    # the kernel-argument surfaces as the reference host uploads them
    subprocess.run([REF, "--w", "96", "--h", "64", "--spp", "0", "--dump", os.path.join(HERE, "cornell_96x64.ptbd")] + cornell, check=True, capture_output=True)
    run("cornell_96x64_spp3_d4", 96, 64, 3, 4, cornell, tex=True)
    run("cornell_96x64_spp2_d1", 96, 64, 2, 1, cornell)
    d = scenes.describe("small_mixed")
    ptbs = os.path.join(HERE, "small_mixed.ptbs")
    scenes.write_ptbs(d, ptbs)
    mixed = ["--tribin", ptbs, "--eye"] + [repr(float(v)) for v in d["eye"]] + ["--fov", repr(float(d["fov"]))]
    subprocess.run([REF, "--w", "80", "--h", "60", "--spp", "0", "--iterations", "6", "--dump", os.path.join(HERE, "small_mixed_80x60.ptbd")] + mixed, check=True, capture_output=True)
    run("small_mixed_80x60_spp4_d6", 80, 60, 4, 6, mixed)
    os.remove(ptbs)
    filt_im_fixture(cornell)


def tie_image(w, h, seed=11):
    """random radiance with what makes the reference's unstable selection sort matter (prog.cl:410-424): runs of pixels with
    EQUAL grey value and different colours, flat black areas, repeated pixels"""
    import numpy as np
    rng = np.random.default_rng(seed)
    img = np.zeros((h, w, 4), dtype=np.float32)
    img[..., :3] = rng.random((h, w, 3), dtype=np.float32) * 2.0
    ties = np.array([(1, 2, 3), (3, 2, 1), (2, 2, 2), (0, 3, 3), (6, 0, 0), (2, 1, 3)], dtype=np.float32)
    img[4:14, 5:25, :3] = ties[rng.integers(0, len(ties), size=(10, 20))]
    img[16:24, 10:30, :3] = 0.0
    img[20:30, 28:44, :3] = ties[rng.integers(0, 3, size=(10, 16))]
    img[26:, :8, :3] = img[26, 0, :3]
    return img


def filt_im_fixture(cornell):
    """filt_im_48x32.{in,out}.f32: the reference's own filt_im kernel (prog.cl:391-427, launched as main.cpp:665-668 would)
    on an image with constructed ties; pixels the kernel does not write stay 0"""
    base = os.path.join(HERE, "filt_im_48x32")
    tie_image(48, 32).tofile(base + ".in.f32")
    subprocess.run([REF, "--w", "48", "--h", "32", "--spp", "0", "--colors-in", base + ".in.f32", "--filt-out", base + ".out.f32"] + cornell,
                   check=True, capture_output=True)
    print("filt_im_48x32 written")


if __name__ == "__main__":
    main()
