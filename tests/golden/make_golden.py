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


if __name__ == "__main__":
    main()
