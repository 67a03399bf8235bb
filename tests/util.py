"""Shared helpers of the test-suite: running the oracles, comparing images."""
import json
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_CPU = os.path.join(ROOT, "oracle", "_ref", "ref_cpu")
GOLDEN = os.path.join(ROOT, "tests", "golden")
CORNELL = os.path.join(ROOT, "scenes", "cornell.obj")


def have_ref_cpu():
    return os.access(REF_CPU, os.X_OK)


def run_ref_cpu(w, h, spp, iterations, obj=None, tribin=None, pos=(555, 0, 0), eye=None, fov=None, yaw=None, pitch=None,
                brute=False, count=False, threads=None, want=("colors", "rnds", "dump"), extra=()):
    """Run the reference's own host+kernels on the CPU (oracle/_ref/ref_cpu).  Returns a dict with the JSON
    line it prints plus the requested raw buffers."""
    tmp = tempfile.mkdtemp(prefix="ptb_ref_")
    cmd = [REF_CPU, "--w", str(w), "--h", str(h), "--spp", str(spp), "--iterations", str(iterations)]
    if obj:
        cmd += ["--obj", obj, "--pos"] + [str(v) for v in pos]
    if tribin:
        cmd += ["--tribin", tribin]
    if eye is not None:
        cmd += ["--eye"] + [repr(float(v)) for v in eye]
    if fov is not None:
        cmd += ["--fov", repr(float(fov))]
    if yaw is not None:
        cmd += ["--cam-yaw", repr(float(yaw))]
    if pitch is not None:
        cmd += ["--cam-pitch", repr(float(pitch))]
    if brute:
        cmd.append("--brute")
    if count:
        cmd.append("--count")
    if threads:
        cmd += ["--threads", str(threads)]
    files = {}
    if "colors" in want:
        files["colors"] = os.path.join(tmp, "colors.bin"); cmd += ["--out", files["colors"]]
    if "rnds" in want:
        files["rnds"] = os.path.join(tmp, "rnds.bin"); cmd += ["--rnds-out", files["rnds"]]
    if "dump" in want:
        files["dump"] = os.path.join(tmp, "surface.ptbd"); cmd += ["--dump", files["dump"]]
    if "tex" in want:
        files["tex"] = os.path.join(tmp, "tex.bin"); cmd += ["--tex-out", files["tex"]]
    cmd += list(extra)
    out = subprocess.run(cmd, capture_output=True, text=True, check=True).stdout
    info = json.loads([ln for ln in out.splitlines() if ln.startswith("{")][-1])
    res = {"info": info, "files": files}
    if "colors" in files:
        res["colors"] = np.fromfile(files["colors"], dtype=np.float32).reshape(h, w, 4)
    if "rnds" in files:
        res["rnds"] = np.fromfile(files["rnds"], dtype=np.int32)
    if "tex" in files:
        res["tex"] = np.fromfile(files["tex"], dtype=np.float32).reshape(h, w, 4)
    return res


def compare(a, b):
    """a, b: (H,W,>=3) float arrays.  Returns statistics of the per-pixel relative error of the rgb lanes."""
    a = np.asarray(a, dtype=np.float64)[..., :3]
    b = np.asarray(b, dtype=np.float64)[..., :3]
    num = np.abs(a - b).max(axis=-1)
    den = np.maximum(np.abs(b).max(axis=-1), 1e-6)
    rel = num / den
    npx = rel.size
    return {
        "bit_equal_frac": float((np.asarray(a, np.float32).view(np.uint32) == np.asarray(b, np.float32).view(np.uint32)).all(axis=-1).mean()),
        "frac_rel_le_1e-4": float((rel <= 1e-4).sum() / npx),
        "frac_rel_le_1e-5": float((rel <= 1e-5).sum() / npx),
        "max_rel": float(rel.max()),
        "rmse_over_mean": float(np.sqrt(((a - b) ** 2).mean()) / max(b.mean(), 1e-12)),
    }
