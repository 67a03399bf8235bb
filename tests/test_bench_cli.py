"""bench.py's contract on the CPU side: the reference arm (the reference's own prog.cl on the host cores) prints ONE
JSON line with the keys the driver reads, on a size that takes a second; the CUDA arm refuses to run without a device
instead of falling back to anything."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, cwd=ROOT, timeout=600)


def test_reference_arm_prints_one_json_line():
    r = _run("--impl", "reference", "--workload", "cornell_512_16spp_d4", "--width", "64", "--height", "48",
             "--steps", "2", "--warmup", "1", "--ref-spp", "2")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True
    assert d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1 and d["gpu_launches"] == 0
    assert d["value"] > 0 and d["ms_per_step"] > 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["config"]["workload"] == "cornell_512_16spp_d4" and d["config"]["triangles"] == 32 and d["config"]["spp"] == 16
    assert "sample" in d["cpu_baseline"] and "fair" in d["cpu_baseline"]


def test_reference_arm_on_a_generated_scene_uses_the_port():
    r = _run("--impl", "reference", "--workload", "spheres100k_1080p_256spp_d4", "--width", "48", "--height", "27",
             "--steps", "1", "--warmup", "0", "--ref-spp", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["cpu_baseline"]["kind"] == "port" and d["value"] > 0


def test_cuda_arm_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = _run("--steps", "1", "--warmup", "0", "--width", "32", "--height", "32", "--spp", "1")
    assert r.returncode != 0
    assert "no CUDA device" in (r.stderr + r.stdout)
