#!/usr/bin/env python
"""bench.py — Mrays/s and spp/s of the path-trace hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this implementation (CUDA, sm_100a)
  python bench.py --impl reference [...]                         # the reference's prog.cl on the host CPU cores

One "step" = one pass of the hot path over the whole frame: `spp` samples of every pixel of the
workload (default: BASELINE configs[1], Cornell box 1920x1080, 1024 spp, depth 4).  A "ray" is one
executed iteration of the bounce loop (prog.cl:317-319), counted by the kernel itself.

N > 1 (torchrun, one rank per GPU): spp-split — every rank renders `spp` samples of every pixel with
its own seed set into a sum buffer, then ONE NCCL reduce(SUM) to rank 0 + a sum/(N*spp) resolve, all
inside the timed step (weak scaling: per-GPU work fixed, the job is an N*spp-sample image).
`--split tile` (BASELINE configs[4]): the SAME frame x spp job is cut into bands of 8 scanlines dealt
round-robin to the ranks, every pixel keeps the reference's seed, and the disjoint pixel sets are summed
at rank 0 by one NCCL reduce per step (strong scaling: total work fixed; the image is bitwise the 1-GPU one).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

EMIT = print          # replaced by _QuietStdout.emit when run as a script
# measured once under ncu (never inside a timed run): DRAM bytes of one launch of the dominant kernel, per workload
NCU_DRAM_BYTES_PER_LAUNCH = {"cornell_1080p_1024spp_d4": 32.8e6,      # 8.35 MB read + 24.46 MB written
                             "spheres100k_1080p_256spp_d4": 87.1e6}   # k_wf_trace, one iteration: 83.3 MB read + 3.8 MB written
NCU_DRAM_SOURCE = {
    "cornell_1080p_1024spp_d4": "dram__bytes_read.sum + dram__bytes_write.sum of one k_render launch, ncu --set full (profiles/r01_render_v7.txt); "
                                "per launch, independent of spp: the per-pixel state is touched once",
    "spheres100k_1080p_256spp_d4": "dram__bytes_read.sum + dram__bytes_write.sum of one k_wf_trace launch (one bounce of every path: 2.07 M rays), "
                                   "ncu --set full (profiles/r01_wf_trace_v9.txt): the 64 B ray records and the queue; the 15 MB tree stays in L2",
}

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FP32_LANES_PER_SM = 128
WORKLOADS = {
    # name: (scene builder name, default W, H, spp, depth)
    "cornell_1080p_1024spp_d4": ("cornell", 1920, 1080, 1024, 4),
    "cornell_512_16spp_d4": ("cornell", 512, 512, 16, 4),
    "spheres100k_1080p_256spp_d4": ("spheres100k", 1920, 1080, 256, 4),
    "grid1m_4k_64spp_d4": ("grid1m", 3840, 2160, 64, 4),
    "grid1m_4k_8192spp_d4": ("grid1m", 3840, 2160, 8192, 4),     # BASELINE configs[4]; use with --split tile (and --spp for a prefix)
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cornell_1080p_1024spp_d4", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel per step")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--math", default="fast", choices=["fast", "parity"])
    ap.add_argument("--accel", default="auto", choices=["auto", "brute", "refbvh", "lbvh"])
    ap.add_argument("--pipeline", default="auto", choices=["auto", "megakernel", "wavefront"])
    ap.add_argument("--split", default="spp", choices=["spp", "tile"],
                    help="N > 1: spp (weak scaling: every rank renders spp samples of every pixel, reduce(SUM)) or tile (strong scaling: "
                         "the frame's bands of 8 scanlines round-robin over the ranks with the reference's seeds, bitwise the 1-GPU image)")
    ap.add_argument("--gather", default="nccl", choices=["nccl", "p2p"],
                    help="--split tile: nccl = one reduce of the disjoint pixel sets per step; p2p = every rank's kernels store their pixels "
                         "straight into rank 0's colors buffer over NVLink (CUDA IPC mapping), a barrier and no collective per step")
    ap.add_argument("--emulate-shard", type=int, default=0, metavar="N",
                    help="N = 1 GPU only, profiling aid: render just rank 0's share of an N-way tile split (what one GPU of --gpus N --split tile does)")
    ap.add_argument("--ref-spp", type=int, default=0, help="reference arm: samples per step (default: sized to a few seconds)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------- scenes (no reference files at run time)
def build_scene(name, scene):
    """Fill a host.Scene through the reference's own Scene interface (add_Obj / add_Triangle / end_Obj)."""
    from opencl_path_tracer_b200 import scenes
    if name == "cornell":
        scene.add_Obj(os.path.join(ROOT, "scenes", "cornell.obj"), (555, 0, 0), (1, 1, 1), 0, 0)
        scene.global_fov, scene.global_yaw, scene.global_pitch = 39.3, 0.0, 0.0
        eye = np.array([277.5, 277.5, -800.0], dtype=np.float32)
    else:
        eye = scenes.populate(scene, name)
    scene.global_shift = np.array([eye[0] - np.float32(500.0), eye[1] - np.float32(500.0), eye[2] + np.float32(1299.037842)], dtype=np.float32)


# ---------------------------------------------------------------- clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 8:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); power.append(float(r[2]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if r[4 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------- the reference arm / CPU baseline
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_cpu_reference(scene_name, W, H, spp, depth, threads):
    """Time the reference's own prog.cl on the host cores: oracle/_ref/ref_cpu when the reference compiled
    here ("reference"), else the committed C restatement oracle/libpt_oracle.so ("port")."""
    ref_cpu = os.path.join(ROOT, "oracle", "_ref", "ref_cpu")
    if scene_name == "cornell" and os.access(ref_cpu, os.X_OK):
        cmd = [ref_cpu, "--w", str(W), "--h", str(H), "--spp", str(spp), "--iterations", str(depth), "--threads", str(threads),
               "--obj", os.path.join(ROOT, "scenes", "cornell.obj"), "--pos", "555", "0", "0"]
        out = subprocess.run(cmd, capture_output=True, text=True, check=True).stdout
        info = json.loads([ln for ln in out.splitlines() if ln.startswith("{")][-1])
        return {"kind": "reference", "seconds": info["seconds"], "rays": info["rays"], "threads": threads}
    from oracle import pt_oracle_py   # the C port behind ctypes (bench.py's cpu leg is allowed to execute oracle/)
    return pt_oracle_py.time_render(scene_name, W, H, spp, depth, threads)


def reference_arm(args, W, H, depth, scene_name, workload):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    spp = args.ref_spp
    if spp <= 0:   # size a step to ~3 s of wall clock
        probe = run_cpu_reference(scene_name, W, H, 1, depth, cores)
        spp = max(1, min(64, int(3.0 / max(probe["seconds"], 1e-3))))
    rays = secs = 0.0
    kind = "reference"
    for i in range(args.warmup + args.steps):
        r = run_cpu_reference(scene_name, W, H, spp, depth, cores)
        kind = r["kind"]
        if i >= args.warmup:
            rays += r["rays"]; secs += r["seconds"]
    v = rays / secs / 1e6
    sample = f"{W}x{H} x {spp} spp per step (of the workload's spp), depth {depth}, {cores} threads"
    line = {"impl": "reference", "metric": "path-trace throughput", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "spp_per_s": spp * args.steps / secs,
            "config": {"workload": workload, "width": W, "height": H, "depth": depth, "sample": sample},
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    EMIT(json.dumps(line))


# ---------------------------------------------------------------- our arm
def main():
    args = parse()
    scene_name, W, H, spp, depth = WORKLOADS[args.workload]
    W, H, spp = args.width or W, args.height or H, args.spp or spp
    if args.impl == "reference":
        return reference_arm(args, W, H, depth, scene_name, args.workload)

    import torch
    import torch.distributed as dist
    from opencl_path_tracer_b200 import capi, host, multigpu

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this implementation has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    ptb = capi.Ptb(local)
    stream = torch.cuda.Stream(device=dev)      # a real (non-default) stream: the kernels and the timing events share it
    torch.cuda.set_stream(stream)
    ptb.set_stream(stream.cuda_stream)
    ptb.set_math(capi.MATH_FAST if args.math == "fast" else capi.MATH_PARITY)
    ptb.set_accel({"auto": 0, "brute": 1, "refbvh": 2, "lbvh": 3}[args.accel])
    ptb.set_pipeline({"auto": capi.PIPE_AUTO, "megakernel": capi.PIPE_MEGAKERNEL, "wavefront": capi.PIPE_WAVEFRONT}[args.pipeline])

    scene = host.Scene(ptb)
    scene.build_trees = args.accel == "refbvh"
    t0 = time.time()
    build_scene(scene_name, scene)
    tile = world > 1 and args.split == "tile"
    seed_skip = 0 if tile else multigpu.rank_seed_skip(rank, W, H)      # tile-split: every pixel keeps the reference's own seed
    scene.init_Scene(W, H, seed_skip=seed_skip)
    frame = multigpu.DeviceFrame(ptb, W, H, dev)
    seeds_host = torch.from_numpy(host.minstd_rand0_seeds(W * H, seed_skip)).pin_memory()
    frame.rnds.copy_(seeds_host)
    frame.attach(scene)
    p2p = tile and args.gather == "p2p"
    if p2p:                                      # every rank's colors argument is rank 0's buffer
        frame.h_colors = scene.buffer_colors = multigpu.share_root_colors(ptb, W, H)
    scene.upload_Triangles()
    scene.upload_Materials()
    scene.iterations = depth
    scene.camera = scene._camera()
    scene._set_trace_args()
    ptb.build_accel()
    ptb.finish()
    setup_s = time.time() - t0
    n_tris = len(scene.tris)
    accel_used = {1: "brute(shared memory)", 2: "reference trees", 3: "lbvh"}[ptb.get_accel()]
    ptb.set_accumulate_sum(1 if world > 1 and not tile else 0)
    if tile:
        ptb.set_pixel_shard(rank, world, multigpu.shard_group(W))
    elif world == 1 and args.emulate_shard > 1:
        ptb.set_pixel_shard(0, args.emulate_shard, multigpu.shard_group(W))

    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def step():
        scene.current_sample = 0
        if p2p:
            scene.render_spp(spp)
            multigpu.finish_peer_gather(ptb)     # stream drained + barrier: rank 0 holds the frame
            return
        if world > 1:
            frame.colors.zero_()
        scene.render_spp(spp)
        if world > 1:
            multigpu.reduce_sum_to_root(frame.colors)     # tile-split: disjoint pixel sets, x + 0 == x, so the sum is an exact gather
            if rank == 0 and not tile:
                ptb.resolve_sum(frame.h_colors, spp * world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    ptb.reset_counters()
    launches0 = ptb.launch_count()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for a, b in ev:
        flush.fill_(1)             # L2 flush between timed iterations (not timed)
        a.record(stream)
        step()
        b.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = ptb.launch_count() - launches0
    ms = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device=dev)
    cnt = ptb.counters()
    rays = torch.tensor([float(cnt["rays"])], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    total_ms, total_rays = float(ms.item()), float(rays.item())
    value = total_rays / (total_ms * 1e-3) / 1e6

    # ---- end to end through the C-ABI with HOST buffers: seeds + scene up, colors down, every step
    e2e = None
    if not args.no_e2e:
        colors_host = torch.empty(W * H, 4, dtype=torch.float32).pin_memory()
        tris_b, mats_b = np.ascontiguousarray(scene.tris), np.stack(scene.mats)
        h2d = seeds_host.numel() * 4 + tris_b.nbytes + mats_b.nbytes
        d2h = colors_host.numel() * 4 if rank == 0 else 0

        def e2e_step():
            ptb.write_ptr(frame.h_rnds, seeds_host.data_ptr(), seeds_host.numel() * 4)
            ptb.write(scene.buffer_tris, tris_b)
            ptb.write(scene.buffer_mats, mats_b)
            step()
            if rank == 0:
                ptb.read_ptr(frame.h_colors, colors_host.data_ptr(), colors_host.numel() * 4)

        e2e_step()
        barrier()
        ptb.reset_counters()
        ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_e2e = max(1, min(args.steps, 3))
        ea.record(stream)
        for _ in range(n_e2e):
            e2e_step()
        eb.record(stream)
        barrier()
        ems = torch.tensor([ea.elapsed_time(eb)], dtype=torch.float64, device=dev)
        erays = torch.tensor([float(ptb.counters()["rays"])], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ems, op=dist.ReduceOp.MAX)
            dist.all_reduce(erays, op=dist.ReduceOp.SUM)
        e2e = {"value": float(erays.item()) / (float(ems.item()) * 1e-3) / 1e6, "unit": "Mrays/s",
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": n_e2e,
               "ms_per_step": float(ems.item()) / n_e2e,
               "what": "ptb_write(seeds, triangles, materials from pinned/pageable host) + ptb_render(spp) + ptb_read(colors to pinned host) per step"}

    # ---- the bit-exact flavour beside the shipping one (rank 0, N == 1): same kernels' structure, prog.cl's literal arithmetic
    exact = None
    if rank == 0 and world == 1 and args.math == "fast":
        n = min(spp, 64)
        ptb.set_math(capi.MATH_PARITY)
        ptb.set_accumulate_sum(0)
        scene.current_sample = 0
        scene.render_spp(n)                                   # warm (stages the parity layouts)
        torch.cuda.synchronize()
        ptb.reset_counters()
        ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        scene.current_sample = 0
        ea.record(stream)
        scene.render_spp(n)
        eb.record(stream)
        torch.cuda.synchronize()
        exact = {"math": "parity", "value": float(ptb.counters()["rays"]) / (ea.elapsed_time(eb) * 1e-3) / 1e6, "unit": "Mrays/s", "spp": n,
                 "note": "PTB_MATH_PARITY: bit-identical to the reference's prog.cl (tests/), timed on the same scene; the headline runs "
                         "PTB_MATH_FAST, which the tests hold to 1e-4 relative on >= 99.9 % of pixels"}
        ptb.set_math(capi.MATH_FAST)

    # ---- algorithmic flops per ray of the shipping traversal, from a counting pass (untimed)
    roof = None
    if rank == 0:
        ptb.set_accumulate_sum(0)
        ptb.set_count_work(1)
        ptb.reset_counters()
        scene.current_sample = 0
        scene.render_spp(min(spp, 4))
        c = ptb.counters()
        ptb.set_count_work(0)
        r = max(c["rays"], 1)
        B, T, T1, T2, T3 = (c[k] / r for k in ("box_tests", "tri_tests", "tri_stage1", "tri_stage2", "tri_stage3"))
        f_ray = 22.0 * B + 14.0 * T + 26.0 * T1 + 20.0 * T2 + 20.0 * T3 + 150.0
        sms = ptb.compute_units()
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        sm_max = float(peaks.get("sm_max_mhz", 1965.0))
        peak_tflops = sms * FP32_LANES_PER_SM * 2 * sm_max * 1e6 / 1e12
        per_gpu_rays_s = total_rays / world / (total_ms * 1e-3)
        achieved = f_ray * per_gpu_rays_s / 1e12
        wavefront = accel_used == "lbvh" and args.pipeline != "megakernel"
        kernel_name = ("k_wf_trace (wavefront closest-hit over the 4-ary LBVH; k_wf_shade is the other ~20 % of a step)" if wavefront
                       else "k_render (fused gen_ray + trace_ray + accumulate)")
        roof = {"bound": "fp32", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
                "traffic": NCU_DRAM_BYTES_PER_LAUNCH.get(args.workload), "traffic_source": NCU_DRAM_SOURCE.get(args.workload),
                "kernel": kernel_name,
                "peak_source": f"{sms} SMs x 128 FP32 lanes x 2 x {sm_max:.0f} MHz (sm_max_mhz of MEASURED_PEAKS.json); the path has no dense contraction "
                               "and ~0 mandatory HBM traffic, so neither the hbm nor the tensor roofline bounds it (SURVEY 8d)",
                "numerator": "algorithmic flops of the reference's closest-hit search for this scene (counting pass over the plain triangles); "
                             "the shipping scan may execute fewer (merged quad primitives, config.scan_primitives)",
                "flops_per_ray": f_ray, "box_tests_per_ray": B, "tri_tests_per_ray": T,
                "tri_stage_fractions": [T1 / max(T, 1e-9), T2 / max(T, 1e-9), T3 / max(T, 1e-9)],
                "hbm_secondary": {"algorithmic_bytes_per_step": W * H * (4 + 4 + 16 + 16), "note": "seed read + state/colors/image write per pixel per launch"}}

    # ---- CPU baseline beside it (rank 0, N == 1 only): a bounded sample of the same workload
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cores = host_cores()
            probe = run_cpu_reference(scene_name, W, H, 1, depth, cores)
            n = max(1, min(spp, int(15.0 / max(probe["seconds"], 1e-3))))
            r = run_cpu_reference(scene_name, W, H, n, depth, cores)
            cpu = {"value": r["rays"] / r["seconds"] / 1e6, "unit": "Mrays/s", "cores": cores, "kind": r["kind"],
                   "sample": f"{W}x{H} x {n} spp of the workload's {spp}, depth {depth} ({r['seconds']:.1f} s)"}
        except Exception as e:   # noqa: BLE001 — the baseline is a report, never a reason to lose the GPU number
            cpu = {"value": None, "unit": "Mrays/s", "cores": host_cores(), "kind": "unavailable", "sample": f"{type(e).__name__}: {e}"}

    if rank == 0:
        line = {
            "metric": "path-trace throughput", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong" if tile else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "spp_per_s": spp * (1 if tile else world) * args.steps / (total_ms * 1e-3),
            "rays_per_step": total_rays / args.steps,
            "config": {"workload": args.workload, "width": W, "height": H, "spp_per_gpu_per_step": spp, "split": (args.split if world > 1 else None), "emulate_shard": args.emulate_shard or None, "depth": depth, "triangles": n_tris,
                       "scan_primitives": int(ptb.scan_prims()) if accel_used.startswith("brute") and args.math == "fast" else None,
                       "accel": accel_used, "math": args.math, "pipeline": args.pipeline,
                       "parallelism": "single GPU" if world == 1 else
                       ((f"tile-split x{world} (bands of 8 scanlines round-robin, reference seeds), gather fused into the path kernel: every rank stores "
                         "its pixels into rank 0's colors buffer over NVLink (CUDA IPC), barrier only" if p2p else
                         f"tile-split x{world} (bands of 8 scanlines round-robin, reference seeds) + NCCL reduce of the disjoint pixel sets per step") if tile
                        else f"spp-split x{world} + NCCL reduce(SUM) of the accumulation buffer per step"),
                       "l2": "512 MB buffer written between timed steps (L2 flush)", "setup_s": setup_s,
                       "lbvh": ptb.accel_stats() if accel_used == "lbvh" else None},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu, "bit_exact_flavour": exact,
        }
        EMIT(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


class _QuietStdout:
    """The contract is ONE JSON line on stdout: while the bench runs, file descriptor 1 points at stderr so that library
    chatter (NCCL's version banner, torchrun notices) cannot land in front of it; `emit` writes to the real stdout."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def emit(self, text):
        sys.stdout.flush()
        os.write(self.saved, (text + "\n").encode())

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)
        return False


if __name__ == "__main__":
    with _QuietStdout() as _out:
        EMIT = _out.emit
        main()
