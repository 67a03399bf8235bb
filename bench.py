#!/usr/bin/env python
"""bench.py — Mrays/s and spp/s of the path-trace hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this implementation (CUDA, sm_100a)
  python bench.py --impl reference [...]                         # the reference's prog.cl on the host CPU cores

One "step" = one pass of the hot path over the whole frame: `spp` samples of every pixel of the workload.  A "ray" is
one executed iteration of the bounce loop (prog.cl:317-319), counted by the kernel itself.

The printed line's headline is BASELINE configs[1] (Cornell 1920x1080, 1024 spp, depth 4).  Its `workloads` object carries
the configurations `north_star`'s target is stated on, measured the same way in the same run: configs[2] (104 k triangles,
1080p, 256 spp, GPU LBVH), configs[3] (1 M triangles, 4K, 64 spp, spp-split + NCCL reduce) and configs[4] (1 M triangles,
4K, tile split with the gather fused into the path kernels' own stores; a prefix of the 8192 spp per step).

N > 1 (torchrun, one rank per GPU): every workload is a FIXED total job cut N ways ("scaling": "strong"):
  spp-split   rank r renders its share of the job's spp (multigpu.split_spp) of every pixel from its own seed set into a sum
              buffer; ONE NCCL reduce(SUM) to rank 0 on the library's stream (ptb_comm_reduce) and the fused
              sum -> mean -> tone-map resolve (ptb_resolve_frame), all inside the timed step;
  tile-split  bands of 8 scanlines dealt round-robin, every pixel keeps the reference's seed (bitwise the 1-GPU image); the
              ranks' kernels store their pixels straight into rank 0's colors buffer over NVLink (CUDA IPC mapping), a
              barrier and no collective per step (`--gather nccl`: one reduce of the disjoint pixel sets instead).
Each N > 1 workload ends with an untimed `parity_check` of the merged frame against a 1-GPU render on rank 0.
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
ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# measured once under ncu (never inside a timed run), per launch of the workload's dominant kernel: DRAM bytes
# (dram__bytes_read.sum + dram__bytes_write.sum, --set full) and executed FP32 flops (fadd + fmul + 2 ffma thread instructions)
NCU = {
    "cornell_1080p_1024spp_d4": {"kernel": "k_render", "dram_bytes": 32.8e6, "source": "profiles/r01_render_v7.txt (one launch; independent of spp: the per-pixel state is touched once)"},
}
try:   # refreshed by the round's captures (profiles/ncu_constants.json, written from the .ncu-rep summaries)
    NCU.update(json.load(open(os.path.join(ROOT, "profiles", "ncu_constants.json"))))
except (OSError, ValueError):
    pass

FP32_LANES_PER_SM = 128
WORKLOADS = {
    # name: (scene, W, H, spp of the job, depth, default split for N > 1, spp actually rendered per step (0 = all))
    "cornell_1080p_1024spp_d4": ("cornell", 1920, 1080, 1024, 4, "spp", 0),
    "cornell_512_16spp_d4": ("cornell", 512, 512, 16, 4, "spp", 0),
    "spheres100k_1080p_256spp_d4": ("spheres100k", 1920, 1080, 256, 4, "spp", 0),
    "grid1m_4k_64spp_d4": ("grid1m", 3840, 2160, 64, 4, "spp", 0),
    "grid1m_4k_8192spp_d4": ("grid1m", 3840, 2160, 8192, 4, "tile", 128),     # BASELINE configs[4]: a 128-spp prefix per step
}
EXTRA_WORKLOADS = ["spheres100k_1080p_256spp_d4", "grid1m_4k_64spp_d4", "grid1m_4k_8192spp_d4"]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cornell_1080p_1024spp_d4", choices=sorted(WORKLOADS))
    ap.add_argument("--workloads", default="default", help="extra workloads for the `workloads` object: default (configs 2-4 when the headline is the "
                                                          "default one), none, or a comma-separated list")
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel per step (of the whole job)")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--math", default="fast", choices=["fast", "parity"])
    ap.add_argument("--accel", default="auto", choices=["auto", "brute", "refbvh", "lbvh"])
    ap.add_argument("--pipeline", default="auto", choices=["auto", "megakernel", "wavefront", "pool"])
    ap.add_argument("--split", default="auto", choices=["auto", "spp", "tile"], help="N > 1: how the fixed job is cut (auto: the workload's own)")
    ap.add_argument("--gather", default="p2p", choices=["nccl", "p2p"],
                    help="--split tile: p2p = every rank's kernels store their pixels straight into rank 0's colors buffer over NVLink (CUDA IPC "
                         "mapping), a barrier and no collective per step; nccl = one reduce of the disjoint pixel sets per step")
    ap.add_argument("--emulate-shard", type=int, default=0, metavar="N",
                    help="N = 1 GPU only, profiling aid: render just rank 0's share of an N-way tile split (what one GPU of --gpus N --split tile does)")
    ap.add_argument("--ref-spp", type=int, default=0, help="reference arm: samples per step (default: sized to a few seconds)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity-check", action="store_true")
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


def scene_triangles(name):
    if name == "cornell":
        return 32
    from opencl_path_tracer_b200 import scenes
    return int(len(scenes.describe(name)["tri_mat"]))


def workload_config(workload, W, H, spp, depth, n_tris):
    """the workload itself, identical in both arms' lines (everything implementation-specific lives in `detail`)"""
    return {"workload": workload, "width": W, "height": H, "spp": spp, "depth": depth, "triangles": n_tris, "seeds": "std::minstd_rand0, pixel i <- i-th output (main.cpp:522-526)"}


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


def run_cpu_reference(scene_name, W, H, spp, depth, threads, fair=False):
    """Time the reference's own prog.cl on the host cores: oracle/_ref/ref_cpu when the reference compiled here
    ("reference": canonical -O2 -ffp-contract=off build; fair=True: the -O3 x86-64-v3 FMA build of the same sources), else
    the committed C restatement oracle/libpt_oracle.so ("port")."""
    ref_cpu = os.path.join(ROOT, "oracle", "_ref", "ref_cpu_fair" if fair else "ref_cpu")
    if scene_name == "cornell" and os.access(ref_cpu, os.X_OK):
        cmd = [ref_cpu, "--w", str(W), "--h", str(H), "--spp", str(spp), "--iterations", str(depth), "--threads", str(threads),
               "--obj", os.path.join(ROOT, "scenes", "cornell.obj"), "--pos", "555", "0", "0"]
        out = subprocess.run(cmd, capture_output=True, text=True, check=True).stdout
        info = json.loads([ln for ln in out.splitlines() if ln.startswith("{")][-1])
        return {"kind": "reference", "seconds": info["seconds"], "rays": info["rays"], "threads": threads}
    if fair:
        raise RuntimeError("no fair build for this scene")
    from oracle import pt_oracle_py   # the C port behind ctypes (bench.py's cpu leg is allowed to execute oracle/)
    return pt_oracle_py.time_render(scene_name, W, H, spp, depth, threads)


def reference_arm(args, W, H, spp_job, depth, scene_name, workload):
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
    sample = f"{W}x{H} x {spp} spp per step (of the workload's {spp_job}), depth {depth}, {cores} threads"
    fair = None
    try:
        f = run_cpu_reference(scene_name, W, H, spp, depth, cores, fair=True)
        fair = {"value": f["rays"] / f["seconds"] / 1e6, "unit": "Mrays/s", "build": "g++ -O3 -march=x86-64-v3 (AVX2 + FMA contraction) of the same reference sources"}
    except Exception as e:   # noqa: BLE001
        fair = {"value": None, "unavailable": f"{type(e).__name__}: {e}"[:200]}
    line = {"impl": "reference", "metric": "path-trace throughput", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "spp_per_s": spp * args.steps / secs,
            "config": workload_config(workload, W, H, spp_job, depth, scene_triangles(scene_name)),
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample,
                             "build": "g++ -O2 -ffp-contract=off (the canonical rounding the parity tests compare against)", "fair": fair},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    EMIT(json.dumps(line))


# ---------------------------------------------------------------- our arm: one workload
class Ctx:
    pass


def run_workload(cx, args, workload, headline):
    """Set the workload up on every rank, time K steps (device events, max over ranks), the end-to-end arm, a counting pass
    for the roofline numerator and — N > 1 — the untimed parity check of the merged frame.  Returns the result dict (rank 0)."""
    import torch
    import torch.distributed as dist
    from opencl_path_tracer_b200 import capi, host, multigpu
    ptb, dev, stream, world, rank = cx.ptb, cx.dev, cx.stream, cx.world, cx.rank
    scene_name, W, H, spp_job, depth, split_default, spp_step = WORKLOADS[workload]
    if headline:
        W, H = args.width or W, args.height or H
        if args.spp:
            spp_job, spp_step = args.spp, 0
    spp_total = spp_step or spp_job                                     # samples of every pixel one step renders (all ranks together)
    split = split_default if (not headline or args.split == "auto") else args.split
    tile = world > 1 and split == "tile"
    spp_mine = spp_total if (world == 1 or tile) else multigpu.split_spp(spp_total, world, rank)
    p2p = tile and args.gather == "p2p"

    ptb.set_math(capi.MATH_FAST if args.math == "fast" else capi.MATH_PARITY)
    ptb.set_accel({"auto": 0, "brute": 1, "refbvh": 2, "lbvh": 3}[args.accel])
    ptb.set_pipeline({"auto": capi.PIPE_AUTO, "megakernel": capi.PIPE_MEGAKERNEL, "wavefront": capi.PIPE_WAVEFRONT, "pool": getattr(capi, "PIPE_POOL", capi.PIPE_AUTO)}[args.pipeline])
    ptb.set_accumulate_sum(0)
    ptb.set_pixel_shard(0, 1, 1)

    scene = host.Scene(ptb)
    scene.build_trees = args.accel == "refbvh"
    t0 = time.time()
    build_scene(scene_name, scene)
    seed_skip = 0 if (tile or world == 1) else multigpu.rank_seed_skip(rank, W, H)      # tile-split: every pixel keeps the reference's own seed
    frame = multigpu.DeviceFrame(ptb, W, H, dev)
    seeds_host = torch.from_numpy(host.minstd_rand0_seeds(W * H, seed_skip)).pin_memory()
    ptb.write_ptr(frame.h_rnds, seeds_host.data_ptr(), seeds_host.numel() * 4)
    scene.screen_width, scene.screen_height, scene.rays_size = W, H, W * H
    frame.attach(scene)
    shared = None
    if p2p:                                      # every rank's colors argument is rank 0's buffer
        shared = multigpu.share_root_colors(ptb, W, H)
        scene.buffer_colors = shared
    colors_h = scene.buffer_colors
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
    sum_mode = world > 1 and not tile

    def configure():
        ptb.set_accumulate_sum(1 if sum_mode else 0)
        if tile:
            ptb.set_pixel_shard(rank, world, multigpu.shard_group(W))
        elif world == 1 and headline and args.emulate_shard > 1:
            ptb.set_pixel_shard(0, args.emulate_shard, multigpu.shard_group(W))
        else:
            ptb.set_pixel_shard(0, 1, 1)

    def step(n_mine=None, n_total=None):
        """one pass of the hot path over the frame: this rank's share, then the exchange step, then the root's resolve"""
        n_mine = spp_mine if n_mine is None else n_mine
        n_total = spp_total if n_total is None else n_total
        scene.current_sample = 0
        if p2p:
            scene.render_spp(n_mine)
            ptb.comm_barrier()                   # every rank's stream drained: rank 0's buffer holds the frame; no pixel data moves here
            return
        if world > 1:
            frame.colors.zero_()
            ptb.invalidate(frame.h_colors)
        if n_mine > 0:
            scene.render_spp(n_mine)
        if world > 1:
            ptb.comm_reduce(frame.h_colors, 0)   # spp-split: the real sum; tile-split: disjoint pixel sets, x + 0 == x, an exact gather
            if rank == 0 and not tile:
                ptb.resolve_frame(frame.h_colors, W, H, divide_by=n_total, mode=0, image=frame.h_image, want_rgba8=False)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    configure()
    for _ in range(args.warmup):
        step()
    barrier()
    ptb.reset_counters()
    launches0 = ptb.launch_count()
    sampler = ClockSampler(cx.local)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for a, b in ev:
        cx.flush.fill_(1)          # L2 flush between timed iterations (not timed)
        a.record(stream)
        step()
        b.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = ptb.launch_count() - launches0
    ms = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device=dev)
    rays = torch.tensor([float(ptb.counters()["rays"])], dtype=torch.float64, device=dev)
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
        tris_pin = torch.from_numpy(tris_b.view(np.uint8).reshape(-1)).pin_memory()       # pinned like the seeds: the copy runs at link speed
        mats_pin = torch.from_numpy(np.ascontiguousarray(mats_b).view(np.uint8).reshape(-1)).pin_memory()
        h2d = seeds_host.numel() * 4 + tris_b.nbytes + mats_b.nbytes
        d2h = colors_host.numel() * 4 if rank == 0 else 0

        def e2e_step():
            ptb.write_ptr(frame.h_rnds, seeds_host.data_ptr(), seeds_host.numel() * 4)
            ptb.write_ptr(scene.buffer_tris, tris_pin.data_ptr(), tris_pin.numel())
            ptb.write_ptr(scene.buffer_mats, mats_pin.data_ptr(), mats_pin.numel())
            step()
            if rank == 0:
                ptb.read_ptr(colors_h, colors_host.data_ptr(), colors_host.numel() * 4)
            if p2p:
                ptb.comm_barrier()               # the peers' next step stores into the buffer rank 0 has just read: not before it is done

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
               "what": "ptb_write(seeds, triangles, materials from pinned host memory) + ptb_render(spp) [+ exchange + resolve] + ptb_read(colors to pinned host) per step"}

    # ---- N > 1: is the merged frame right?  (untimed)
    parity = None
    if world > 1 and not args.no_parity_check:
        chk = 4
        ptb.write_ptr(frame.h_rnds, seeds_host.data_ptr(), seeds_host.numel() * 4)
        if tile:
            step(chk, chk)
            barrier()
            merged = None
            if rank == 0:
                merged = torch.from_numpy(ptb.read(colors_h, np.float32, W * H * 4).reshape(W * H, 4))
            barrier()
            if rank == 0:                         # the same 4 samples of the same pixel sets on this GPU alone, reference seeds: shard after shard, so
                # that every shard takes the pipeline decision its rank took (AUTO goes by owned pixels; in the FAST flavour the pool and the
                # wavefront kernels agree only to rounding — different FMA contraction of the same source —, in the parity flavour bit for bit)
                ptb.write_ptr(frame.h_rnds, seeds_host.data_ptr(), seeds_host.numel() * 4)
                scene.buffer_colors = frame.h_colors
                for r in range(world):
                    ptb.set_pixel_shard(r, world, multigpu.shard_group(W))
                    scene.current_sample = 0
                    scene.render_spp(chk)
                ptb.finish()
                ptb.set_pixel_shard(rank, world, multigpu.shard_group(W))
                one = torch.from_numpy(ptb.read(frame.h_colors, np.float32, W * H * 4).reshape(W * H, 4))
                same = bool((merged.view(torch.int32) == one.view(torch.int32)).all().item())
                rel = float(((merged[:, :3] - one[:, :3]).abs().max(dim=1).values / one[:, :3].abs().max(dim=1).values.clamp_min(1e-6)).max().item())
                parity = {"kind": "tile-split frame gathered at rank 0 == the same pixel shards rendered one after the other on one GPU, bit for bit",
                          "spp": chk, "pixels": W * H, "bitwise_equal": same, "max_rel_diff": rel, "ok": same}
                scene.buffer_colors = colors_h
            barrier()
        else:
            n_each = chk
            step(n_each, n_each * world)
            barrier()
            if rank == 0:
                merged = torch.from_numpy(ptb.read(frame.h_colors, np.float32, W * H * 4).reshape(W * H, 4))[:, :3].double()
                ptb.set_accumulate_sum(0)
                ptb.write_ptr(frame.h_rnds, seeds_host.data_ptr(), seeds_host.numel() * 4)
                scene.current_sample = 0
                half = max(1, n_each * world // 2)
                scene.render_spp(half)
                ptb.finish()
                one = torch.from_numpy(ptb.read(frame.h_colors, np.float32, W * H * 4).reshape(W * H, 4))[:, :3].double()
                scene.render_spp(n_each * world - half)
                ptb.finish()
                two = torch.from_numpy(ptb.read(frame.h_colors, np.float32, W * H * 4).reshape(W * H, 4))[:, :3].double()
                mean2 = float(two.mean())
                rel_mean = abs(float(merged.mean()) / mean2 - 1.0)
                rmse = float(((merged - two) ** 2).mean().sqrt()) / mean2
                noise = float(((one - two) ** 2).mean().sqrt()) / mean2
                ok = rel_mean <= 0.02 and rmse <= 3.0 * noise + 1e-3
                parity = {"kind": "spp-split merged frame vs a 1-GPU render of the same total spp (different random numbers: Monte-Carlo bound "
                                  "rel. mean <= 2 %, RMSE <= 3 x the 1-GPU half-vs-full RMSE)", "spp_total": n_each * world,
                          "rel_mean": rel_mean, "rmse_over_mean": rmse, "noise_over_mean": noise, "ok": bool(ok)}
            barrier()
        configure()

    # ---- the bit-exact flavour beside the shipping one (rank 0, N == 1): same kernels' structure, prog.cl's literal arithmetic
    exact = None
    if rank == 0 and world == 1 and args.math == "fast":
        n = min(spp_total, 64)
        ptb.set_math(capi.MATH_PARITY)
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
        ptb.set_math(capi.MATH_FAST if args.math == "fast" else capi.MATH_PARITY)

    # ---- algorithmic flops per ray of the SHIPPING traversal: a counting build of the same kernels (untimed)
    roof = None
    if rank == 0:
        ptb.set_accumulate_sum(0)
        ptb.set_pixel_shard(0, 1, 1)
        ptb.set_count_work(1)
        ptb.reset_counters()
        scene.current_sample = 0
        scene.render_spp(min(spp_total, 4))
        c = ptb.counters()
        ptb.set_count_work(0)
        r = max(c["rays"], 1)
        B, T, T1, T2, T3 = (c[k] / r for k in ("box_tests", "tri_tests", "tri_stage1", "tri_stage2", "tri_stage3"))
        f_ray = 22.0 * B + 14.0 * T + 26.0 * T1 + 20.0 * T2 + 20.0 * T3 + 150.0
        sms = ptb.compute_units()
        sm_max = float(cx.peaks.get("sm_max_mhz", 1965.0))
        peak_tflops = sms * FP32_LANES_PER_SM * 2 * sm_max * 1e6 / 1e12
        per_gpu_rays_s = total_rays / world / (total_ms * 1e-3)
        achieved = f_ray * per_gpu_rays_s / 1e12
        ncu = NCU.get(workload, {})
        executed = None
        if ncu.get("fp32_flops_per_ray"):
            executed = {"flops_per_ray": ncu["fp32_flops_per_ray"], "tflops": ncu["fp32_flops_per_ray"] * per_gpu_rays_s / 1e12,
                        "frac": ncu["fp32_flops_per_ray"] * per_gpu_rays_s / 1e12 / peak_tflops,
                        "what": "EXECUTED FP32 flops per ray from ncu (fadd + fmul + 2 x ffma thread instructions, FFMA2 counted twice) x this run's rays/s"}
        roof = {"bound": "fp32", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
                "traffic": ncu.get("dram_bytes"), "traffic_source": ncu.get("source"), "kernel": ncu.get("kernel") or cx.kernel_name(accel_used),
                "peak_source": f"{sms} SMs x 128 FP32 lanes x 2 x {sm_max:.0f} MHz (sm_max_mhz of MEASURED_PEAKS.json); the path has no dense contraction "
                               "and ~0 mandatory HBM traffic, so neither the hbm nor the tensor roofline bounds it (SURVEY 8d)",
                "numerator": "SURVEY 8d: F_ray = 22 B + 14 T + 26 T1 + 20 T2 + 20 T3 + 150 with B, T, T1..T3 counted by a COUNT_WORK build of the "
                             "SHIPPING traversal (merged quads count as one primitive test each; LBVH: wide-node child boxes, leaf triangles, the outlier scan)",
                "flops_per_ray": f_ray, "box_tests_per_ray": B, "tri_tests_per_ray": T,
                "tri_stage_fractions": [T1 / max(T, 1e-9), T2 / max(T, 1e-9), T3 / max(T, 1e-9)],
                "executed_flops": executed,
                "hbm_secondary": {"algorithmic_bytes_per_step": W * H * (4 + 4 + 16 + 16), "note": "seed read + state/colors/image write per pixel per launch"}}
        configure()

    # ---- CPU baseline beside it (headline only, rank 0, N == 1): a bounded sample of the same workload
    cpu = None
    if headline and rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cores = host_cores()
            probe = run_cpu_reference(scene_name, W, H, 1, depth, cores)
            n = max(1, min(spp_total, int(15.0 / max(probe["seconds"], 1e-3))))
            r = run_cpu_reference(scene_name, W, H, n, depth, cores)
            cpu = {"value": r["rays"] / r["seconds"] / 1e6, "unit": "Mrays/s", "cores": cores, "kind": r["kind"],
                   "sample": f"{W}x{H} x {n} spp of the workload's {spp_total}, depth {depth} ({r['seconds']:.1f} s)",
                   "build": "g++ -O2 -ffp-contract=off (canonical rounding)"}
            try:
                f = run_cpu_reference(scene_name, W, H, n, depth, cores, fair=True)
                cpu["fair"] = {"value": f["rays"] / f["seconds"] / 1e6, "unit": "Mrays/s", "build": "g++ -O3 -march=x86-64-v3 (AVX2 + FMA contraction), same sources"}
            except Exception as e:   # noqa: BLE001
                cpu["fair"] = {"value": None, "unavailable": f"{type(e).__name__}: {e}"[:200]}
        except Exception as e:   # noqa: BLE001 — the baseline is a report, never a reason to lose the GPU number
            cpu = {"value": None, "unit": "Mrays/s", "cores": host_cores(), "kind": "unavailable", "sample": f"{type(e).__name__}: {e}"}

    res = None
    if rank == 0:
        res = {
            "value": value, "unit": "Mrays/s", "ms_per_step": total_ms / args.steps, "scaling": "strong",
            "spp_per_s": spp_total * args.steps / (total_ms * 1e-3), "rays_per_step": total_rays / args.steps,
            "config": workload_config(workload, W, H, spp_job, depth, n_tris),
            "detail": {"spp_per_step": spp_total, "spp_this_rank": spp_mine, "split": (split if world > 1 else None),
                       "gather": (args.gather if tile else ("nccl reduce + fused resolve" if world > 1 else None)),
                       "emulate_shard": (args.emulate_shard or None) if headline else None,
                       "scan_primitives": int(ptb.scan_prims()) if accel_used.startswith("brute") and args.math == "fast" else None,
                       "accel": accel_used, "math": args.math, "pipeline": args.pipeline,
                       "parallelism": "single GPU" if world == 1 else
                       ((f"tile-split x{world} (bands of 8 scanlines round-robin, reference seeds), gather fused into the path kernel: every rank stores "
                         "its pixels into rank 0's colors buffer over NVLink (CUDA IPC), barrier only" if p2p else
                         f"tile-split x{world} (bands of 8 scanlines round-robin, reference seeds) + NCCL reduce of the disjoint pixel sets per step") if tile
                        else f"spp-split x{world} of the fixed job + NCCL reduce(SUM) of the accumulation buffer + fused sum->mean->tone-map resolve per step"),
                       "l2": "512 MB buffer written between timed steps (L2 flush)", "setup_s": setup_s,
                       "lbvh": ptb.accel_stats() if accel_used == "lbvh" else None},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof, "parity_check": parity,
            "bit_exact_flavour": exact,
        }
        if cpu is not None:
            res["cpu_baseline"] = cpu
    # ---- give the memory back before the next workload
    barrier()
    ptb.set_accumulate_sum(0)
    ptb.set_pixel_shard(0, 1, 1)
    if shared is not None:
        multigpu.release_shared_colors(ptb, shared)
    for h in (scene.buffer_tris, scene.buffer_mats, scene.buffer_kd_tree, scene.buffer_kd_tree_shift, frame.h_colors, frame.h_rnds, frame.h_image):
        if h:
            ptb.release(h)
    del frame, scene
    torch.cuda.empty_cache()
    return res


def main():
    args = parse()
    scene_name, W, H, spp, depth, _split, _step = WORKLOADS[args.workload]
    W, H, spp = args.width or W, args.height or H, args.spp or spp
    if args.impl == "reference":
        return reference_arm(args, W, H, spp, depth, scene_name, args.workload)

    import torch
    import torch.distributed as dist
    from opencl_path_tracer_b200 import capi, multigpu

    cx = Ctx()
    cx.world = int(os.environ.get("WORLD_SIZE", "1"))
    cx.rank = int(os.environ.get("RANK", "0"))
    cx.local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this implementation has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(cx.local)
    cx.dev = torch.device("cuda", cx.local)
    if cx.world > 1:
        dist.init_process_group("nccl", device_id=cx.dev)      # control plane: the NCCL id, barriers around the timed region, max-over-ranks of the timings
    cx.ptb = capi.Ptb(cx.local)
    cx.stream = torch.cuda.Stream(device=cx.dev)      # a real (non-default) stream: the kernels, the library's collective and the timing events share it
    torch.cuda.set_stream(cx.stream)
    cx.ptb.set_stream(cx.stream.cuda_stream)
    multigpu.comm_init_from_dist(cx.ptb)                        # data plane: the library's own NCCL communicator (ptb_comm_*)
    cx.flush = torch.empty(512 << 20, dtype=torch.uint8, device=cx.dev)   # > 126 MB L2
    cx.peaks = {}
    try:
        cx.peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    cx.kernel_name = lambda accel_used: ("k_render (fused gen_ray + trace_ray + accumulate)" if accel_used != "lbvh" else
                                         "LBVH path kernel (closest hit over the wide LBVH + shading)")

    head = run_workload(cx, args, args.workload, True)
    extra = {}
    names = EXTRA_WORKLOADS if (args.workloads == "default" and args.workload == "cornell_1080p_1024spp_d4" and not (args.width or args.height or args.spp)) else \
        ([] if args.workloads in ("default", "none") else [w for w in args.workloads.split(",") if w])
    for w in names:
        extra[w] = run_workload(cx, args, w, False)

    if cx.rank == 0:
        line = {"metric": "path-trace throughput", "value": head["value"], "unit": "Mrays/s", "n_gpus": cx.world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
        line.update({k: v for k, v in head.items() if k not in ("value", "unit", "ms_per_step", "scaling")})
        if extra:
            line["workloads"] = extra
            line["gpu_launches"] = int(head["gpu_launches"] + sum(v["gpu_launches"] for v in extra.values()))
            line["gpu_launches_headline"] = int(head["gpu_launches"])
        EMIT(json.dumps(line))
    if cx.world > 1:
        cx.ptb.comm_destroy()
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
