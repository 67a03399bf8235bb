"""Two real GPUs, one process per GPU, NCCL.  Skipped on a single-GPU box; the host-side logic of the same
code runs world-size 2 over gloo in tests/test_multigpu_cpu.py.

* tile-split: the union of the two ranks' pixel shards after the SUM all-reduce is BITWISE the 1-GPU image (parity
  flavour, reference seeds).
* tile-split with the peer-memory gather (multigpu.share_root_colors): rank 1's kernels store into rank 0's colors
  buffer through a CUDA IPC mapping; after a barrier rank 0's buffer is BITWISE the 1-GPU image, also across three
  progressive launches (the later ones read the running mean back through the mapping).
* spp-split: rank r renders spp samples of every pixel from seed offset r*W*H into a SUM buffer, one NCCL reduce to rank
  0, resolve by the total: rank 0's own half is bitwise the 1-GPU render of spp samples, and the merged image agrees
  with a 1-GPU render of 2*spp samples to Monte-Carlo accuracy (different random numbers: RMSE bound, not equality).
"""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H, SPP, DEPTH = 256, 192, 16, 4


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _scene(ptb, seed_skip=0):
    from opencl_path_tracer_b200 import host
    s = host.Scene(ptb)
    s.init_Scene(W, H, seed_skip)
    s.add_Obj(os.path.join(ROOT, "scenes", "cornell.obj"), (555, 0, 0), (1, 1, 1), 0, 0)
    s.upload_Triangles(); s.upload_Materials()
    s.iterations = DEPTH
    cam = host.camera_at(W, H, 39.3, 0, 0, (277.5, 277.5, -800))
    s._camera = lambda: cam
    return s


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from opencl_path_tracer_b200 import capi, multigpu
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    ptb = capi.Ptb(rank)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    ptb.set_stream(stream.cuda_stream)
    ptb.set_accel(capi.ACCEL_BRUTE); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_MEGAKERNEL)

    # ---- tile split: reference seeds, my bands of 8 scanlines, SUM all-reduce of disjoint pixel sets
    group = multigpu.shard_group(W, rows=8)
    ptb.set_pixel_shard(rank, world, group)
    s = _scene(ptb)
    frame = multigpu.DeviceFrame(ptb, W, H, dev)
    ptb.write(frame.h_rnds, ptb.read(s.buffer_rnds, np.int32, W * H))      # the seeds init_Scene made
    frame.attach(s)
    s.render_spp(SPP)
    ptb.finish()
    multigpu.merge_tile_split(frame.colors)
    torch.cuda.synchronize()
    tile = frame.colors.cpu().numpy().reshape(H, W, 4)

    # ---- the same tile split with the gather fused into the kernels' own stores: every rank's colors argument IS rank 0's
    # buffer (CUDA IPC mapping over NVLink); after the barrier rank 0 holds the whole frame, no collective moved a pixel
    s1 = _scene(ptb)
    shared = multigpu.share_root_colors(ptb, W, H)
    s1.buffer_colors = shared
    refreshes, frames = [], []                    # progressive: 3 launches of 6/6/4 samples, the running mean is read back through the mapping

    def on_refresh(done):                         # the root reads the WHOLE shared frame at every refresh: it must hold `done` samples of every pixel
        refreshes.append(done)
        if rank == 0:
            frames.append(ptb.read(shared, np.float32, W * H * 4).reshape(H, W, 4).copy())

    multigpu.render_progressive(s1, ptb, SPP, 6, on_refresh=on_refresh)
    assert refreshes == [6, 12, 16]
    peer = ptb.read(shared, np.float32, W * H * 4).reshape(H, W, 4) if rank == 0 else np.zeros((H, W, 4), np.float32)
    dist.barrier()
    ptb.release(shared)
    ptb.set_pixel_shard(0, 1, 1)

    # ---- spp split: my own seed set, SUM buffer, one reduce to rank 0, resolve by the total
    ptb.set_accumulate_sum(1)
    s2 = _scene(ptb, multigpu.rank_seed_skip(rank, W, H))
    frame2 = multigpu.DeviceFrame(ptb, W, H, dev)
    ptb.write(frame2.h_rnds, ptb.read(s2.buffer_rnds, np.int32, W * H))
    frame2.attach(s2)
    s2.render_spp(SPP)
    ptb.finish()
    mine = frame2.colors.cpu().numpy().reshape(H, W, 4).copy()
    # the same exchange step through the LIBRARY's communicator (ptb_comm_reduce: NCCL on the library's stream) on a copy
    lib_buf = frame2.colors.clone()
    h_lib = ptb.wrap_device(lib_buf.data_ptr(), W * H * 16)
    assert multigpu.comm_init_from_dist(ptb) and ptb.comm_ranks() == world and ptb.comm_rank() == rank
    ptb.comm_reduce(h_lib, 0)
    ptb.comm_barrier()
    multigpu.reduce_sum_to_root(frame2.colors)
    torch.cuda.synchronize()
    if rank == 0:
        assert torch.equal(lib_buf.view(torch.int32), frame2.colors.view(torch.int32))      # a + b either way: the same bits
    ptb.release(h_lib)
    if rank == 0:
        ptb.resolve_sum(frame2.h_colors, SPP * world)
        ptb.finish()
    merged = frame2.colors.cpu().numpy().reshape(H, W, 4)
    ptb.set_accumulate_sum(0)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), tile=tile, mine=mine, merged=merged, peer=peer,
             refresh=np.stack(frames) if frames else np.zeros((0, H, W, 4), np.float32))
    dist.barrier()
    ptb.comm_destroy()
    dist.destroy_process_group()


def test_two_gpus_tile_split_is_bitwise_and_spp_split_converges(ptb, tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    from opencl_path_tracer_b200 import capi
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    r0, r1 = (np.load(tmp_path / f"rank{r}.npz") for r in (0, 1))

    def bits(a):
        return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)

    # single-GPU renders in this process
    ptb.set_accel(capi.ACCEL_BRUTE); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_MEGAKERNEL)
    s = _scene(ptb)
    partial = []
    for n in (6, 6, 4):                                    # the refresh points of the progressive run
        s.render_spp(n)
        partial.append(s.download_colors().copy())
    one = partial[-1]
    s.render_spp(SPP)                                      # ... continued to 2*SPP samples
    two = s.download_colors().copy()
    ptb.finish()
    assert (bits(r0["tile"]) == bits(one)).all() and (bits(r1["tile"]) == bits(one)).all()
    assert (bits(r0["peer"]) == bits(one)).all()           # peer-memory gather, three progressive launches: the same bits on rank 0
    assert len(r0["refresh"]) == 3                         # ... and every refresh frame is the 1-GPU frame at that sample count (no torn reads)
    for got, want in zip(r0["refresh"], partial):
        assert (bits(got) == bits(want)).all()
    # rank 0's SUM buffer / SPP is the 1-GPU mean up to the rounding of sum-vs-running-mean
    assert np.abs(r0["mine"][..., :3] / SPP - one[..., :3]).max() <= 2e-5 * max(1.0, float(one[..., :3].max()))
    # the merged image is a 2*SPP estimate from different random numbers: same expectation, Monte-Carlo distance
    merged = r0["merged"][..., :3]
    rel_mean = abs(float(merged.mean()) / float(two[..., :3].mean()) - 1.0)
    rmse = float(np.sqrt(((merged - two[..., :3]) ** 2).mean()) / two[..., :3].mean())
    noise = float(np.sqrt(((one[..., :3] - two[..., :3]) ** 2).mean()) / two[..., :3].mean())    # SPP vs 2*SPP of one stream
    assert rel_mean <= 0.02, rel_mean
    assert rmse <= 3.0 * noise + 1e-3, (rmse, noise)


# ---------------------------------------------------------------- several GPUs inside ONE process: ptb_set_gpus
def _two_gpus():
    import torch
    return torch.cuda.device_count() >= 2


def test_set_gpus_tile_split_in_one_process_is_bitwise_the_one_gpu_frame(ptb):
    """ptb_set_gpus(2, PTB_SPLIT_TILE): the host talks to one device, the second GPU stages its own scene copy through peer access
    and stores its bands of pixels into the primary's rnds / colors / image over NVLink.  Brute force (megakernel) and LBVH (pool
    kernel), per-sample launches like the reference host's and the in-kernel sample loop: all bitwise the 1-GPU result."""
    if not _two_gpus():
        pytest.skip("needs two GPUs")
    from opencl_path_tracer_b200 import capi, scenes, host

    def bits(a):
        return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)

    def render(n_gpus, accel, pipe, launches, scene_name):
        ptb.set_accel(accel); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(pipe)
        ptb.set_gpus(n_gpus, capi.SPLIT_TILE)
        ptb.reset_counters()                         # (after set_gpus: every participating GPU's counters)
        try:
            if scene_name == "cornell":
                s = _scene(ptb)
            else:
                d = scenes.describe(scene_name)
                s = host.Scene(ptb)
                s.build_trees = False
                s.init_Scene(W, H)
                scenes.populate(s, d)
                s.upload_Triangles(); s.upload_Materials()
                s.iterations = DEPTH
                cam = host.camera_at(W, H, d["fov"], d["yaw"], d["pitch"], d["eye"])
                s._camera = lambda cam=cam: cam
            if launches:
                for _ in range(3):
                    s.render()
            else:
                s.render_spp(5)
                s.render_spp(3)
            ptb.finish()
            out = s.download_colors().copy(), s.download_rnds().copy(), s.download_image().copy()
            rays = ptb.counters()["rays"]
        finally:
            ptb.set_gpus(1, capi.SPLIT_TILE)
        return out, rays

    for accel, pipe, launches, scene_name in ((capi.ACCEL_BRUTE, capi.PIPE_AUTO, False, "cornell"), (capi.ACCEL_BRUTE, capi.PIPE_AUTO, True, "cornell"),
                                              (capi.ACCEL_LBVH, capi.PIPE_POOL, False, "spheres100k"), (capi.ACCEL_LBVH, capi.PIPE_MEGAKERNEL, False, "small_mixed")):
        ptb.reset_counters()
        one, rays1 = render(1, accel, pipe, launches, scene_name)
        ptb.reset_counters()
        two, rays2 = render(2, accel, pipe, launches, scene_name)
        assert (bits(one[0]) == bits(two[0])).all() and (one[1] == two[1]).all() and (bits(one[2]) == bits(two[2])).all(), (accel, pipe, launches, scene_name)
        assert rays1 == rays2 > 0                        # the work counters of both GPUs add up to the 1-GPU count


def test_set_gpus_spp_split_in_one_process_converges_and_resolves(ptb):
    """ptb_set_gpus(2, PTB_SPLIT_SPP): ptb_render(2k) = k samples on each GPU from its own seed set, sums added by ONE kernel on the
    primary through peer loads and tone-mapped in the same pass.  Same expectation as a 1-GPU render of 2k samples."""
    if not _two_gpus():
        pytest.skip("needs two GPUs")
    from opencl_path_tracer_b200 import capi
    ptb.set_accel(capi.ACCEL_BRUTE); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_AUTO)
    s = _scene(ptb)
    s.render_spp(SPP)
    one = s.download_colors().copy()
    s.render_spp(SPP)
    two = s.download_colors().copy()
    ptb.set_gpus(2, capi.SPLIT_SPP)
    try:
        s2 = _scene(ptb)
        s2.render_spp(SPP)                               # 8 + 8
        s2.render_spp(SPP)                               # ... continued to 16 + 16: the sums persist across calls
        ptb.finish()
        merged, tex = s2.download_colors().copy(), s2.download_image().copy()
    finally:
        ptb.set_gpus(1, capi.SPLIT_TILE)
    m2 = float(two[..., :3].mean())
    rel_mean = abs(float(merged[..., :3].mean()) / m2 - 1.0)
    rmse = float(np.sqrt(((merged[..., :3] - two[..., :3]) ** 2).mean())) / m2
    noise = float(np.sqrt(((one[..., :3] - two[..., :3]) ** 2).mean())) / m2
    assert rel_mean <= 0.02 and rmse <= 3.0 * noise + 1e-3, (rel_mean, rmse, noise)
    # the image is the tone map of exactly the mean the resolve wrote (one pass: sum -> mean -> reinhard + sRGB)
    from oracle import pt_oracle_py as po
    fn = po.lib().pto_reinhard
    idx = np.random.default_rng(1).integers(0, W * H, 500)
    flat_c, flat_t = np.ascontiguousarray(merged.reshape(-1, 4)[:, :3]), tex.reshape(-1, 4)
    for i in idx:
        want = np.zeros(4, dtype=np.float32)
        fn(flat_c[i].ctypes.data, want.ctypes.data)
        same = (want.view(np.uint32) == flat_t[i].view(np.uint32)) | (np.isnan(want) & np.isnan(flat_t[i]))      # L == 0: a NaN, whichever payload
        assert same.all(), (i, want, flat_t[i])


def test_reference_host_renders_on_two_gpus_through_the_c_abi(tmp_path):
    """oracle/_ref/ref_host_b200 (the reference's own main.cpp over the shim) with the one added call ptb_set_gpus(2, tile):
    the image its Scene::render loop produces on two GPUs is bitwise its 1-GPU image."""
    import json, subprocess
    if not _two_gpus():
        pytest.skip("needs two GPUs")
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_host_b200")
    if not os.access(exe, os.X_OK):
        pytest.skip("oracle/_ref/ref_host_b200 not built")
    outs = []
    for gpus in (1, 2):
        out = str(tmp_path / f"g{gpus}.bin")
        r = subprocess.run([exe, "--w", "160", "--h", "120", "--spp", "4", "--iterations", "4", "--obj", os.path.join(ROOT, "scenes", "cornell.obj"),
                            "--pos", "555", "0", "0", "--math", "0", "--gpus", str(gpus), "--split", "tile", "--out", out], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-2000:]
        info = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
        assert info["error"] == ""
        outs.append((np.fromfile(out, dtype=np.float32), info["rays"]))
    assert (outs[0][0].view(np.uint32) == outs[1][0].view(np.uint32)).all() and outs[0][1] == outs[1][1]
