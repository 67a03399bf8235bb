"""N > 1 host logic on CPU: world_size-2 process groups over gloo (the GPU box runs the same code over NCCL).
The render call itself is stood in by the CPU oracle — only the sharding / seeding / merge logic is under test."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from opencl_path_tracer_b200 import host, multigpu
from tests import util

W, H, SPP = 48, 32, 3


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _oracle_render(rank_skip, spp):
    from oracle import pt_oracle_py as po
    surf = dict(host.load_ptbd(os.path.join(util.GOLDEN, "cornell_96x64.ptbd")))
    surf["W"], surf["H"] = W, H
    surf["camera"] = host.camera_at(W, H, 39.3, 0, 0, (277.5, 277.5, -800))
    o = po.OracleScene(surf, seeds=host.minstd_rand0_seeds(W * H, rank_skip))
    return o.render(spp, threads=2).reshape(-1, 4).copy()


def _worker(rank, world, port, mode, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        if mode == "tile":
            full = _oracle_render(0, SPP)                            # every rank has the reference's seeds
            owner = multigpu.pixel_owner(W, H, world, multigpu.shard_group(W, rows=8))
            local = torch.from_numpy(np.where((owner == rank)[:, None], full, np.float32(0.0)))
            merged = multigpu.merge_tile_split(local)
            np.save(os.path.join(out_dir, f"tile_{rank}.npy"), merged.numpy())
        else:
            mean = _oracle_render(multigpu.rank_seed_skip(rank, W, H), SPP)
            buf = torch.from_numpy(mean * np.float32(SPP))           # what the sum-mode kernel leaves in colors[]
            multigpu.reduce_sum_to_root(buf)
            if rank == 0:
                np.save(os.path.join(out_dir, "spp.npy"), multigpu.merge_spp_split(buf, SPP * world).numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["tile", "spp"])
def test_world_size_2_gloo(tmp_path, mode):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, mode, str(tmp_path)), nprocs=2, join=True)
    if mode == "tile":
        full = _oracle_render(0, SPP)
        for r in range(2):       # tile split is BITWISE the single-rank image, on every rank
            assert np.load(tmp_path / f"tile_{r}.npy").tobytes() == full.tobytes()
    else:
        m0 = _oracle_render(0, SPP)
        m1 = _oracle_render(multigpu.rank_seed_skip(1, W, H), SPP)
        want = (m0.astype(np.float64) + m1.astype(np.float64)) / 2
        got = np.load(tmp_path / "spp.npy")
        assert np.allclose(got, want, rtol=2e-6, atol=1e-7)
        assert not np.array_equal(m0, m1)                            # rank 1 really used a different seed set


def test_pixel_owner_partitions_the_image():
    for n in (1, 2, 4, 8):
        own = multigpu.pixel_owner(1920, 1080, n, multigpu.shard_group(1920, rows=8))
        counts = np.bincount(own, minlength=n)
        assert counts.sum() == 1920 * 1080 and counts.min() > 0
        assert counts.max() - counts.min() <= 1920 * 8               # at most one band of imbalance
    assert multigpu.rank_seed_skip(3, 1920, 1080) == 3 * 1920 * 1080


def test_progressive_cadence_and_single_rank_sharing():
    """Host logic of the progressive tile-split render without a device: launches of `refresh_spp` samples (the last one
    shorter), a drain after each, `on_refresh` with the running sample count; without a process group
    share_root_colors is a plain allocation of W*H float4."""
    class FakePtb:
        def __init__(self): self.finishes, self.allocs = 0, []
        def finish(self): self.finishes += 1
        def alloc(self, n): self.allocs.append(n); return 42

    class FakeScene:
        def __init__(self): self.launches = []
        def render_spp(self, n): self.launches.append(n)

    ptb, scene, seen = FakePtb(), FakeScene(), []
    assert multigpu.render_progressive(scene, ptb, 20, 8, on_refresh=seen.append) == 20
    assert scene.launches == [8, 8, 4] and seen == [8, 16, 20] and ptb.finishes == 3
    assert multigpu.share_root_colors(ptb, 16, 9) == 42 and ptb.allocs == [16 * 9 * 16]
    assert multigpu.render_progressive(scene, ptb, 0, 8) == 0 and scene.launches == [8, 8, 4]


def test_pool_cursor_mapping_matches_the_pixel_owner():
    """the persistent pool kernel hands out "the k-th pixel this rank owns": count and mapping must enumerate exactly the
    rank's pixels of pixel_owner(), in increasing id order, also when the last band of the image is short"""
    for (w, h, n, rows) in ((64, 40, 2, 8), (37, 19, 3, 8), (130, 9, 4, 8), (17, 129, 8, 8), (5, 3, 2, 8)):
        group = multigpu.shard_group(w, rows)
        own = multigpu.pixel_owner(w, h, n, group)
        total = 0
        for r in range(n):
            cnt = multigpu.owned_count(w * h, r, n, group)
            ids = [multigpu.owned_pixel(k, r, n, group) for k in range(cnt)]
            assert ids == np.nonzero(own == r)[0].tolist()
            total += cnt
        assert total == w * h


def test_spp_split_schedules():
    # one process per GPU: a FIXED job of spp samples cut n ways
    for spp in (1, 7, 64, 1024):
        for n in (1, 2, 3, 8):
            assert sum(multigpu.split_spp(spp, n, r) for r in range(n)) == spp
            assert max(multigpu.split_spp(spp, n, r) for r in range(n)) - min(multigpu.split_spp(spp, n, r) for r in range(n)) <= 1
    # single process (ptb_set_gpus): sample with running index i goes to GPU i % n, across calls of any size
    for n in (2, 3, 8):
        done = 0
        per_dev = [0] * n
        for spp in (1, 1, 5, 16, 3):
            counts = [multigpu.spp_split_count(done, spp, n, d) for d in range(n)]
            want = [sum(1 for j in range(spp) if multigpu.spp_split_device(done + j, n) == d) for d in range(n)]
            assert counts == want and sum(counts) == spp
            per_dev = [a + b for a, b in zip(per_dev, counts)]
            done += spp
        assert max(per_dev) - min(per_dev) <= 1
