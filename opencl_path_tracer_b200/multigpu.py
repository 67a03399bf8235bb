"""Multi-GPU drivers: one process per GPU, torch.distributed for the plumbing.

The path shards naturally (pixels and samples are independent, the scene is read-only), so there is
exactly one exchange step per render:

* **spp-split** — every rank renders ``spp`` samples of EVERY pixel into a per-pixel SUM buffer with its
  own seed set (rank r takes outputs ``[r*W*H, (r+1)*W*H)`` of the host's ``minstd_rand0`` sequence, so
  rank 0 alone reproduces the reference's seeds), then one ``reduce(SUM)`` of the float buffer to rank 0
  over NCCL/NVLink and a ``sum / total_spp`` resolve.  Not stream-identical to a 1-GPU render of
  ``N*spp`` samples (different random numbers), validated by converged-image RMSE.
* **tile-split** — rank r renders the pixel groups ``(id // group) % N == r`` for ALL samples with the
  reference's per-pixel seeds, then the disjoint pixel sets are summed (every pixel is non-zero on
  exactly one rank) — bitwise identical to the 1-GPU image.

* **tile-split with a peer-memory gather** (`share_root_colors`) — rank 0 allocates the frame's ``colors`` buffer, exports
  it as a CUDA IPC handle, and every other rank binds the imported mapping as ITS ``colors`` kernel argument: the path
  kernel's own end-of-pixel store lands in rank 0's HBM over NVLink, pixel by pixel as paths finish, so the gather is
  fused into the kernel's epilogue and only a barrier (no data) follows the render.  Same bits as the all-reduce merge.

The collective runs on whatever backend the process group has (NCCL on the GPU box, gloo in the CPU
tests); only the render call needs the CUDA library.
"""
import numpy as np
import torch
import torch.distributed as dist


def rank_seed_skip(rank, width, height):
    """Offset into the host generator's output sequence for ``rank`` (main.cpp:522-526 continues)."""
    return rank * width * height


def shard_group(width, rows=8):
    """Pixel-group size of the tile split: bands of ``rows`` scanlines (matches the kernel's 16x8 blocks)."""
    return width * rows


def pixel_owner(width, height, nranks, group):
    """(H*W,) array: which rank renders which pixel id under ptb_set_pixel_shard(rank, nranks, group)."""
    ids = np.arange(width * height, dtype=np.int64)
    return ((ids // group) % nranks).astype(np.int32)


def owned_count(n_pixels, rank, nranks, group):
    """how many pixel ids rank owns under ptb_set_pixel_shard(rank, nranks, group) — what the persistent pool kernel's pixel
    cursor runs to (launch_pool, csrc/pt_pool.cuh): whole bands of `group` ids dealt round-robin, the image's last band may be short"""
    if nranks <= 1:
        return n_pixels
    bands = (n_pixels + group - 1) // group
    return sum(group if (b + 1) * group <= n_pixels else n_pixels - b * group for b in range(rank, bands, nranks))


def owned_pixel(k, rank, nranks, group):
    """pixel id of the k-th pixel rank owns (pool_item_pixel, csrc/pt_pool.cuh)"""
    if nranks <= 1:
        return k
    band, off = divmod(k, group)
    return (band * nranks + rank) * group + off


def spp_split_device(sample_index, nranks):
    """single-process spp split (ptb_set_gpus(n, PTB_SPLIT_SPP)): the GPU that renders the sample with this running index"""
    return sample_index % nranks


def spp_split_count(sample0, spp, nranks, d):
    """... and how many of the `spp` samples starting at index sample0 GPU d renders (render_multi, csrc/ptb_abi.cu)"""
    first = (d - sample0) % nranks
    return (spp - first + nranks - 1) // nranks if first < spp else 0


def reduce_sum_to_root(buf, group=None):
    """The one collective of the spp-split: SUM of the per-rank accumulation buffers at rank 0."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(buf, dst=0, op=dist.ReduceOp.SUM, group=group)
    return buf


def merge_spp_split(sum_buf, total_spp):
    """sum → mean on the root (the CUDA path uses ptb_resolve_sum; this is the host-side statement of it)."""
    return sum_buf / float(total_spp)


def merge_tile_split(local_colors, group=None):
    """Disjoint pixel sets: a SUM all-reduce is an exact gather (x + 0 == x for every finite x, and the
    non-owned pixels are exactly +0)."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(local_colors, op=dist.ReduceOp.SUM, group=group)
    return local_colors


def comm_init_from_dist(ptb, group=None):
    """Bring up the LIBRARY's communicator (ptb_comm_init: NCCL on the library's own stream) for the ranks of a
    torch.distributed process group: rank 0 makes the 128-byte id, the group carries it to the others (control plane
    only).  After this the data-path exchange is ptb.comm_reduce / ptb.comm_barrier — no torch collective touches a pixel."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return False
    rank = dist.get_rank(group)
    payload = [ptb.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(payload, src=0, group=group)
    ptb.comm_init(rank, dist.get_world_size(group), payload[0])
    return True


def split_spp(total_spp, nranks, rank):
    """samples of a FIXED total job that rank renders under the spp-split (strong scaling): the first total % n ranks take one more"""
    return total_spp // nranks + (1 if rank < total_spp % nranks else 0)


def release_shared_colors(ptb, handle, group=None):
    """Counterpart of share_root_colors: nobody may still be storing into the root's buffer when it is freed, and the peers'
    mappings must be closed before the allocation goes away (CUDA IPC rules)."""
    ptb.finish()
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.barrier(group=group)
        rank = dist.get_rank(group)
        if rank != 0:
            ptb.release(handle)
        dist.barrier(group=group)
        if rank == 0:
            ptb.release(handle)
    else:
        ptb.release(handle)


def share_root_colors(ptb, width, height, group=None, root=0):
    """Peer-memory gather of the tile split: returns a ptb handle to bind as ``colors`` (``scene.buffer_colors``).  On
    ``root`` it is a fresh ptb.alloc'ed W*H float4 buffer; on every other rank it is that same memory mapped through
    CUDA IPC (ptb_ipc_export / ptb_ipc_import), so the rank's kernels write their pixels straight into the root's frame.
    The 64-byte handle travels through the process group (control plane only)."""
    nbytes = width * height * 16
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return ptb.alloc(nbytes)
    payload = [None]
    h = None
    if rank == root:
        h = ptb.alloc(nbytes)
        ptb.finish()                      # the allocation's zero-fill has completed before any peer can store into it
        payload = [ptb.ipc_export(h)]
    dist.broadcast_object_list(payload, src=root, group=group)
    if rank != root:
        h = ptb.ipc_import(payload[0], nbytes)
    return h


def finish_peer_gather(ptb, group=None):
    """After a render into shared colors: every rank's stores are complete once its stream has drained; the barrier tells
    the root that all of them have.  No pixel data moves here."""
    ptb.finish()
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.barrier(group=group)


def render_progressive(scene, ptb, total_spp, refresh_spp, on_refresh=None, group=None):
    """Progressive tile-split render with a refresh cadence (BASELINE configs[4]; the reference refreshes its window every
    sample, main.cpp:670-681): ``refresh_spp`` samples per launch into the shared colors buffer (share_root_colors), then
    drain + barrier, then ``on_refresh(samples_done)`` on every rank (the root resolves / writes the image there — its
    buffer holds the whole frame's running mean).  The launches continue each pixel's sample index and RNG stream, so
    the final image is bitwise the one-launch, one-GPU render of ``total_spp`` samples."""
    done = 0
    while done < total_spp:
        n = min(refresh_spp, total_spp - done)
        scene.render_spp(n)
        finish_peer_gather(ptb, group)
        done += n
        if on_refresh is not None:
            on_refresh(done)
            # the root reads the shared frame inside on_refresh: nobody launches the next samples (which store into that
            # same buffer over NVLink) before it is done — otherwise a refresh could mix pixels at `done` and `done + n` samples
            if dist.is_initialized() and dist.get_world_size(group) > 1:
                dist.barrier(group=group)
    return done


class DeviceFrame:
    """colors / rnds / image buffers as torch CUDA tensors wrapped into ptb handles, so the same memory is
    visible to the path kernel (through the C-ABI) and to NCCL (through torch.distributed)."""

    def __init__(self, ptb, width, height, device):
        self.ptb, self.W, self.H = ptb, width, height
        n = width * height
        self.colors = torch.zeros(n, 4, dtype=torch.float32, device=device)
        self.rnds = torch.zeros(n, dtype=torch.int32, device=device)
        self.image = torch.zeros(n, 4, dtype=torch.float32, device=device)
        self.h_colors = ptb.wrap_device(self.colors.data_ptr(), n * 16)
        self.h_rnds = ptb.wrap_device(self.rnds.data_ptr(), n * 4)
        self.h_image = ptb.wrap_device(self.image.data_ptr(), n * 16)

    def attach(self, scene):
        """Point a host.Scene at these buffers instead of its own."""
        scene.buffer_colors, scene.buffer_rnds, scene.cl_screen = self.h_colors, self.h_rnds, self.h_image
        return scene
