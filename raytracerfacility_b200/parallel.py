"""Multi-GPU sharding of the three entry points: one process per GPU (torch.distributed), BVH replicated,
work split, at most one collective per job.  The reference is single-GPU (CUDAModule.cpp:25
`cudaSetDevice(0)`), so this layer is new.  The sharding and the NCCL collectives live in the C library
(csrc/multi.cu); DistributedRayTracer below is a thin mirror of its rtf_*_sharded entry points, and the pure functions
of this file restate the partitioning rules for the CPU (gloo) tests of the host logic.

* illumination: SAMPLE-sharded (rank r traces samples [r*S/G, (r+1)*S/G) of every triangle), partial sums
  {sum E, sum D} per triangle, ONE all-reduce(sum) of 6*nTris floats, then finalize (IlluminationEstimation.cu:125-128).
  Alternative `triangle` mode: contiguous triangle ranges + all-gather, bit-identical to one GPU.
* LiDAR: interleaved blocks of the ray grid (block k of about 64 per rank belongs to rank k mod G; see
  rtf_scan_grid_shard_layout), no data-path collective (each rank writes its rays' own elements), optional grouped
  broadcasts to complete the arrays on every rank.
* camera: interleaved row bands (band k belongs to rank k mod G), all of a rank's bands traced as one job, grouped
  broadcasts of the float4 bands.

`compute` callbacks make the wiring testable on CPU with the gloo backend (tests inject the oracle there).
"""
import numpy as np


def split_range(n, world, rank):
    """Contiguous, balanced [begin, end) of n units for `rank` of `world` (the first n % world ranks get one more)."""
    base, rem = divmod(int(n), int(world))
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def sample_shard(samples, world, rank):
    return split_range(samples, world, rank)


def bind_to_gpu_numa_node(local_rank):
    """Pin this process to the CPU cores NVML reports as local to GPU `local_rank` (first-touch then places pinned host
    buffers on that NUMA node).  With several ranks feeding their GPUs from host memory (the host-buffer entry points)
    this keeps the H2D / D2H traffic off the inter-socket link.  Best effort: returns the core list, or None."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cores = [64 * w + b for w, mask in enumerate(words) for b in range(64) if (int(mask) >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cores = [c for c in cores if c in allowed]
        if cores:
            os.sched_setaffinity(0, cores)
            return cores
    except Exception:
        pass
    return None


def init_distributed(backend=None):
    """Initialise torch.distributed from the torchrun environment; returns (rank, world, local_rank)."""
    import os

    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and torch.cuda.is_available():
        bind_to_gpu_numa_node(local)
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, world, local


def illumination_sample_sharded(n_tris, samples, faces, compute_partial, finalize, all_reduce_sum, world, rank, accum):
    """accum: zero-initialised 6*n_tris float buffer (torch tensor, device or CPU).
    compute_partial(s0, s1, accum) adds the raw sums of samples [s0, s1) for all triangles;
    all_reduce_sum(accum) sums across ranks in place; finalize(accum, faces, samples) normalises in place."""
    s0, s1 = sample_shard(samples, world, rank)
    if s1 > s0:
        compute_partial(s0, s1, accum)
    if world > 1:
        all_reduce_sum(accum)
    finalize(accum, faces, samples)
    return accum


def illumination_triangle_sharded(n_tris, compute_range, all_gather, world, rank):
    """compute_range(t0, t1) -> finalised (t1-t0, 6) array for its triangle range; all_gather(list-of-sizes, local)
    returns the concatenation over ranks (bit-identical to a single-GPU run)."""
    t0, t1 = split_range(n_tris, world, rank)
    local = compute_range(t0, t1)
    if world == 1:
        return local
    sizes = [split_range(n_tris, world, r) for r in range(world)]
    return all_gather([e - b for b, e in sizes], local)


class DistributedRayTracer:
    """One RayTracer per rank (one process per GPU).  A thin mirror: sharding AND collectives run inside librtf_b200.so on
    the library's own NCCL communicator (csrc/multi.cu: rtf_comm_init_rank + the rtf_*_sharded entry points);
    torch.distributed only carries the 128-byte communicator id from rank 0 to the others at start-up.  Scene calls are
    replicated (every rank commits the same scene).  Stream contract: the context and every tensor handed to it live on
    ONE non-default torch stream (self.stream), so allocation, kernels and collectives are ordered; before a method
    returns a tensor, the caller's current stream is made to wait for self.stream (the `_device` entry points only enqueue),
    so `.cpu()` or any torch op on the result is ordered after the job whatever stream the caller runs on."""

    def __init__(self, rt, rank=0, world=1, group=None, stream=None):
        import torch
        self.rt, self.rank, self.world, self.group = rt, rank, world, group
        dev = torch.device("cuda", rt.device)
        if stream is None:
            stream = torch.cuda.Stream(device=dev)
            rt.set_stream(stream.cuda_stream)
        self.stream = stream
        cur_rank, cur_n = rt.comm_rank()
        if world > 1 and cur_n != world:
            import torch.distributed as dist
            box = [rt.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(box, src=0, group=group)
            rt.comm_init_rank(world, rank, box[0])

    def estimate_illumination_mesh(self, env, ray_props, seed, push=0.001, front=True, back=True, mode="sample"):
        """Returns a CUDA float32 tensor (n_tris, 6) = {energy, direction}, identical on all ranks."""
        import torch
        rt = self.rt
        n = rt.triangle_count
        with torch.cuda.stream(self.stream):
            out = torch.empty(n * 6, dtype=torch.float32, device=torch.device("cuda", rt.device))
            rt.estimate_illumination_mesh_sharded_device(env, ray_props, seed, push, front, back,
                                                         0 if mode == "sample" else 1, out.data_ptr())
        self._hand_over(out)
        return out.view(n, 6)

    def _hand_over(self, *tensors):
        """Results were allocated and written on self.stream and are about to be read on the caller's stream: make that
        stream wait for the job, and tell the caching allocator about the second stream (record_stream) so that a block the
        caller drops is not handed to the next job on self.stream while the caller's stream is still reading it."""
        import torch
        cur = torch.cuda.current_stream(tensors[0].device)
        cur.wait_stream(self.stream)
        if cur != self.stream:
            for x in tensors:
                x.record_stream(cur)

    def scan_grid(self, grid, gather=False):
        """Each rank traces its INTERLEAVED blocks of the ray grid (block k belongs to rank k mod world) into full-size
        device tensors; gather completes them on every rank.  Returns (t, prim, mine): t and prim are (n_blocks, block)
        views of the padded arrays (ray i = element [i // block, i % block]; the tail of the last block is padding),
        mine = slice(rank, None, world) selects this rank's blocks (`local_rays` flattens them without the padding)."""
        import torch
        rt = self.rt
        total = rt.scan_grid_ray_count(grid)
        with torch.cuda.stream(self.stream):
            dev = torch.device("cuda", rt.device)
            # one block more than any layout needs (blocks are at most 2^24 rays), so that the padded view always fits
            t = torch.empty(total + (1 << 24), dtype=torch.float32, device=dev)
            prim = torch.empty(total + (1 << 24), dtype=torch.int32, device=dev)
            block, n_blocks = rt.scan_grid_sharded_device(grid, gather, t.data_ptr(), prim.data_ptr())
        self._hand_over(t, prim)
        return t[:n_blocks * block].view(n_blocks, block), prim[:n_blocks * block].view(n_blocks, block), \
            slice(self.rank, None, self.world)

    def local_rays(self, total, arr2d):
        """This rank's rays of a (n_blocks, block) scan array as one flat tensor (block after block, padding cut off) and
        the first ray index of every block in it."""
        n_blocks, block = arr2d.shape
        rows = arr2d[self.rank::self.world]
        flat = rows.reshape(-1)
        owns_last = n_blocks > 0 and (n_blocks - 1) % self.world == self.rank
        n = flat.numel() - ((n_blocks * block - total) if owns_last else 0)
        return flat[:n], [k * block for k in range(self.rank, n_blocks, self.world)]

    def render_to_camera(self, env, camera, ray_props, outputs=(0,)):
        mask = 0
        for w in outputs:
            mask |= 1 << int(w)
        return self.rt.render_to_camera_sharded(env, camera, ray_props, mask)


# ---- camera: interleaved row bands (load balance: sky rows are cheap, canopy rows are not) + ONE all-gather --------
BAND_ROWS = 8


def camera_bands(height, world, rank, band_rows=BAND_ROWS):
    """Row ranges [(r0, r1), ...] rendered by `rank`: bands of `band_rows` rows dealt round-robin."""
    n_bands = (int(height) + band_rows - 1) // band_rows
    return [(b * band_rows, min((b + 1) * band_rows, int(height))) for b in range(rank, n_bands, world)]


def image_bands_sharded(height, width, channels, render_rows, all_gather_into, world, rank, frame, band_rows=BAND_ROWS):
    """frame: (height, width, channels) float32 tensor that render_rows(r0, r1) fills for its rows (in place).
    all_gather_into(out, local) gathers equal-size tensors from all ranks into out[world, ...].  Returns the full frame,
    identical on all ranks and bit-identical to a single-rank render (pixels keep their global RNG streams)."""
    import torch
    for r0, r1 in camera_bands(height, world, rank, band_rows):
        render_rows(r0, r1)
    if world == 1:
        return frame
    n_bands = (height + band_rows - 1) // band_rows
    k = (n_bands + world - 1) // world  # bands per rank, padded
    local = torch.zeros((k, band_rows, width, channels), dtype=frame.dtype, device=frame.device)
    for j, (r0, r1) in enumerate(camera_bands(height, world, rank, band_rows)):
        local[j, :r1 - r0] = frame[r0:r1]
    gathered = torch.empty((world, k, band_rows, width, channels), dtype=frame.dtype, device=frame.device)
    all_gather_into(gathered, local)
    # band b = j * world + r lives at gathered[r, j]
    full = gathered.permute(1, 0, 2, 3, 4).reshape(k * world * band_rows, width, channels)
    frame.copy_(full[:height])
    return frame


def default_band_rows(height, world):
    """16 bands per rank: enough interleaving to balance sky against canopy, few enough calls to keep batches large."""
    return int(height) if world == 1 else max(BAND_ROWS, int(height) // (16 * world))
