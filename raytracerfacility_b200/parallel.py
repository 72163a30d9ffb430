"""Multi-GPU sharding of the three entry points: one process per GPU (torch.distributed), BVH replicated,
work split, at most one collective per job.  The reference is single-GPU (CUDAModule.cpp:25
`cudaSetDevice(0)`), so this layer is new; it only ever calls the C ABI through module.RayTracer.

* illumination: SAMPLE-sharded (rank r traces samples [r*S/G, (r+1)*S/G) of every triangle), partial sums
  {sum E, sum D} per triangle, ONE all-reduce(sum) of 6*nTris floats, then finalize (IlluminationEstimation.cu:125-128).
  Alternative `triangle` mode: contiguous triangle ranges + all-gather, bit-identical to one GPU.
* LiDAR: contiguous ray ranges, no data-path collective (each rank writes its own slice), optional all-gather.
* camera: contiguous row bands per rank, all-gather of the float4 bands.

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
    """One RayTracer per rank + the collectives.  Scene calls are replicated (every rank builds the same BVH)."""

    def __init__(self, rt, rank=0, world=1, group=None):
        self.rt, self.rank, self.world, self.group = rt, rank, world, group

    def estimate_illumination_mesh(self, env, ray_props, seed, push=0.001, front=True, back=True, mode="sample"):
        """Returns a CUDA float32 tensor (n_tris, 6) = {energy, direction}, identical on all ranks."""
        import torch
        import torch.distributed as dist
        rt = self.rt
        n = rt.triangle_count
        faces = int(bool(front)) + int(bool(back))
        dev = torch.device("cuda", rt.device)
        if mode == "sample":
            accum = torch.zeros(n * 6, dtype=torch.float32, device=dev)

            def compute_partial(s0, s1, acc):
                rt.estimate_illumination_mesh_device(env, ray_props, seed, push, front, back, 0, n, s0, s1, acc.data_ptr())

            def finalize(acc, f, s):
                rt.illumination_finalize_device(acc.data_ptr(), n, f, s)

            def all_reduce_sum(acc):
                dist.all_reduce(acc, op=dist.ReduceOp.SUM, group=self.group)

            illumination_sample_sharded(n, ray_props.samples, faces, compute_partial, finalize, all_reduce_sum,
                                        self.world, self.rank, accum)
            return accum.view(n, 6)
        # triangle mode
        out = torch.zeros(n * 6, dtype=torch.float32, device=dev)
        t0, t1 = split_range(n, self.world, self.rank)
        if t1 > t0:
            rt.estimate_illumination_mesh_device(env, ray_props, seed, push, front, back, t0, t1, 0, ray_props.samples,
                                                 out.data_ptr())
            rt.illumination_finalize_device(out.data_ptr() + 24 * t0, t1 - t0, faces, ray_props.samples)
        if self.world > 1:
            parts = [out[6 * b:6 * e] for b, e in (split_range(n, self.world, r) for r in range(self.world))]
            if len({p.numel() for p in parts}) == 1:
                dist.all_gather(parts, parts[self.rank].clone(), group=self.group)
            else:
                for r, p in enumerate(parts):
                    dist.broadcast(p, src=r, group=self.group)
        return out.view(n, 6)

    def scan_grid(self, grid, gather=False):
        """Each rank traces its contiguous ray range into device SoA tensors (t, prim); optional all-gather."""
        import torch
        import torch.distributed as dist
        rt = self.rt
        total = rt.scan_grid_ray_count(grid)
        b, e = split_range(total, self.world, self.rank)
        dev = torch.device("cuda", rt.device)
        t = torch.empty(e - b, dtype=torch.float32, device=dev)
        prim = torch.empty(e - b, dtype=torch.int32, device=dev)
        if e > b:
            rt.scan_grid_device(grid, b, e, t.data_ptr(), prim.data_ptr())
        if not gather or self.world == 1:
            return (b, e), t, prim
        sizes = [split_range(total, self.world, r) for r in range(self.world)]
        full_t = torch.empty(total, dtype=torch.float32, device=dev)
        full_p = torch.empty(total, dtype=torch.int32, device=dev)
        for r, (rb, re) in enumerate(sizes):
            if r == self.rank:
                full_t[rb:re], full_p[rb:re] = t, prim
            dist.broadcast(full_t[rb:re], src=r, group=self.group)
            dist.broadcast(full_p[rb:re], src=r, group=self.group)
        return (0, total), full_t, full_p


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


class _DeviceArray:
    """Zero-copy torch view of a raw device pointer (__cuda_array_interface__)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def _camera_frame_tensor(camera, which, device):
    import torch
    h, w = camera.props.height, camera.props.width
    return torch.as_tensor(_DeviceArray(camera.device_buffer(which), (h, w, 4)), device=device)


def render_to_camera_sharded(drt, env, camera, ray_props, outputs=(0,)):
    """DistributedRayTracer camera render: every rank renders its interleaved row bands into its own camera buffers,
    then one all-gather per requested output (structs.OUT_COLOR / OUT_NORMAL / OUT_ALBEDO) completes the frame on all
    ranks.  Progressive accumulation state stays rank-local per band."""
    import torch
    import torch.distributed as dist
    rt = drt.rt
    dev = torch.device("cuda", rt.device)
    h, w = camera.props.height, camera.props.width
    band_rows = default_band_rows(h, drt.world)
    bands = camera_bands(h, drt.world, drt.rank, band_rows)
    modified = camera.props.modified
    ok = True
    for i, (r0, r1) in enumerate(bands):
        # every band of one frame must see the same frame id / modified flag (the C call advances them per call)
        camera.props.modified = modified if i == 0 else 0
        if i > 0:
            camera.props.frame_id = first_frame_id
        ok = rt.render_to_camera_rows(env, camera, ray_props, r0, r1) and ok
        if i == 0:
            first_frame_id = camera.props.frame_id - 1
    if not bands:
        rt.render_to_camera_rows(env, camera, ray_props, 0, 0)
    if drt.world > 1:
        for which in outputs:
            frame = _camera_frame_tensor(camera, which, dev)
            image_bands_sharded(h, w, 4, lambda r0, r1: None,
                                lambda out, local: dist.all_gather_into_tensor(out, local, group=drt.group), drt.world,
                                drt.rank, frame, band_rows)
    return ok


DistributedRayTracer.render_to_camera = render_to_camera_sharded
