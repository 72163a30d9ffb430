"""Host-side sharding logic of raytracerfacility_b200.parallel on CPU: world_size-2 gloo processes, with the
oracle injected as the compute function (tests may use the oracle; the product never does)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from .conftest import ROOT


def test_split_range_partitions():
    from raytracerfacility_b200.parallel import split_range
    for n in (0, 1, 7, 64, 256, 20002, 10 ** 8):
        for w in (1, 2, 4, 8):
            parts = [split_range(n, w, r) for r in range(w)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(w - 1))
            sizes = [e - b for b, e in parts]
            assert max(sizes) - min(sizes) <= 1


def test_interleaved_lidar_shard_layout():
    """rtf_scan_grid_shard_layout (host arithmetic of the library, what the sharded LiDAR entry points use): the blocks
    of all ranks cover every ray of the grid exactly once, block k belongs to rank k mod world, only the grid's last block
    is partial, about 64 blocks per rank, and a rank's blocks are what DistributedRayTracer.local_rays cuts out."""
    from raytracerfacility_b200 import module, scenes
    from raytracerfacility_b200.parallel import DistributedRayTracer
    for side_x, side_y in ((400, 200), (130, 140), (10000, 10000), (1, 1), (4097, 3)):
        grid = scenes.lidar_grid(center=(0.0, 5.0, 0.0), size=(float(side_x), float(side_y)), distance=(1.0, 1.0))
        total = side_x * side_y
        for world in (1, 2, 3, 8):
            lay = [module.scan_grid_shard_layout(grid, world, r) for r in range(world)]
            block, n_blocks = lay[0][0], lay[0][1]
            assert all(l[0] == block and l[1] == n_blocks for l in lay)
            assert block & (block - 1) == 0 and 1 << 12 <= block <= 1 << 24
            assert n_blocks == (total + block - 1) // block
            assert n_blocks <= 64 * world or block in (1 << 12, 1 << 24)
            assert sum(l[2] for l in lay) == total
            seen = np.zeros(n_blocks, dtype=np.int32)
            for r in range(world):
                mine = list(range(r, n_blocks, world))
                seen[mine] += 1
                rays = sum(min(block, total - k * block) for k in mine)
                assert rays == lay[r][2], (side_x, side_y, world, r)
                if total > 1 << 20:
                    continue
                # the mirror's view of the same layout (no device needed for the index arithmetic)
                drt = DistributedRayTracer.__new__(DistributedRayTracer)
                drt.rank, drt.world = r, world
                arr = torch.arange(n_blocks * block, dtype=torch.int64).view(n_blocks, block)
                flat, starts = drt.local_rays(total, arr)
                assert starts == [k * block for k in mine] and flat.numel() == lay[r][2]
                if flat.numel():
                    assert int(flat.max()) < total and int(flat[0]) == (starts[0] if starts else 0)
            assert (seen == 1).all()
    with pytest.raises(module.RtfError):
        module.scan_grid_shard_layout(grid, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from oracle import pyoracle as orc
    from raytracerfacility_b200 import parallel, scenes, structs as S
    r, w, _ = parallel.init_distributed("gloo")
    assert (r, w) == (rank, world)
    sd = scenes.random_soup(60, seed=2, size=0.4)
    osc = orc.OracleScene(sd)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=33)  # odd: ragged sample shards
    n = osc.n_tris

    def compute_partial(s0, s1, acc):
        part, _ = osc.illum_mesh(env, rp, 9, 1e-3, s0=s0, s1=s1, normalise=False, nthreads=1)
        acc += torch.from_numpy(part.reshape(-1))

    def finalize(acc, faces, samples):
        a = acc.numpy().reshape(-1, 6)
        cnt = np.float32(faces * samples)
        a[:, :3] = a[:, :3] / cnt
        d = a[:, 3:]
        inv = np.float32(1.0) / np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2])
        a[:, 3:] = d * inv[:, None]

    acc = torch.zeros(n * 6, dtype=torch.float32)
    parallel.illumination_sample_sharded(n, rp.samples, 2, compute_partial, finalize,
                                         lambda a: dist.all_reduce(a, op=dist.ReduceOp.SUM), world, rank, acc)

    def compute_range(t0, t1):
        part, _ = osc.illum_mesh(env, rp, 9, 1e-3, tri_begin=t0, tri_end=t1, nthreads=1)
        return torch.from_numpy(part)

    def all_gather(sizes, local):
        outs = [torch.empty(s, 6) for s in sizes]
        for src in range(world):
            if src == rank:
                outs[src].copy_(local)
            dist.broadcast(outs[src], src=src)
        return torch.cat(outs)

    tri = parallel.illumination_triangle_sharded(n, compute_range, all_gather, world, rank)
    np.save(os.path.join(out_dir, f"sample_{rank}.npy"), acc.numpy().reshape(-1, 6))
    np.save(os.path.join(out_dir, f"tri_{rank}.npy"), tri.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_library_shard_rule_equals_the_python_restatement(mod):
    """csrc/multi.cu shards by rtf_shard_range; parallel.split_range restates it for the gloo tests of the host logic."""
    from raytracerfacility_b200 import parallel
    for n in (0, 1, 7, 8, 9, 1000, 10_000_002, 100_000_000, (1 << 40) + 3):
        for world in (1, 2, 3, 4, 8):
            for r in range(world):
                assert mod.shard_range(n, world, r) == parallel.split_range(n, world, r), (n, world, r)


def test_world2_gloo_sharded_illumination(tmp_path, orc, S, scenes):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    sd = scenes.random_soup(60, seed=2, size=0.4)
    osc = orc.OracleScene(sd)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=33)
    single, _ = osc.illum_mesh(env, rp, 9, 1e-3)
    s0, s1 = np.load(tmp_path / "sample_0.npy"), np.load(tmp_path / "sample_1.npy")
    assert np.array_equal(s0.view(np.uint32), s1.view(np.uint32))  # all ranks hold the same result
    assert np.allclose(s0[:, :3], single[:, :3], rtol=1e-5) and np.allclose(s0[:, 3:], single[:, 3:], rtol=1e-5, atol=1e-5)
    t0, t1 = np.load(tmp_path / "tri_0.npy"), np.load(tmp_path / "tri_1.npy")
    assert np.array_equal(t0.view(np.uint32), single.view(np.uint32))  # triangle sharding: bit-identical to 1 rank
    assert np.array_equal(t1.view(np.uint32), single.view(np.uint32))


def _camera_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from oracle import pyoracle as orc
    from raytracerfacility_b200 import parallel, scenes, structs as S
    from tests.golden.make_golden import camera_props
    parallel.init_distributed("gloo")
    osc = orc.OracleScene(scenes.random_soup(300, seed=11, extent=1.0, size=0.35))
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=2)
    cp = camera_props(20, 27)  # 27 rows: 4 bands of 8 rows (the last one ragged), dealt round-robin over 2 ranks
    frame = torch.zeros((27, 20, 4), dtype=torch.float32)

    def render_rows(r0, r1):
        col, _, _, _ = osc.camera(env, rp, cp, pixel_begin=r0 * 20, pixel_end=r1 * 20, nthreads=1)
        frame[r0:r1] = torch.from_numpy(col.reshape(r1 - r0, 20, 4))

    def all_gather_into(out, local):  # gloo: list form (NCCL uses all_gather_into_tensor in the product)
        parts = [torch.empty_like(local) for _ in range(world)]
        dist.all_gather(parts, local)
        out.copy_(torch.stack(parts))

    parallel.image_bands_sharded(27, 20, 4, render_rows, all_gather_into, world, rank, frame)
    np.save(os.path.join(out_dir, f"cam_{rank}.npy"), frame.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_world2_gloo_camera_bands(tmp_path, orc, S, scenes):
    from raytracerfacility_b200.parallel import camera_bands
    from tests.golden.make_golden import camera_props
    for h in (1, 8, 27, 1080, 2160):
        for w in (1, 2, 4, 8):
            rows = sorted(r for k in range(w) for b in camera_bands(h, w, k) for r in range(*b))
            assert rows == list(range(h))  # every row rendered exactly once
    world = 2
    mp.spawn(_camera_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    osc = orc.OracleScene(scenes.random_soup(300, seed=11, extent=1.0, size=0.35))
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=2)
    col, _, _, _ = osc.camera(env, rp, camera_props(20, 27))
    c0, c1 = np.load(tmp_path / "cam_0.npy"), np.load(tmp_path / "cam_1.npy")
    assert np.array_equal(c0.view(np.uint32), c1.view(np.uint32))
    assert np.array_equal(c0.reshape(-1, 4).view(np.uint32), col.view(np.uint32))  # bit-identical to one rank
