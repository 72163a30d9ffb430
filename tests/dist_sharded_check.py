"""Run under torchrun (one rank per GPU) by tests/test_gpu_multi.py: the library's NCCL path against one GPU."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch

    from raytracerfacility_b200 import module, parallel, scenes, structs as S
    rank, world, local = parallel.init_distributed("nccl")
    torch.cuda.set_device(local)
    rt = module.RayTracer(local)
    drt = parallel.DistributedRayTracer(rt, rank, world)
    sd = scenes.sorghum_scene(1)
    rt.load_scene(sd, 63)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=3, samples=16)
    whole = rt.estimate_illumination_mesh(env, rp, 0x5EED)  # every rank: the full job on its own GPU, plain entry point
    a = drt.estimate_illumination_mesh(env, rp, 0x5EED, mode="sample").cpu().numpy()
    b = drt.estimate_illumination_mesh(env, rp, 0x5EED, mode="triangle").cpu().numpy()
    err = np.abs(a[:, :3] - whole[:, :3]) / np.maximum(np.abs(whole[:, :3]), 1e-30)
    assert err.max() <= 1e-5, err.max()
    assert np.array_equal(b.view(np.uint32), whole.view(np.uint32)), "triangle-sharded result differs from one GPU"
    grid = scenes.lidar_grid(center=(0.0, 5.0, 0.0), size=(4.0, 4.0), distance=(0.01, 0.01))  # 160 000 rays: 40 blocks of 4096
    t1, p1, _ = rt.scan_grid(grid)
    tg, pg, mine = drt.scan_grid(grid, gather=True)  # interleaved blocks, completed on every rank
    n = len(t1)
    assert np.array_equal(tg.reshape(-1)[:n].cpu().numpy().view(np.uint32), t1.view(np.uint32))
    assert np.array_equal(pg.reshape(-1)[:n].cpu().numpy().view(np.uint32), p1)
    tl, pl, _ = drt.scan_grid(grid, gather=False)  # only this rank's blocks are written
    loc, starts = drt.local_rays(n, pl)
    blk = pl.shape[1]
    ref = np.concatenate([p1[b:b + blk] for b in starts]) if starts else p1[:0]
    assert np.array_equal(loc.cpu().numpy().view(np.uint32), ref), "interleaved blocks differ from one GPU"
    assert tl.shape[0] >= world, "grid too small to give every rank a block"
    cam1, camn = rt.create_camera(), rt.create_camera()
    for c in (cam1, camn):
        c.resize(256, 160)
        c.props.accumulate = 0
        c.set((0.0, 1.0, 3.0), (1.0, 0.0, 0.0, 0.0))
    rp2 = S.RayProperties(bounces=2, samples=4)
    rt.render_to_camera(env, cam1, rp2)
    drt.render_to_camera(env, camn, rp2, outputs=(S.OUT_COLOR, S.OUT_NORMAL, S.OUT_ALBEDO))
    for which in (S.OUT_COLOR, S.OUT_NORMAL, S.OUT_ALBEDO):
        assert np.array_equal(cam1.download(which).view(np.uint32), camn.download(which).view(np.uint32)), which
    ph = rt.phase_timings()
    mx = rt.comm_reduce([1.0 + rank], "max")
    assert mx[0] == float(world)
    rt.comm_barrier()
    if rank == 0:
        print("SHARDED OK", world, "ranks", ph, flush=True)
    cam1.destroy(), camn.destroy()
    rt.destroy()
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
