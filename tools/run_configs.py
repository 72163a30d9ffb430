#!/usr/bin/env python
"""Times the five BASELINE.json configurations at their full sizes through the public API (one JSON line each).

    python tools/run_configs.py [--configs C1,C2,C3,C4,C5] [--reps 2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/run_configs.py --configs C3,C4,C5

C1  per-triangle illumination, 20 002-triangle sorghum + ground, 64 samples/triangle (also run on the host cores by the
    CPU oracle: parity <= 1e-5 and the CPU time are reported)
C2  camera render 1920x1080, 16 spp, 4 bounces, 1 M-triangle field (single GPU)
C3  illumination of the 10 M-triangle field, 256 samples/triangle, sample-sharded + one NCCL all-reduce
C4  LiDAR scan, 10^8 rays over the 5 M-triangle field, ray ranges per GPU, no collective
C5  compressed-BTF render 3840x2160, 64 spp, interleaved row bands per GPU + one all-gather
Times are device-side (CUDA events on the library's stream), max over ranks, scene already committed."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="C1,C2,C3,C4,C5")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--c3-samples", type=int, default=256)
    ap.add_argument("--c5-spp", type=int, default=64)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    from raytracerfacility_b200 import module, parallel, scenes, structs as S
    rank, world, local = parallel.init_distributed()
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        t = torch.tensor([float(x)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        t = torch.tensor([float(x)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def timed(fn):
        fn()  # warm-up (allocations, first touches)
        best = None
        for _ in range(args.reps):
            barrier()
            e0.record(stream)
            fn()
            e1.record(stream)
            barrier()
            ms = max_over_ranks(e0.elapsed_time(e1))
            best = ms if best is None else min(best, ms)
        return best

    def fresh(sd):
        rt = module.RayTracer(local)
        rt.set_stream(stream.cuda_stream)
        t0 = time.perf_counter()
        rt.load_scene(sd, 63)
        torch.cuda.synchronize()
        return rt, time.perf_counter() - t0

    def emit(d):
        if rank == 0:
            d["n_gpus"] = world
            print(json.dumps(d), flush=True)

    env = S.EnvironmentProperties()
    for cfg in args.configs.split(","):
        if cfg == "C1":
            sd = scenes.sorghum_scene(1)
            rt, bw = fresh(sd)
            rp = S.RayProperties(bounces=4, samples=64)
            drt = parallel.DistributedRayTracer(rt, rank, world)
            res = {}
            ms = timed(lambda: res.__setitem__("r", drt.estimate_illumination_mesh(env, rp, 0x5EED)))
            n = rt.triangle_count
            out = {"config": "C1", "what": "illumination, 20 002 triangles x 64 samples x 2 faces, 4 bounces, sky light",
                   "ms": ms, "triangle_samples_per_s": n * 64 * 2 / (ms * 1e-3), "commit_wall_s": bw}
            if rank == 0:  # the reference's own CPU-runnable case: the oracle on the host cores, full job
                from oracle import pyoracle as orc
                osc = orc.OracleScene(sd, 63)
                t0 = time.perf_counter()
                cpu, st = osc.illum_mesh(env, rp, seed=0x5EED, push=1e-3)
                cs = time.perf_counter() - t0
                gpu = res["r"].cpu().numpy()
                err = np.abs(gpu[:, :3] - cpu[:, :3]) / np.maximum(np.abs(cpu[:, :3]), 1e-30)
                out.update(cpu_ms=cs * 1e3, cpu_cores=orc.lib().orc_hardware_threads(), cpu_kind="port",
                           max_rel_err_vs_cpu=float(err.max()), rays=int(st.rays))
            emit(out)
            rt.destroy()
        elif cfg == "C2" and world == 1:
            rt, bw = fresh(scenes.sorghum_scene(50))
            cam = rt.create_camera()
            cam.resize(1920, 1080)
            cam.props.accumulate = 0
            cam.set((0.0, 1.5, 4.0), (0.9950042, -0.0998334, 0.0, 0.0))
            rp = S.RayProperties(bounces=4, samples=16)
            ms = timed(lambda: rt.render_to_camera(env, cam, rp))
            rays, paths = rt.last_job_stats()
            emit({"config": "C2", "what": "camera 1920x1080, 16 spp, 4 bounces, reference BRDF, 1 000 002 triangles", "ms": ms,
                  "Mpaths_per_s": paths / ms / 1e3, "Mrays_per_s": rays / ms / 1e3, "commit_wall_s": bw})
            cam.destroy(), rt.destroy()
        elif cfg == "C3":
            rt, bw = fresh(scenes.sorghum_scene(500))
            rp = S.RayProperties(bounces=4, samples=args.c3_samples)
            drt = parallel.DistributedRayTracer(rt, rank, world)
            res = {}
            ms = timed(lambda: res.__setitem__("r", drt.estimate_illumination_mesh(env, rp, 0x5EED)))
            n = rt.triangle_count
            rays = sum_over_ranks(rt.last_job_stats()[0])
            emit({"config": "C3", "what": f"illumination, {n} triangles x {rp.samples} samples x 2 faces, 4 bounces, "
                  f"sample-sharded x{world} + all-reduce of {n * 24} B", "ms": ms,
                  "triangle_samples_per_s": n * rp.samples * 2 / (ms * 1e-3), "Mrays_per_s": rays / ms / 1e3,
                  "mean_energy": float(res["r"][:, 0].mean().item()), "commit_wall_s": bw, "build_ms": rt.build_timings()})
            rt.destroy()
        elif cfg == "C4":
            sd = scenes.sorghum_scene(250)
            rt, bw = fresh(sd)
            lo, hi = scenes.scene_bounds(sd)
            ext = hi - lo
            grid = scenes.lidar_grid(center=((lo[0] + hi[0]) / 2, 30.0, (lo[2] + hi[2]) / 2), size=(float(ext[0]), float(ext[2])),
                                     distance=(float(ext[0]) / 10000.0, float(ext[2]) / 10000.0))
            drt = parallel.DistributedRayTracer(rt, rank, world)
            total = rt.scan_grid_ray_count(grid)
            res = {}
            ms = timed(lambda: res.__setitem__("r", drt.scan_grid(grid)))
            (b, e), t, prim = res["r"]
            hits = sum_over_ranks(int((prim != -1).sum().item()))
            # end to end: the rank's slice lands in host memory (12 B/ray incl. instance ids would be 1.2 GB; here t + prim)
            host_t = torch.empty(e - b, dtype=torch.float32).pin_memory()
            host_p = torch.empty(e - b, dtype=torch.int32).pin_memory()

            def e2e():
                (_, _), tt, pp = drt.scan_grid(grid)
                host_t.copy_(tt, non_blocking=True), host_p.copy_(pp, non_blocking=True)
            ms2 = timed(e2e)
            emit({"config": "C4", "what": f"LiDAR {total} rays over 5 000 002 triangles, ray ranges x{world}, no collective", "ms": ms,
                  "Mrays_per_s": total / ms / 1e3, "ms_incl_d2h": ms2, "Mrays_per_s_incl_d2h": total / ms2 / 1e3,
                  "hit_fraction": hits / total, "commit_wall_s": bw})
            rt.destroy()
        elif cfg == "C5":
            sd = scenes.sorghum_scene(50)
            sd.materials[1] = S.Material(material_type=S.MAT_COMPRESSED_BTF)
            sd.btfs[1] = scenes.synthetic_btf(seed=4, multiplier=0.0487)
            sd.material_btf[1] = 1
            rt, bw = fresh(sd)
            cam = rt.create_camera()
            cam.resize(3840, 2160)
            cam.props.accumulate = 0
            cam.set((0.0, 1.5, 4.0), (0.9950042, -0.0998334, 0.0, 0.0))
            rp = S.RayProperties(bounces=4, samples=args.c5_spp)
            drt = parallel.DistributedRayTracer(rt, rank, world)
            ms = timed(lambda: drt.render_to_camera(env, cam, rp, outputs=(S.OUT_COLOR,)))
            rays = sum_over_ranks(rt.last_job_stats()[0])
            img = cam.download(S.OUT_COLOR)
            emit({"config": "C5", "what": f"compressed-BTF camera 3840x2160, {rp.samples} spp, 4 bounces, synthetic MLVQ database "
                  f"(13x7x7x16, 256x256 texels), row bands x{world} + all-gather", "ms": ms,
                  "Mpaths_per_s": 3840 * 2160 * rp.samples / ms / 1e3, "image_mean": float(img[..., :3].mean()),
                  "commit_wall_s": bw})
            cam.destroy(), rt.destroy()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
