#!/usr/bin/env python
"""Stand-in for racecheck / synccheck where compute-sanitizer cannot be run (it is closed on this GPU pool): the kernels
with hand-made synchronisation — k_sort_pass (decoupled look-back), k_collapse (grid barrier), k_trace_rays / k_trace_queue
(TMA tile re-armed after one __syncwarp, mbarrier double buffering), the wavefront queues (warp-aggregated appends) — are
run N times on the same input and every output is compared bit for bit with the first run's.  A data race or a missed
barrier that changes a result changes it differently from run to run; the parity tests pin run 1 to the oracle, this pins
runs 2..N to run 1 at sizes the oracle cannot redo (1 M and 10 M triangles, 2^24 rays).
    python tools/determinism_check.py [--plants 50] [--reps 20] [--log2-rays 24]"""
import argparse
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def digest(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).view(np.uint8).tobytes())
    return h.hexdigest()[:16]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--plants", type=int, default=50)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--log2-rays", type=int, default=24)
    ap.add_argument("--illum-tris", type=int, default=65536)
    args = ap.parse_args()
    import torch

    from raytracerfacility_b200 import module, scenes, structs as S
    rt = module.RayTracer(0)
    sd = scenes.sorghum_scene(args.plants)
    lo, hi = scenes.scene_bounds(sd)
    n = 1 << args.log2_rays
    rays = scenes.incoherent_rays_torch(n, lo, hi, seed=5, start=0, device=torch.device("cuda", 0))
    hits = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=4, samples=8)
    ext = hi - lo
    grid = scenes.lidar_grid(center=((lo[0] + hi[0]) / 2, 30.0, (lo[2] + hi[2]) / 2), size=(float(ext[0]), float(ext[2])),
                             distance=(float(ext[0]) / 2048, float(ext[2]) / 2048))  # 2048 x 2048 rays over the whole field
    ih = max(sd.instances)
    g_, m_, x_, c_ = sd.instances[ih]
    x_moved = np.array(x_, dtype=np.float32).copy()
    x_moved.reshape(-1)[12] += np.float32(0.125)
    first, bad = None, 0
    for rep in range(args.reps):
        if rep == 0:
            rt.load_scene(sd, 63)
        else:  # a full rebuild of the same scene: move an instance away and back (two commits, sort + collapse each)
            rt.set_instance(ih, g_, m_, x_moved, c_)
            rt.build_ias(63)
            rt.set_instance(ih, g_, m_, x_, c_)
            rt.build_ias(63)
        d_build = digest(rt.debug(S.DBG_SORTED_IDS), rt.debug(S.DBG_WIDE_NODES), rt.debug(S.DBG_WIDE_TRIS))
        torch.cuda.synchronize()
        rt.trace_closest_device(rays.data_ptr(), n, hits.data_ptr())
        torch.cuda.synchronize()
        d_trace = digest(hits.cpu().numpy())
        nt = min(args.illum_tris, rt.triangle_count)
        e6 = rt.estimate_illumination_mesh(env, rp, seed=0x5EED, push=1e-3, tri_begin=0, tri_end=nt)
        d_illum = digest(e6, np.asarray(rt.last_job_stats(), dtype=np.uint64))
        t, prim, inst = rt.scan_grid(grid)
        d_lidar = digest(t, prim, inst)
        cur = (d_build, d_trace, d_illum, d_lidar)
        if first is None:
            first = cur
            print(f"run 0: build {d_build} trace {d_trace} illum {d_illum} lidar {d_lidar}  ({rt.triangle_count} triangles, "
                  f"{n} rays, {nt} probe triangles, {len(prim)} LiDAR rays)", flush=True)
        elif cur != first:
            bad += 1
            print(f"run {rep}: DIFFERS from run 0: {cur}", flush=True)
    print(f"DETERMINISM {'OK' if bad == 0 else 'FAILED'}: {args.reps} runs, {bad} differing, plants={args.plants}", flush=True)
    rt.destroy()
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
