#!/usr/bin/env python
"""Developer loop: time the illumination entry point (mesh mode) on the C2 scene.
    python tools/quick_illum.py [--plants 50] [--samples 64] [--tris N] [--batch-log2 24]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--plants", type=int, default=50)
    ap.add_argument("--samples", type=int, default=64)
    ap.add_argument("--tris", type=int, default=0)
    ap.add_argument("--tri-begin", type=int, default=0, help="first triangle of the probed range (profiling: a range in the middle of the field)")
    ap.add_argument("--batch-log2", type=int, default=0)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--sort-rays", type=int, default=0, help="option sort_rays: sort bounce rays by (octant, origin cell)")
    ap.add_argument("--oracle-bvh", type=int, default=0, help="build the BVH with the CPU oracle and upload it (experiments)")
    ap.add_argument("--presplit", type=int, default=0, help="option presplit (0 = library default)")
    args = ap.parse_args()
    import torch

    from raytracerfacility_b200 import module, scenes, structs as S
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    rt = module.RayTracer(0)
    rt.set_stream(stream.cuda_stream)
    if args.batch_log2:
        rt.set_option("batch_paths", 1 << args.batch_log2)
    rt.set_option("sort_rays", args.sort_rays)
    sd = scenes.sorghum_scene(args.plants)
    rt.load_scene(sd, 63, presplit=args.presplit or None)
    if args.oracle_bvh:
        from oracle import pyoracle as _orc
        from raytracerfacility_b200 import structs as _S
        _o = _orc.OracleScene(sd, 63)
        _nodes, _tris = _o.debug(_S.DBG_WIDE_NODES), _o.debug(_S.DBG_WIDE_TRIS)
        print(f"oracle BVH: {len(_nodes)} nodes, {len(_tris)} triangle refs, depth {_o.depth}", flush=True)
        rt.debug_upload_bvh(_nodes, _tris, _o.depth)
    n = args.tris or rt.triangle_count
    t0 = min(args.tri_begin, rt.triangle_count - n)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=4, samples=args.samples)
    acc = torch.zeros(rt.triangle_count * 6, dtype=torch.float32, device=dev)
    rt.estimate_illumination_mesh_device(env, rp, 0x5EED, 1e-3, True, True, t0, t0 + n, 0, rp.samples, acc.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(args.reps):
        rt.estimate_illumination_mesh_device(env, rp, 0x5EED, 1e-3, True, True, t0, t0 + n, 0, rp.samples, acc.data_ptr())
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.reps
    rays, paths = rt.last_job_stats()
    print(f"illum: {ms:.2f} ms, {paths} paths, {rays} rays ({rays / paths:.2f}/path) -> {n * rp.samples * 2 / ms / 1e3:.1f} M tri-samples/s, "
          f"{rays / ms / 1e3:.1f} Mrays/s")
    rt.destroy()


if __name__ == "__main__":
    main()
