#!/usr/bin/env python
"""Developer loop: time k_trace_rays on the C2 scene and check a sub-sample bit for bit against the CPU oracle.
    python tools/quick_trace.py [--plants 50] [--log2-rays 26] [--reps 5] [--check 131072]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--plants", type=int, default=50)
    ap.add_argument("--log2-rays", type=int, default=26)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--check", type=int, default=1 << 17)
    ap.add_argument("--sort", type=int, default=0, help="sort the rays by (octant, origin cell of 2^SORT per axis) first")
    ap.add_argument("--oracle-bvh", type=int, default=0, help="build the BVH with the CPU oracle and upload it (experiments)")
    ap.add_argument("--presplit", type=int, default=0, help="option presplit (0 = library default)")
    args = ap.parse_args()
    import torch

    from raytracerfacility_b200 import module, scenes
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    rt = module.RayTracer(0)
    rt.set_stream(stream.cuda_stream)
    sd = scenes.sorghum_scene(args.plants)
    rt.load_scene(sd, 63, presplit=args.presplit or None)
    if args.oracle_bvh:
        from oracle import pyoracle as _orc
        from raytracerfacility_b200 import structs as _S
        _o = _orc.OracleScene(sd, 63)
        _nodes, _tris = _o.debug(_S.DBG_WIDE_NODES), _o.debug(_S.DBG_WIDE_TRIS)
        print(f"oracle BVH: {len(_nodes)} nodes, {len(_tris)} triangle refs, depth {_o.depth}", flush=True)
        rt.debug_upload_bvh(_nodes, _tris, _o.depth)
    lo, hi = scenes.scene_bounds(sd)
    n = 1 << args.log2_rays
    rays = scenes.incoherent_rays_torch(n, lo, hi, seed=3, device=dev)
    if args.sort:
        g = 1 << args.sort
        lo_t = torch.tensor(lo, dtype=torch.float32, device=dev)
        ext = torch.tensor(hi - lo, dtype=torch.float32, device=dev)
        cell = ((rays[:, 0:3] - lo_t) / ext * g).clamp(0, g - 1).to(torch.int64)
        key = torch.zeros(n, dtype=torch.int64, device=dev)
        for b in range(args.sort):
            for a in range(3):
                key |= ((cell[:, a] >> b) & 1) << (3 * b + (2 - a))
        octant = ((rays[:, 4] < 0).to(torch.int64) << 2) | ((rays[:, 5] < 0).to(torch.int64) << 1) | (rays[:, 6] < 0).to(torch.int64)
        key |= octant << (3 * args.sort)
        rays = rays[torch.argsort(key)].contiguous()
        del key, cell, octant
    hits = torch.empty((n, 4), dtype=torch.float32, device=dev)
    for _ in range(3):
        rt.trace_closest_device(rays.data_ptr(), n, hits.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(args.reps):
        rt.trace_closest_device(rays.data_ptr(), n, hits.data_ptr())
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.reps
    print(f"k_trace_rays: {ms:.3f} ms for 2^{args.log2_rays} rays = {n / ms / 1e3:.1f} Mrays/s", flush=True)
    if args.check:
        from oracle import pyoracle as orc
        osc = orc.OracleScene(sd, 63)
        m = args.check
        sub = rays[:m].cpu().numpy()
        gh, gnv, gtv = rt.trace_closest(sub, counts=True)
        oh, onv, otv = osc.trace(sub, counts=True)
        ok_h = np.array_equal(gh.view(np.uint32), oh.view(np.uint32))
        ok_d = np.array_equal(hits[:m].cpu().numpy().view(np.uint32).reshape(-1), gh.view(np.uint32).reshape(-1))
        ok_c = np.array_equal(gnv, onv) and np.array_equal(gtv, otv)
        print(f"parity on {m} rays: hits bit-exact={ok_h} device-path==host-path={ok_d} visit counts equal={ok_c} "
              f"(mean nodes {gnv.mean():.2f}, tris {gtv.mean():.2f})")
        if not (ok_h and ok_d and ok_c):
            sys.exit(1)
    rt.destroy()


if __name__ == "__main__":
    main()
