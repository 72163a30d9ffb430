#!/usr/bin/env python
"""Developer loop: end-to-end rtf_trace_closest (pinned host rays in, host hits out) vs the pipeline chunk size.
    python tools/quick_e2e.py [--log2-rays 26] [--chunks 18,19,20,21,22,23]"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2-rays", type=int, default=26)
    ap.add_argument("--chunks", default="18,19,20,21,22,23")
    ap.add_argument("--reps", type=int, default=4)
    args = ap.parse_args()
    import torch

    from raytracerfacility_b200 import module, scenes, structs as S
    dev = torch.device("cuda", 0)
    rt = module.RayTracer(0)
    sd = scenes.sorghum_scene(50)
    rt.load_scene(sd, 63)
    lo, hi = scenes.scene_bounds(sd)
    n = 1 << args.log2_rays
    rays = scenes.incoherent_rays_torch(n, lo, hi, seed=3, device=dev)
    h_rays = torch.empty((n, 8), dtype=torch.float32, pin_memory=True)
    h_hits = torch.empty((n, 4), dtype=torch.float32, pin_memory=True)
    h_rays.copy_(rays)
    torch.cuda.synchronize()
    np_rays, np_hits = h_rays.numpy(), h_hits.numpy().view(S.HIT_DTYPE).reshape(-1)
    # plain copies for scale: H2D of the rays, D2H of the hits
    d_r = torch.empty_like(rays)
    for name, fn in (("h2d", lambda: d_r.copy_(h_rays, non_blocking=True)), ("d2h", lambda: h_hits.copy_(d_r[:, :4], non_blocking=True))):
        fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        nbytes = n * (32 if name == "h2d" else 16)
        print(f"{name}: {dt * 1e3:.2f} ms = {nbytes / dt / 1e9:.1f} GB/s", flush=True)
    for c in [int(x) for x in args.chunks.split(",")]:
        rt.set_option("pipe_chunk_rays", 1 << c)
        rt.trace_closest(np_rays, out=np_hits)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.reps):
            rt.trace_closest(np_rays, out=np_hits)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / args.reps
        print(f"chunk 2^{c}: {dt * 1e3:.2f} ms = {n / dt / 1e6:.0f} Mrays/s", flush=True)


if __name__ == "__main__":
    main()
