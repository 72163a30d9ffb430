#!/usr/bin/env python
"""Trace-side estimate of a two-level (instance) BVH — the reference's form, RayTracer.cu:1618-1715 — against the flat
world-space BVH this library builds, on the instanced field scenes (ONE 20 k-triangle plant under P transforms + ground).

What is measured, on the GPU, with the product's own traversal kernel and its visit counters:
  flat       the field scene as committed today: node visits and triangle tests per ray (instrumented kernel), Mrays/s.
  two-level  a LOWER BOUND of the work a two-level traversal does for the SAME rays: a ray must enter every instance whose
             world box it reaches before its closest hit (front-to-back order cannot cull those), so for every such
             (ray, instance) pair the ray is taken to the plant's object space (o' = M^-1 o, d' = M^-1 d, t unchanged) and
             traced through the single-plant BVH with tmax = the ray's final hit distance (the best culling any order
             could give).  Bottom-level visits are those counts summed; the top level adds at least 1 + entries / 8 node
             visits (an 8-wide tree over the instance boxes), and every entry costs a ray transform (about 30 instructions).
  check      the nearest pair hit must be the flat hit (same triangle up to float noise in t): the emulation traces the
             same geometry.
The single-plant BVH (about 1.4 MB) is L2-resident, so the pair trace also shows the per-visit speed of an L2-resident
structure with this kernel.
    python tools/exp_two_level.py [--plants 500] [--log2-sample 16]"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def world_boxes(sd):
    """[(instance handle, lo, hi, M(4x4 row-major float64))] from the scene description."""
    out = []
    for ih in sorted(sd.instances):
        g, _, x, _ = sd.instances[ih]
        v = sd.geometries[g][0]["position"].astype(np.float64)
        M = np.asarray(x, dtype=np.float64).reshape(4, 4).T  # column-major -> row-major
        w = v @ M[:3, :3].T + M[:3, 3]
        out.append((ih, g, w.min(0), w.max(0), M))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--plants", type=int, default=500)
    ap.add_argument("--log2-sample", type=int, default=16)
    ap.add_argument("--log2-rays", type=int, default=24)
    args = ap.parse_args()
    import torch

    from raytracerfacility_b200 import module, scenes
    dev = torch.device("cuda", 0)
    sd = scenes.sorghum_scene(args.plants)
    rt = module.RayTracer(0)
    rt.load_scene(sd, 63)
    lo, hi = scenes.scene_bounds(sd)
    n = 1 << args.log2_rays
    rays_d = scenes.incoherent_rays_torch(n, lo, hi, seed=3, device=dev)
    hits_d = torch.empty((n, 4), dtype=torch.float32, device=dev)
    for _ in range(3):
        rt.trace_closest_device(rays_d.data_ptr(), n, hits_d.data_ptr())
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        rt.trace_closest_device(rays_d.data_ptr(), n, hits_d.data_ptr())
    torch.cuda.synchronize()
    flat_ms = (time.perf_counter() - t0) / 5 * 1e3
    m = 1 << args.log2_sample
    rays = rays_d[:m].cpu().numpy()
    fh, fnv, ftv = rt.trace_closest(rays, counts=True)
    S = module.S if hasattr(module, "S") else None
    bvh_mb = (rt.debug_bytes(5) + rt.debug_bytes(6)) / 1e6
    print(f"flat: {rt.triangle_count} triangles, BVH {bvh_mb:.0f} MB, {n / flat_ms / 1e3:.0f} Mrays/s on 2^{args.log2_rays} rays; "
          f"per ray {fnv.mean():.2f} node visits, {ftv.mean():.2f} triangle tests (sample of {m})", flush=True)
    rt.destroy()

    # ---- which instances must a two-level traversal enter? ----
    t_hit = np.where(fh["prim"] != module.MISS, fh["t"], rays[:, 7]).astype(np.float64)
    o, d = rays[:, 0:3].astype(np.float64), rays[:, 4:7].astype(np.float64)
    inv = 1.0 / np.where(np.abs(d) < 1e-30, 1e-30, d)
    boxes = world_boxes(sd)
    pair_ray, pair_inst = [], []
    ground_entries = np.zeros(m, dtype=np.int64)
    for k, (ih, g, blo, bhi, M) in enumerate(boxes):
        t1, t2 = (blo - o) * inv, (bhi - o) * inv
        tn = np.minimum(t1, t2).max(1)
        tf = np.maximum(t1, t2).min(1)
        enter = (tn <= tf) & (tf >= rays[:, 3]) & (np.maximum(tn, rays[:, 3]) <= t_hit)
        if g != 1:  # the ground: two triangles, one leaf
            ground_entries += enter
            continue
        idx = np.nonzero(enter)[0]
        pair_ray.append(idx)
        pair_inst.append(np.full(len(idx), k))
    pair_ray, pair_inst = np.concatenate(pair_ray), np.concatenate(pair_inst)
    entries = np.bincount(pair_ray, minlength=m) + ground_entries
    print(f"two-level: a ray enters {entries.mean():.2f} instances on average before its closest hit (median {np.median(entries):.0f}, "
          f"95th percentile {np.percentile(entries, 95):.0f}, max {entries.max()}); {len(pair_ray)} (ray, plant) pairs", flush=True)

    # ---- bottom level: the pairs through the single-plant BVH, object-space rays ----
    one = scenes.sorghum_scene(1, ground=False)
    rt1 = module.RayTracer(0)
    rt1.load_scene(one, 63)
    Minv = np.stack([np.linalg.inv(b[4]) for b in boxes])
    A = Minv[pair_inst]
    po = np.einsum("nij,nj->ni", A[:, :3, :3], o[pair_ray]) + A[:, :3, 3]
    pd = np.einsum("nij,nj->ni", A[:, :3, :3], d[pair_ray])
    pr = np.zeros((len(pair_ray), 8), dtype=np.float32)
    pr[:, 0:3], pr[:, 4:7] = po, pd
    pr[:, 3] = rays[pair_ray, 3]
    pr[:, 7] = (t_hit[pair_ray] * (1 + 1e-5)).astype(np.float32)
    ph, pnv, ptv = rt1.trace_closest(pr, counts=True)
    b_nodes = np.bincount(pair_ray, weights=pnv, minlength=m) + ground_entries
    b_tris = np.bincount(pair_ray, weights=ptv, minlength=m) + 2 * ground_entries
    top = 1 + entries / 8.0
    # emulation check: nearest pair hit = flat hit
    best = np.full(m, np.inf)
    np.minimum.at(best, pair_ray, np.where(ph["prim"] != module.MISS, ph["t"], np.inf))
    plant_hit = (fh["prim"] != module.MISS) & (fh["prim"] >= 2)  # primitives 0, 1 are the ground
    agree = np.abs(best[plant_hit] - fh["t"][plant_hit]) <= 1e-4 * np.maximum(1.0, fh["t"][plant_hit])
    print(f"check: nearest object-space hit equals the flat hit for {agree.mean() * 100:.2f} % of the {plant_hit.sum()} rays that hit a plant")
    # speed of the L2-resident single-plant structure with the same kernel (pairs replicated to 2^22 rays)
    reps = max(1, (1 << 22) // len(pr))
    big = torch.from_numpy(np.tile(pr, (reps, 1))).to(dev)
    bh = torch.empty((len(big), 4), dtype=torch.float32, device=dev)
    for _ in range(3):
        rt1.trace_closest_device(big.data_ptr(), len(big), bh.data_ptr())
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        rt1.trace_closest_device(big.data_ptr(), len(big), bh.data_ptr())
    torch.cuda.synchronize()
    pair_ms = (time.perf_counter() - t0) / 5 * 1e3
    one_mb = (rt1.debug_bytes(5) + rt1.debug_bytes(6)) / 1e6
    ns_flat = flat_ms * 1e6 / n
    ns_pair = pair_ms * 1e6 / len(big)
    print(f"bottom level (single plant, BVH {one_mb:.1f} MB, L2-resident): {pnv.mean():.2f} node visits, {ptv.mean():.2f} triangle tests per "
          f"entry; {len(big) / pair_ms / 1e3:.0f} M entries/s = {ns_pair:.3f} ns per entry ({ns_pair / max(pnv.mean(), 1e-9) * 1e3:.1f} ps per node visit; "
          f"flat: {ns_flat:.3f} ns per ray, {ns_flat / fnv.mean() * 1e3:.1f} ps per node visit)")
    print(f"two-level lower bound per ray: {b_nodes.mean() + top.mean():.2f} node visits ({b_nodes.mean():.2f} bottom + {top.mean():.2f} top), "
          f"{b_tris.mean():.2f} triangle tests, {entries.mean():.2f} ray transforms   |   flat: {fnv.mean():.2f} node visits, {ftv.mean():.2f} triangle tests")
    est_ns = entries.mean() * ns_pair
    print(f"two-level time, bottom level alone at the measured L2-resident rate: {est_ns:.3f} ns per ray = {1e3 / est_ns:.0f} Mrays/s "
          f"against flat {n / flat_ms / 1e3:.0f} Mrays/s")
    rt1.destroy()


if __name__ == "__main__":
    main()
