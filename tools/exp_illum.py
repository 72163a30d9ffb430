#!/usr/bin/env python
"""Developer loop: the C3 illumination job (10 M triangles) at a reduced sample count under option variants.
    python tools/exp_illum.py [--plants 500] [--samples 16]"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--plants", type=int, default=500)
    ap.add_argument("--samples", type=int, default=16)
    ap.add_argument("--variants", default="base,sort,presplit4,presplit16,presplit1,batch21,batch23,batch24,bits30")
    args = ap.parse_args()
    import torch

    from raytracerfacility_b200 import module, scenes, structs as S
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    rt = module.RayTracer(0)
    rt.set_stream(stream.cuda_stream)
    sd = scenes.sorghum_scene(args.plants)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=4, samples=args.samples)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = sd.triangle_count()
    out = torch.empty(n * 6, dtype=torch.float32, device="cuda")

    def run(tag):
        rt.estimate_illumination_mesh_sharded_device(env, S.RayProperties(bounces=4, samples=2), 0x5EED, 1e-3, True, True, 0, out.data_ptr())
        rt.set_option("profile", 1)
        torch.cuda.synchronize()
        e0.record(stream)
        rt.estimate_illumination_mesh_sharded_device(env, rp, 0x5EED, 1e-3, True, True, 0, out.data_ptr())
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        ph = rt.profile_read()
        rt.set_option("profile", 0)
        rays, paths = rt.last_job_stats()
        print(f"{tag:12s} {ms:9.2f} ms  {n * args.samples * 2 / ms / 1e3:8.1f} M tri-samples/s  {rays / ms / 1e3:8.1f} Mrays/s  "
              f"refs {len(rt.debug(S.DBG_REF_TRI))} nodes {len(rt.debug(S.DBG_WIDE_NODES))} "
              f"trace {ph['trace_ms']:.1f} shade {ph['shade_ms']:.1f} gen {ph['gen_ms']:.1f} sort {ph['sort_ms']:.1f} "
              f"mean E {float(out.view(n, 6)[:, 0].mean()):.6f}", flush=True)

    for v in args.variants.split(","):
        bits = 63
        rt.set_option("sort_rays", 0)
        rt.set_option("presplit", 8)
        rt.set_option("batch_paths", 1 << 22)
        if v == "sort":
            rt.set_option("sort_rays", 1)
        elif v.startswith("presplit"):
            rt.set_option("presplit", int(v[8:]))
        elif v.startswith("batch"):
            rt.set_option("batch_paths", 1 << int(v[5:]))
        elif v == "bits30":
            bits = 30
        t0 = time.perf_counter()
        rt.load_scene(sd, bits)
        run(v)
    rt.destroy()


if __name__ == "__main__":
    main()
