#!/usr/bin/env python
"""Developer loop: time the camera entry point on C2 / C5 (per-phase device times) — the profiling target for k_shade.
    python tools/quick_camera.py [--config C5] [--spp 8] [--reps 2] [--sss 0]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C5")
    ap.add_argument("--spp", type=int, default=8)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--sss", type=int, default=0, help="give the surface materials a subsurface term (k_shade<SSS>)")
    args = ap.parse_args()
    import torch

    from raytracerfacility_b200 import configs, module, structs as S
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    rt = module.RayTracer(0)
    rt.set_stream(stream.cuda_stream)
    cfg = configs.c5(args.spp) if args.config == "C5" else configs.c2()
    cfg.rp.samples = args.spp
    if args.sss:
        for m in cfg.scene.materials.values():
            if m.material_type != S.MAT_COMPRESSED_BTF:
                m.subsurface_factor = 0.4
                m.subsurface_radius[0] = m.subsurface_radius[1] = m.subsurface_radius[2] = 0.02
    rt.set_option("brdf", cfg.brdf)
    rt.load_scene(cfg.scene, 63)
    cam = configs.setup_camera(rt.create_camera(), cfg)
    rt.render_to_camera(cfg.env, cam, cfg.rp)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rt.set_option("profile", 1)
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(args.reps):
        rt.render_to_camera(cfg.env, cam, cfg.rp)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.reps
    ph = {k: round(v / args.reps, 2) for k, v in rt.profile_read().items() if v}
    rays, paths = rt.last_job_stats()
    print(f"{args.config} {cfg.width}x{cfg.height} {args.spp} spp sss={args.sss}: {ms:.2f} ms/frame, {paths / ms / 1e3:.1f} Mpaths/s, "
          f"{rays / ms / 1e3:.1f} Mrays/s, phases {ph}")
    rt.destroy()


if __name__ == "__main__":
    main()
