#!/usr/bin/env python
"""The reduced workload tools/sanitize.sh runs under compute-sanitizer: every kernel family once, at sizes an instrumented
run finishes in a minute or two.  No oracle, no assertions beyond "the calls return RTF_OK and sane counts" — the tool's
report is the result.  Kernels reached: the builder (bounds, split, Morton, k_sort_pass with its decoupled look-back,
k_karras, k_collapse with its grid barrier, encode), refit, k_trace_rays (TMA ray tiles + mbarrier), k_scan_grid,
k_gen_illum / k_trace_queue / k_shade<0,0> (illumination), k_shade<1,0> (BSSRDF scene), k_shade<0,1> (BTF material, camera),
the flux tree (both entry points), probes, the instanced pre-kernel."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch

    from raytracerfacility_b200 import module, scenes, structs as S
    small = "--small" in sys.argv  # racecheck is the slowest tool
    rt = module.RayTracer(0)
    sd = scenes.sorghum_scene(1)
    rt.load_scene(sd, 63)
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(1 << (13 if small else 16), lo, hi, seed=3)
    hits = rt.trace_closest(rays)
    print("trace:", int((hits["prim"] != module.MISS).sum()), "hits of", len(rays), flush=True)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=3, samples=4 if small else 16)
    e6 = rt.estimate_illumination_mesh(env, rp, seed=0x5EED, push=1e-3, tri_begin=0, tri_end=256 if small else 2048)
    print("illum:", float(e6[:, :3].sum()), rt.last_job_stats(), flush=True)
    grid = scenes.lidar_grid(center=(0.0, 5.0, 0.0), size=(2.0, 2.0), distance=(0.04, 0.04) if small else (0.01, 0.01))
    t, prim, inst = rt.scan_grid(grid)
    print("lidar:", int((prim != module.MISS).sum()), "of", len(prim), flush=True)
    # probes + flux, allocating and static-scratch forms
    pr = rt.prepare_light_probes(0, 3000)
    rt.estimate_illumination(env, S.RayProperties(bounces=2, samples=2), pr, seed=7)
    print("flux:", rt.illumination_flux(pr)[:4], flush=True)
    d_e6 = torch.ones((rt.triangle_count, 6), dtype=torch.float32, device="cuda")
    d_out = torch.zeros(8, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    rt.illumination_flux_mesh_device_static(d_e6.data_ptr(), 0, rt.triangle_count, d_out.data_ptr())
    torch.cuda.synchronize()
    print("flux (static scratch):", d_out.cpu().numpy()[:4], rt.illumination_flux_mesh_device(d_e6.data_ptr(), 0, rt.triangle_count)[:4])
    # refit path: move an instance, commit with refit allowed
    rt.set_option("refit", 1)
    ih = max(sd.instances)
    g, m, x, comp = sd.instances[ih]
    x2 = np.array(x, dtype=np.float32).copy()
    x2.reshape(-1)[12] += np.float32(0.05)
    rt.set_instance(ih, g, m, x2, comp)
    rt.build_ias(63)
    print("refit: commit stats", rt.commit_stats() if hasattr(rt, "commit_stats") else "-", flush=True)
    rt.trace_closest(rays[:4096])
    rt.destroy()

    # BSSRDF + BTF + textures + cube map: the materials scene through the camera and the illumination entry point
    rt = module.RayTracer(0)
    sd = scenes.materials_scene()
    sd.materials[1].subsurface_factor = 0.4
    sd.materials[1].subsurface_radius = (0.2, 0.2, 0.2)
    sd.materials[1].subsurface_color = (0.9, 0.5, 0.4)
    rt.load_scene(sd, 63)
    env = S.EnvironmentProperties()
    e6 = rt.estimate_illumination_mesh(env, S.RayProperties(bounces=3, samples=4), seed=1, push=1e-3)
    print("materials illum:", float(e6[:, :3].sum()), flush=True)
    cam = rt.create_camera()
    cam.resize(48 if small else 96, 32 if small else 54)
    cam.props.accumulate, cam.props.gamma, cam.props.fov = 1, 2.2, 90.0
    cam.set((0.0, 1.5, 3.0), (0.9950042, -0.0998334, 0.0, 0.0))
    assert rt.render_to_camera(env, cam, S.RayProperties(bounces=3, samples=2))
    print("camera:", float(cam.download(S.OUT_COLOR).sum()), flush=True)
    cam.destroy()
    # instanced pre-kernel
    v, tri = scenes.sorghum_scene(1, ground=False).geometries[1]
    mats = np.stack([np.eye(4, dtype=np.float32).T.copy() for _ in range(3)])
    mats[1, 3, 0] = 1.0
    mats[2, 3, 2] = 1.0
    rt.set_geometry_instanced(9, v, tri, mats.reshape(3, 16), None)
    rt.destroy()
    print("SANITIZE WORKLOAD DONE", flush=True)


if __name__ == "__main__":
    main()
