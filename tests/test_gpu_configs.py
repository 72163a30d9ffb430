"""BASELINE.json configs C3-C5 at their full geometry / ray / pixel sizes, checked through size-independent properties
(the oracle covers the same entry points at sizes it finishes in seconds: test_gpu_entrypoints.py, test_gpu_materials.py).

* C3  10 M-triangle field, illumination: triangle-range shards are bit-identical to the whole job; sample shards sum to
      the whole job (1e-5); energies are bounded by the sky.
* C4  100 M LiDAR rays over the 5 M-triangle field: ray-range shards concatenate to the whole scan bit for bit; rays that
      reach the ground report t = scanner height; primitive ids are valid.
* C5  3840x2160 compressed-BTF render: row bands reproduce the full frame bit for bit.

AND against the CPU oracle on a sub-sample of the same full-size jobs (oracle/subsample.py; the counter RNG makes every
triangle / ray / pixel independent): every 1024-th triangle of C3 (<= 1e-5), every 1024-th ray of C4 (distances and
primitive ids bit-exact), every 4096-th pixel of C2 (both BRDFs) and C5 (<= 1e-5).  bench.py asserts nothing but reports
the same checks at the full sample counts."""
import numpy as np
import pytest

from .conftest import assert_bit_equal

pytestmark = pytest.mark.gpu


def _torch():
    import torch
    return torch


def test_c3_ten_million_triangles_shards(rt, scenes, S):
    torch = _torch()
    sd = scenes.sorghum_scene(500)
    rt.load_scene(sd, 63)
    n = rt.triangle_count
    assert n == 10_000_002
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=4, samples=4)
    dev = torch.device("cuda", 0)

    def run(t0, t1, s0, s1, finalize=True):
        acc = torch.zeros(n * 6, dtype=torch.float32, device=dev)
        torch.cuda.synchronize()  # the fill runs on torch's stream, the job on the library's
        rt.estimate_illumination_mesh_device(env, rp, 0x5EED, 1e-3, True, True, t0, t1, s0, s1, acc.data_ptr())
        torch.cuda.synchronize()
        if finalize:
            rt.illumination_finalize_device(acc.data_ptr() + 24 * t0, t1 - t0, 2, s1 - s0)
            torch.cuda.synchronize()
        return acc.view(n, 6)

    whole = run(0, n, 0, 4)
    e = whole[:, :3]
    assert torch.isfinite(e).all() and float(e.min()) >= 0.0 and float(e.max()) <= 1.5  # sky radiance 1, albedo < 1, |N| <= 1.1 (instance scale)
    assert 0.05 < float(e.mean()) < 0.6
    # triangle-range sharding (all-gather mode): bit-identical
    cut = 4_321_987
    a, b = run(0, cut, 0, 4), run(cut, n, 0, 4)
    assert torch.equal(a[:cut].view(torch.int32), whole[:cut].view(torch.int32))
    assert torch.equal(b[cut:].view(torch.int32), whole[cut:].view(torch.int32))
    # sample sharding (all-reduce mode): raw sums add up, normalisation after the reduce
    raw = run(0, n, 0, 1, finalize=False) + run(0, n, 1, 4, finalize=False)
    rt.illumination_finalize_device(raw.data_ptr(), n, 2, 4)
    torch.cuda.synchronize()
    assert torch.allclose(raw[:, :3], whole[:, :3], rtol=1e-5, atol=1e-7)


def test_c4_hundred_million_lidar_rays(rt, scenes, S):
    torch = _torch()
    sd = scenes.sorghum_scene(250)
    rt.load_scene(sd, 63)
    assert rt.triangle_count == 5_000_002
    lo, hi = scenes.scene_bounds(sd)
    ext = (hi - lo)
    grid = scenes.lidar_grid(center=((lo[0] + hi[0]) / 2, 30.0, (lo[2] + hi[2]) / 2), size=(float(ext[0]), float(ext[2])),
                             distance=(float(ext[0]) / 10000.0, float(ext[2]) / 10000.0))
    total = rt.scan_grid_ray_count(grid)
    assert 0.99e8 <= total <= 1.01e8
    dev = torch.device("cuda", 0)
    t = torch.empty(total, dtype=torch.float32, device=dev)
    prim = torch.empty(total, dtype=torch.int32, device=dev)
    rt.scan_grid_device(grid, 0, total, t.data_ptr(), prim.data_ptr())
    torch.cuda.synchronize()
    hit = prim != -1
    assert 0.5 < float(hit.float().mean()) <= 1.0
    assert int(prim[hit].min()) >= 0 and int(prim[hit].max()) < 5_000_002
    ground = (prim == 0) | (prim == 1)  # lowest instance handle = ground = primitives 0, 1
    assert int(ground.sum()) > total // 20
    assert float((t[ground] - 30.0).abs().max()) < 1e-4
    assert float(t[hit & ~ground].max()) < 30.0 and float(t[hit].min()) > 25.0  # plants are < 5 m tall
    # ray-range shards (8 ranks' worth) concatenate to the whole scan, bit for bit
    from raytracerfacility_b200.parallel import split_range
    for r in (0, 3, 7):
        b, e = split_range(total, 8, r)
        ts = torch.empty(e - b, dtype=torch.float32, device=dev)
        ps = torch.empty(e - b, dtype=torch.int32, device=dev)
        rt.scan_grid_device(grid, b, e, ts.data_ptr(), ps.data_ptr())
        torch.cuda.synchronize()
        assert torch.equal(ts.view(torch.int32), t[b:e].view(torch.int32)) and torch.equal(ps, prim[b:e])


def test_c5_btf_4k_row_bands(rt, scenes, S):
    sd = scenes.sorghum_scene(50)
    sd.materials[1] = S.Material(material_type=S.MAT_COMPRESSED_BTF)
    # beta 13, alpha 7, theta 7, phi 16, 256x256 texels, 4096 1-D slices; multiplier keeps (m*rgb)^2.2 / 256 below 1
    sd.btfs[1] = scenes.synthetic_btf(seed=4, multiplier=0.0487)
    sd.material_btf[1] = 1
    rt.load_scene(sd, 63)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=1)

    def camera():
        cam = rt.create_camera()
        cam.resize(3840, 2160)
        cam.props.accumulate, cam.props.gamma, cam.props.fov = 0, 2.2, 120.0
        cam.set((0.0, 1.5, 4.0))
        return cam
    full, bands = camera(), camera()
    assert rt.render_to_camera(env, full, rp)
    from raytracerfacility_b200.parallel import camera_bands
    for r in range(8):  # eight ranks' interleaved bands, rendered one after the other into one frame buffer
        for r0, r1 in camera_bands(2160, 8, r)[::9]:  # every 9th band of each rank: 1/9 of the frame
            assert rt.render_to_camera_rows(env, bands, rp, r0, r1)
    a, b = full.download(S.OUT_COLOR), bands.download(S.OUT_COLOR)
    checked = 0
    for r in range(8):
        for r0, r1 in camera_bands(2160, 8, r)[::9]:
            assert_bit_equal(b[r0:r1], a[r0:r1], f"rows {r0}:{r1}")
            checked += r1 - r0
    assert checked >= 2160 // 10
    img = a[..., :3]
    assert np.isfinite(img).all() and img.std() > 0.02  # BTF-shaded plants against the sky
    alb = full.download(S.OUT_ALBEDO)[..., :3]
    assert alb.std() > 0.01  # first-hit BTF colours vary with the view/illumination angles
    full.destroy(), bands.destroy()


# ---- oracle sub-sample parity at full size (recipes: raytracerfacility_b200/configs.py) ---------------------------------
def test_c3_oracle_subsample_every_1024th_triangle(rt, S):
    from oracle import subsample as sub
    from raytracerfacility_b200 import configs
    cfg = configs.c3(samples=16)  # full 10 000 002-triangle geometry; 16 of the 256 samples keep the oracle to seconds
    rt.load_scene(cfg.scene, 63)
    assert rt.triangle_count == cfg.triangles == 10_000_002
    gpu = rt.estimate_illumination_mesh(cfg.env, cfg.rp, configs.SEED, configs.PUSH)
    osc, _ = sub.build(cfg)
    par = sub.illumination(osc, cfg, gpu, 1024, configs.SEED, configs.PUSH)
    assert par["ok"], par
    # the sample-sharded entry point of the library (one rank: the all-reduce is skipped) gives the same numbers
    sharded = rt.estimate_illumination_mesh_sharded(cfg.env, cfg.rp, configs.SEED, configs.PUSH)
    assert sub.illumination(osc, cfg, sharded, 1024, configs.SEED, configs.PUSH)["ok"]


def test_c4_oracle_subsample_every_1024th_ray_bit_exact(rt, S):
    from oracle import subsample as sub
    from raytracerfacility_b200 import configs, module
    cfg = configs.c4()
    rt.load_scene(cfg.scene, 63)
    assert rt.triangle_count == 5_000_002
    total = rt.scan_grid_ray_count(cfg.grid)
    assert 0.99e8 <= total <= 1.01e8
    osc, _ = sub.build(cfg)
    b, e = module.shard_range(total, 8, 5)  # what rank 5 of 8 traces: a range that does not start on the stride
    for rb, re in ((0, total), (b, e)):
        t, prim, _ = rt.scan_grid(cfg.grid, rb, re)
        par = sub.lidar(osc, cfg, t, prim, 1024, ray_begin=rb)
        assert par["ok"] and par["hit_fraction_on_sample"] > 0.5, par


@pytest.mark.parametrize("which", ["C2 lambert_ggx", "C2 reference", "C5"])
def test_camera_configs_oracle_subsample_every_4096th_pixel(rt, S, which):
    from oracle import subsample as sub
    from raytracerfacility_b200 import configs
    cfg = {"C2 lambert_ggx": lambda: configs.c2(S.BRDF_LAMBERT_GGX), "C2 reference": lambda: configs.c2(S.BRDF_REFERENCE),
           "C5": lambda: configs.c5(spp=16)}[which]()  # C5: full 3840x2160 frame, 16 of the 64 spp
    rt.set_option("brdf", cfg.brdf)
    rt.load_scene(cfg.scene, 63)
    cam = configs.setup_camera(rt.create_camera(), cfg)
    assert rt.render_to_camera(cfg.env, cam, cfg.rp)
    img = cam.download(S.OUT_COLOR)
    osc, _ = sub.build(cfg)
    par = sub.camera(osc, cfg, cam.props, img, 4096)
    assert par["ok"], par
    assert img[..., :3].std() > 0.02
    cam.destroy()
