"""The three entry points through the C ABI vs the oracle: illumination estimation (IlluminationEstimation.cu),
virtual LiDAR (PointCloudScanning.cu, PointCloudScanner.cpp) and camera rendering (CameraRendering.cu).
Tolerances: hit ids / LiDAR t bit-exact; energies, directions, colours <= 1e-5 relative (north_star)."""
import os

import numpy as np
import pytest

from .conftest import GOLDEN, assert_bit_equal

pytestmark = pytest.mark.gpu

RTOL = 1e-5  # BASELINE.json north_star: "within 1e-5 relative for the same RNG sequence"


def close(a, b, what, rtol=RTOL, atol=0.0):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    err = np.abs(a - b) - (atol + rtol * np.abs(b))
    assert np.all(err <= 0), f"{what}: max excess {err.max():.3e} at {np.unravel_index(err.argmax(), err.shape)}"


def test_golden_small_cases(rt, S):
    from tests.golden.make_golden import camera_props, small_cases
    g = np.load(os.path.join(GOLDEN, "oracle_small.npz"))
    c = small_cases()
    rt.load_scene(c["illum_scene"], 63)
    irr = rt.estimate_illumination_mesh(c["env"], c["rp"], seed=0x5EED, push=1e-3)
    gold = g["irradiance"].view(np.float32)
    close(irr[:, :3], gold[:, :3], "energy")
    close(irr[:, 3:], gold[:, 3:], "direction", atol=1e-5)
    rays, paths = rt.last_job_stats()
    assert [rays, paths] == g["illum_rays"].tolist()
    t, prim, inst = rt.scan_grid(c["grid"])
    assert_bit_equal(t.view(np.uint32), g["lidar_t"], "lidar t")
    assert np.array_equal(prim, g["lidar_prim"])
    cam = rt.create_camera()
    cp = camera_props()
    cam.resize(cp.width, cp.height)
    for f, _ in S.CameraProps._fields_:
        setattr(cam.props, f, getattr(cp, f))
    assert rt.render_to_camera(c["env"], cam, S.RayProperties(bounces=2, samples=4))
    close(cam.download(S.OUT_COLOR).reshape(-1, 4), g["camera_color"].view(np.float32), "colour")
    close(cam.download(S.OUT_NORMAL).reshape(-1, 4), g["camera_normal"].view(np.float32), "normal", atol=1e-6)
    close(cam.download(S.OUT_ALBEDO).reshape(-1, 4), g["camera_albedo"].view(np.float32), "albedo")
    cam.destroy()


def test_c1_illumination_parity(rt, orc, S, scenes):
    """BASELINE config 1: 20k-triangle sorghum + ground, sky light, 64 samples/triangle, fixed seed."""
    sd = scenes.sorghum_scene(1)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=4, samples=64)
    gpu = rt.estimate_illumination_mesh(env, rp, seed=0x5EED, push=1e-3)
    cpu, st = osc.illum_mesh(env, rp, seed=0x5EED, push=1e-3)
    close(gpu[:, :3], cpu[:, :3], "energy")
    close(gpu[:, 3:], cpu[:, 3:], "direction", atol=1e-5)
    rays, paths = rt.last_job_stats()
    assert (rays, paths) == (st.rays, st.paths) and paths == 20002 * 64 * 2
    # flux integration of TriangleIlluminationEstimator::SampleLightProbeGroup (TriangleIlluminationEstimator.cpp:84-91)
    assert 0.05 < gpu[:, 0].mean() < 0.5


def test_illumination_aos_entry_point(rt, orc, S, scenes, mod):
    """CudaModule::EstimateIlluminationRayTracing signature: 268 B AoS probes in-out."""
    sd = scenes.random_soup(700, seed=6, size=0.3)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    env, rp = S.EnvironmentProperties(color=(1.0, 0.8, 0.6)), S.RayProperties(bounces=3, samples=32)
    probes = mod.probes_from_scene(sd)
    probes["front_face"][::3] = 0
    probes["back_face"][::5] = 0
    ref = probes.copy()
    rt.estimate_illumination(env, rp, probes, 77, 1e-3)
    osc.illum_probes(env, rp, ref, 77, 1e-3)
    both_off = (ref["front_face"] == 0) & (ref["back_face"] == 0)
    assert both_off.sum() > 10
    close(probes["energy"], ref["energy"], "energy")
    live = ~both_off
    close(probes["direction"][live], ref["direction"][live], "direction", atol=1e-5)
    assert_bit_equal(probes["a"], ref["a"], "inputs untouched")
    rt.estimate_illumination(env, rp, probes[:0], 77, 1e-3)  # empty input: no-op (RayTracer.cu:612-615)


def test_illumination_sharding_and_batches(rt, S, scenes):
    """Sample-range partial sums add up; triangle ranges and small batches are bit-identical (multi-GPU sharding)."""
    sd = scenes.random_soup(900, seed=2, size=0.4)
    rt.load_scene(sd, 63)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=64)
    whole = rt.estimate_illumination_mesh(env, rp, 9, normalise=False)
    a = rt.estimate_illumination_mesh(env, rp, 9, s0=0, s1=32, normalise=False)
    b = rt.estimate_illumination_mesh(env, rp, 9, s0=32, s1=64, normalise=False)
    close(a + b, whole, "sample shards", rtol=1e-5, atol=5e-5)
    c = rt.estimate_illumination_mesh(env, rp, 9, tri_begin=100, tri_end=333, normalise=False)
    assert_bit_equal(c, whole[100:333], "triangle shard")
    rt.set_option("batch_paths", 4096)
    d = rt.estimate_illumination_mesh(env, rp, 9, normalise=False)
    assert_bit_equal(d, whole, "batch size independence")
    rt.set_option("sort_rays", 0)
    e = rt.estimate_illumination_mesh(env, rp, 9, normalise=False)
    assert_bit_equal(e, whole, "ray sorting does not change results")


def test_illumination_materials_alpha_emission_skydome(rt, orc, S, scenes):
    sd = scenes.random_soup(400, seed=12, size=0.35)
    sd.materials[1] = S.Material(albedo=(0.8, 0.7, 0.6), metallic=0.4, roughness=0.6, emission=0.2, transmission=0.5)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    rp = S.RayProperties(bounces=3, samples=16)
    for env in (S.EnvironmentProperties(gamma=2.2, color=(0.9, 0.8, 1.0)),
                S.EnvironmentProperties(lighting_type=S.ENV_SINGLE_LIGHT_SOURCE, light_size=0.1, sun_direction=(0.3, 0.9, 0.1))):
        gpu = rt.estimate_illumination_mesh(env, rp, 5)
        cpu, st = osc.illum_mesh(env, rp, 5, 1e-3)
        close(gpu[:, :3], cpu[:, :3], f"energy env {env.lighting_type}", rtol=2e-5)
        assert rt.last_job_stats()[0] == st.rays
    sky = S.EnvironmentProperties(lighting_type=S.ENV_SKYDOME)  # expf/powf: CUDA vs glibc differ by ulps
    gpu = rt.estimate_illumination_mesh(sky, rp, 5)
    cpu, _ = osc.illum_mesh(sky, rp, 5, 1e-3)
    close(gpu[:, :3], cpu[:, :3], "energy skydome", rtol=1e-4)


def test_lidar_grid_and_aos(rt, orc, S, scenes):
    sd = scenes.sorghum_scene(4)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    grid = scenes.lidar_grid(center=(0.0, 30.0, 0.0), size=(2.6, 2.8), distance=(0.02, 0.02))
    n = rt.scan_grid_ray_count(grid)
    assert n == orc.scan_grid_count(grid) == 130 * 140
    t, prim, inst = rt.scan_grid(grid)
    ot, oprim, oinst, _ = osc.scan_grid(grid)
    assert_bit_equal(t, ot, "t")
    assert np.array_equal(prim, oprim) and np.array_equal(inst, oinst)
    assert (prim != 0xFFFFFFFF).mean() > 0.5
    # a ray range (multi-GPU sharding by ray range) is the same slice
    t2, prim2, _ = rt.scan_grid(grid, 2345, 13456)
    assert_bit_equal(t2, t[2345:13456], "range t") and np.array_equal(prim2, prim[2345:13456])
    # CudaModule::SamplePointCloud signature: 112 B AoS in-out
    rays = orc.scan_grid_rays(grid, 0, n)[::7]
    smp = np.zeros(len(rays) + 2, dtype=S.PC_SAMPLE_DTYPE)
    smp["start"][:-2], smp["direction"][:-2] = rays[:, 0:3], rays[:, 4:7]
    smp["start"][-2:], smp["direction"][-2:] = (0, 50, 0), (0, 1, 0)
    ref = smp.copy()
    p = rt.sample_point_cloud(S.EnvironmentProperties(), smp, want_prim_ids=True)
    op = osc.scan_point_cloud(ref)
    assert np.array_equal(p, op)
    assert np.array_equal(smp["hit"], ref["hit"]) and np.array_equal(smp["handle"], ref["handle"])
    for f in ("position", "normal", "tangent", "color", "tex_coord", "data"):
        assert_bit_equal(smp["hit_info"][f], ref["hit_info"][f], f"hit_info.{f}")
    assert smp["hit"][-1] == 0 and smp["handle"][-1] == 0


def test_point_cloud_against_the_reference_programs(mod, S):
    """The device's SamplePointCloud against tests/golden/programs_ref.npz = what the REFERENCE's own __raygen__PCS /
    __closesthit__PCS_R / __miss__PCS_R programs (ptx/PointCloudScanning.cu, RayDataDefinations.hpp) write when compiled
    unmodified for the host (oracle/ref_optix_driver.cpp) — no oracle in between, and no random numbers on this path.
    The whole 112 B record byte for byte, with or without instance transforms."""
    import os
    from .golden import make_golden as G
    ref = np.load(os.path.join(os.path.dirname(__file__), "golden", "programs_ref.npz"))
    all_scenes = G.programs_ref_scenes()
    for name in ("plants", "mixed"):
        sd = all_scenes[name]
        rt = mod.RayTracer(0)  # one context per scene
        rt.load_scene(sd)
        smp = ref[f"pcs_{name}_in"].view(S.PC_SAMPLE_DTYPE).copy()
        want = ref[f"pcs_{name}"].view(S.PC_SAMPLE_DTYPE)
        rt.sample_point_cloud(S.EnvironmentProperties(), smp)
        rt.destroy()
        assert np.array_equal(smp["hit"], want["hit"]) and np.array_equal(smp["handle"], want["handle"]), name
        assert smp.tobytes() == want.tobytes(), name


def test_config1_statistics_against_the_reference_programs(rt, S, scenes):
    """BASELINE config 1 on the device against tests/golden/c1_ref.npz = the REFERENCE's own illumination programs run on
    the host with the reference's own LCG under two seeds: the device's estimate (keyed Philox draws) must be
    statistically indistinguishable from them (tests/test_ref_programs.py: c1_statistics)."""
    import os
    from .test_ref_programs import c1_statistics
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "c1_ref.npz"))
    rt.load_scene(scenes.sorghum_scene(1))
    e = rt.estimate_illumination_mesh(S.EnvironmentProperties(), S.RayProperties(bounces=4, samples=64), 5)
    c1_statistics(e[:, :3], rt.last_job_stats()[0], g)


def test_camera_render_parity_and_accumulation(rt, orc, S, scenes):
    sd = scenes.sorghum_scene(1)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=4, samples=4)
    cam = rt.create_camera()
    cam.resize(96, 54)
    cam.props.accumulate, cam.props.gamma, cam.props.fov = 1, 2.2, 120.0
    cam.set((0.0, 1.5, 4.0), (0.9950042, -0.0998334, 0.0, 0.0))  # slight downward pitch
    assert rt.render_to_camera(env, cam, rp) and cam.props.frame_id == 1
    cp = S.CameraProps.from_buffer_copy(cam.props)
    cp.frame_id = 0
    col, nrm, alb, st = osc.camera(env, rp, cp)
    g = cam.download(S.OUT_COLOR).reshape(-1, 4)
    close(g, col, "colour frame 0")
    close(cam.download(S.OUT_NORMAL).reshape(-1, 4), nrm, "normal", atol=1e-6)
    close(cam.download(S.OUT_ALBEDO).reshape(-1, 4), alb, "albedo")
    assert rt.last_job_stats() == (st.rays, st.paths)
    assert g[:, :3].std() > 0.05  # an actual image: plant, ground and sky
    # progressive accumulation in gamma space (CameraRendering.cu:113-128): frame ids 1 and 2
    for fid in (1, 2):
        prev = cam.download(S.OUT_COLOR).reshape(-1, 4)
        assert rt.render_to_camera(env, cam, rp) and cam.props.frame_id == fid + 1
        cp.frame_id = fid
        cp.accumulate = 1
        col, _, _, _ = osc.camera(env, rp, cp, prev_color=prev)
        close(cam.download(S.OUT_COLOR).reshape(-1, 4), col, f"colour frame {fid}", rtol=2e-5)
    # a changed setting resets the accumulation (RayTracer.cu:253-269)
    rp2 = S.RayProperties(bounces=2, samples=4)
    assert rt.render_to_camera(env, cam, rp2) and cam.props.frame_id == 1
    # depth output goes to the albedo buffer (CameraRendering.cu:135-140)
    cam.props.output_type = S.OUT_DEPTH
    cam.props.modified = 1
    assert rt.render_to_camera(env, cam, rp2)
    cp = S.CameraProps.from_buffer_copy(cam.props)
    cp.frame_id = 0
    _, _, alb, _ = osc.camera(env, rp2, cp)
    close(cam.download(S.OUT_ALBEDO).reshape(-1, 4), alb, "depth", rtol=2e-5)
    cam.destroy()


def test_cuda_module_facade(mod, orc, S, scenes):
    """The static façade with the reference's names (CUDAModule.hpp:30-45)."""
    mod.CudaModule.Init()
    rt = mod.CudaModule.GetRayTracer()
    sd = scenes.random_soup(200, seed=3, size=0.4)
    rt.load_scene(sd)
    osc = orc.OracleScene(sd)
    probes = mod.probes_from_scene(sd)
    ref = probes.copy()
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=1, samples=8)
    mod.CudaModule.EstimateIlluminationRayTracing(env, rp, probes, 1, 0.001)
    osc.illum_probes(env, rp, ref, 1, 0.001)
    close(probes["energy"], ref["energy"], "energy")
    mod.CudaModule.Terminate()
    with pytest.raises(mod.RtfError):
        mod.CudaModule.GetRayTracer()


def test_ray_sort_option_changes_nothing_but_order(rt, S, scenes):
    """Option "sort_rays": bounce rays sorted by (direction octant, origin cell) between wavefronts.  Per-path state is
    addressed by the path index a ray carries, so irradiance, camera buffers and ray counts are bit-identical."""
    sd = scenes.sorghum_scene(1)
    rt.load_scene(sd, 63)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=4, samples=16)
    cam = rt.create_camera()
    cam.resize(96, 54)
    cam.set((0.0, 1.5, 4.0), (0.9950042, -0.0998334, 0.0, 0.0))
    out = {}
    for mode in (0, 1):
        rt.set_option("sort_rays", mode)
        irr = rt.estimate_illumination_mesh(env, rp, seed=0x5EED, push=1e-3, tri_begin=0, tri_end=4000)
        st = rt.last_job_stats()
        cam.props.frame_id = 0
        assert rt.render_to_camera(env, cam, S.RayProperties(bounces=3, samples=4))
        out[mode] = (irr.copy(), st, cam.download(S.OUT_COLOR).copy(), rt.last_job_stats())
    assert_bit_equal(out[0][0], out[1][0], "irradiance")
    assert_bit_equal(out[0][2], out[1][2], "colour")
    assert out[0][1] == out[1][1] and out[0][3] == out[1][3]
    cam.destroy()


# ---- state-machine edges the advisor asked for ---------------------------------------------------------------------------
def test_replacing_a_texture_needs_a_new_commit(rt, mod, S, scenes):
    """rtf_set_texture on an existing handle may move the texels: the committed material tables are stale until the next
    rtf_commit, so jobs are refused (as after rtf_remove_texture) instead of reading freed memory."""
    sd = scenes.materials_scene(seed=9)
    rt.load_scene(sd, 63)
    th = next(iter(sd.textures))
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(1024, lo, hi, seed=2)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=2)
    before = rt.estimate_illumination_mesh(env, rp, seed=3)
    h, w = sd.textures[th].shape[:2]
    rt.set_texture(th, scenes.checker_texture(4 * w, 4 * h, seed=77))  # larger: the old allocation is released
    with pytest.raises(mod.RtfError) as e:
        rt.estimate_illumination_mesh(env, rp, seed=3)
    assert e.value.code == mod.ERR_NOT_READY
    with pytest.raises(mod.RtfError):
        rt.trace_closest(rays)
    rt.build_ias(63)
    st = rt.commit_stats()
    assert st["full"] == 1 and st["materials"] == 1  # material tables re-resolved, BVH kept
    after = rt.estimate_illumination_mesh(env, rp, seed=3)
    assert after.shape == before.shape and np.isfinite(after).all()
    rt.set_texture(th, sd.textures[th])
    rt.build_ias(63)
    assert_bit_equal(rt.estimate_illumination_mesh(env, rp, seed=3), before, "texture restored")


def test_camera_accumulation_restarts_after_a_commit(rt, S, scenes):
    """The reference resets the progressive accumulation when the scene was rebuilt (RayTracer.cu:254 statusChanged from
    m_requireUpdate, raised after BuildIAS, RayTracerLayer.cpp:382-389)."""
    sd = scenes.sorghum_scene(1)
    rt.load_scene(sd, 63)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=2)
    cam = rt.create_camera()
    cam.resize(64, 36)
    cam.props.accumulate = 1
    cam.set((0.0, 1.5, 4.0), (0.9950042, -0.0998334, 0.0, 0.0))
    for k in range(3):
        assert rt.render_to_camera(env, cam, rp) and cam.props.frame_id == k + 1
    first = None
    ih = max(sd.instances)
    g, m, x, comp = sd.instances[ih]
    x2 = x.copy()
    x2[12] += 0.3
    for refit in (0, 1):  # full rebuild, then the refit path
        rt.set_option("refit", refit)
        rt.set_instance(ih, g, m, x2 if refit == 0 else x, comp)
        rt.build_ias(63)
        assert rt.render_to_camera(env, cam, rp) and cam.props.frame_id == 1, "accumulation must restart"
        assert rt.render_to_camera(env, cam, rp) and cam.props.frame_id == 2
    # material-only commit
    rt.set_material(1, S.Material(albedo=(0.9, 0.1, 0.1)))
    rt.build_ias(63)
    assert rt.commit_stats()["materials"] == 1
    assert rt.render_to_camera(env, cam, rp) and cam.props.frame_id == 1
    # a skipped commit (nothing changed) keeps accumulating
    rt.build_ias(63)
    assert rt.render_to_camera(env, cam, rp) and cam.props.frame_id == 2
    cam.destroy()


def test_scan_grid_rejects_degenerate_grids(rt, mod, S, scenes):
    sd = scenes.random_soup(100, seed=3, size=0.4)
    rt.load_scene(sd, 63)
    good = scenes.lidar_grid(center=(0.0, 5.0, 0.0), size=(1.0, 1.0), distance=(0.1, 0.1))
    assert rt.scan_grid_ray_count(good) > 0
    for size, dist in (((1.0, 1.0), (0.0, 0.1)), ((1.0, 1.0), (-0.1, 0.1)), ((1.0, 1.0), (float("nan"), 0.1)),
                       ((-1.0, 1.0), (0.1, 0.1)), ((float("inf"), 1.0), (0.1, 0.1)), ((1.0, 1.0), (1e-12, 1e-12))):
        bad = scenes.lidar_grid(center=(0.0, 5.0, 0.0), size=(1.0, 1.0), distance=(0.1, 0.1))
        bad.size[:] = size
        bad.distance[:] = dist
        with pytest.raises(mod.RtfError) as e:
            rt.scan_grid_ray_count(bad)
        assert e.value.code == mod.ERR_INVALID, (size, dist)
        with pytest.raises(mod.RtfError):
            rt.scan_grid(bad, 0, 1)
    empty = scenes.lidar_grid(center=(0.0, 5.0, 0.0), size=(0.01, 0.01), distance=(0.1, 0.1))  # fewer than one cell: no rays
    assert rt.scan_grid_ray_count(empty) == 0
    t, prim, inst = rt.scan_grid(empty)
    assert len(t) == 0 and len(prim) == 0


def test_aos_entry_points_chunk_pipeline_is_bit_identical(rt, S, scenes, mod):
    """The host-buffer entry points run as H2D / compute / D2H pipelines over record chunks; RNG streams are keyed by the
    record's index in the caller's array, so any chunk size gives the bytes of the single-chunk call."""
    sd = scenes.random_soup(900, seed=8, size=0.3)
    rt.load_scene(sd, 63)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=3, samples=8)
    probes = mod.probes_from_scene(sd)
    whole = probes.copy()
    rt.estimate_illumination(env, rp, whole, 5, 1e-3)
    stats = rt.last_job_stats()
    assert np.abs(whole["energy"]).sum() > 0
    for chunk in (1, 7, 128, 899, 900):
        rt.set_option("pipe_chunk_probes", chunk)
        got = probes.copy()
        rt.estimate_illumination(env, rp, got, 5, 1e-3)
        assert_bit_equal(got.view(np.uint8), whole.view(np.uint8), f"probe chunk {chunk}")
        assert rt.last_job_stats() == stats, chunk
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(5000, lo - 0.2, hi + 0.2, seed=4)
    smp = np.zeros(len(rays), dtype=S.PC_SAMPLE_DTYPE)
    smp["start"], smp["direction"] = rays[:, 0:3], rays[:, 4:7]
    whole_s = smp.copy()
    whole_p = rt.sample_point_cloud(env, whole_s, want_prim_ids=True)
    assert 0 < whole_s["hit"].sum() < len(smp)
    for chunk in (1, 33, 1024, 4999, 1 << 20):
        rt.set_option("pipe_chunk_samples", chunk)
        got = smp.copy()
        gp = rt.sample_point_cloud(env, got, want_prim_ids=True)
        assert_bit_equal(got.view(np.uint8), whole_s.view(np.uint8), f"sample chunk {chunk}")
        assert np.array_equal(gp, whole_p)
