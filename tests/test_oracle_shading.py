"""Closed-form pins of the oracle's restatement of IlluminationEstimation.cu / PointCloudScanning.cu /
CameraRendering.cu / RayFunctions.cuh (CPU only)."""
import math

import numpy as np

from .conftest import assert_bit_equal


def _single_triangle(scenes, S, y=0.0, scale=1.0, material=None):
    sd = scenes.SceneDesc()
    pos = np.array([[-1, y, -1], [1, y, -1], [0, y, 1]], dtype=np.float64) * [scale, 1, scale]
    nrm = np.tile([0.0, 1.0, 0.0], (3, 1))
    sd.geometries[1] = (scenes.make_vertices(pos, nrm), np.array([[0, 2, 1]], dtype=np.uint32))
    sd.materials[1] = material or S.Material(albedo=(0.5, 0.5, 0.5))
    sd.instances[1] = (1, 1, scenes.IDENTITY.copy(), 77)
    return sd


def test_single_triangle_under_constant_sky(orc, S, scenes):
    # E[energy] = sky * E[cos] = sky/2 for the unoccluded front face: hemisphere uniform in cos(theta), weight |n.d|,
    # no pdf division (IlluminationEstimation.cu:61,81-82; RayTracerUtilities.cuh:138).  The back face looks down
    # (nothing below) and gets the same.
    sd = _single_triangle(scenes, S)
    osc = orc.OracleScene(sd)
    env = S.EnvironmentProperties(color=(1.0, 0.5, 0.25), skylight_intensity=2.0)
    rp = S.RayProperties(bounces=4, samples=4096)
    out, st = osc.illum_mesh(env, rp, seed=1, push=1e-3, front=True, back=False)
    sigma = 2.0 * math.sqrt(1 / 12) / math.sqrt(4096)
    assert abs(out[0, 0] - 1.0) < 4 * sigma
    assert np.allclose(out[0, :3] / out[0, 0], [1.0, 0.5, 0.25], rtol=1e-5)
    assert st.paths == 4096 and st.rays == 4096  # every ray misses: one ray per path
    d = out[0, 3:]
    assert abs(np.linalg.norm(d) - 1) < 1e-5 and d[1] > 0.95  # mean direction = +y


def test_occluded_probe_gets_nothing(orc, S, scenes):
    # a black emission-free lid right above the probe: every front-face path hits it and dies dark
    sd = _single_triangle(scenes, S, y=0.0, scale=0.2)
    lid = _single_triangle(scenes, S, y=0.05, scale=50.0, material=S.Material(albedo=(0.0, 0.0, 0.0)))
    sd.geometries[2], sd.materials[2] = lid.geometries[1], lid.materials[1]
    sd.instances[2] = (2, 2, scenes.IDENTITY.copy(), 78)
    osc = orc.OracleScene(sd)
    out, _ = osc.illum_mesh(S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=256), seed=3, push=1e-3,
                            front=True, back=False, tri_begin=0, tri_end=1)
    assert np.all(out[0, :3] == 0.0)


def test_emissive_lid_closed_form(orc, S, scenes):
    # lid with emission e and albedo a, bounces = 0: radiance = e*a for every hit (RayFunctions.cuh:168-171),
    # probe energy -> e*a*E[cos] = e*a/2
    sd = _single_triangle(scenes, S, y=0.0, scale=0.2)
    lid = _single_triangle(scenes, S, y=0.05, scale=50.0, material=S.Material(albedo=(0.5, 1.0, 0.25), emission=3.0))
    sd.geometries[2], sd.materials[2] = lid.geometries[1], lid.materials[1]
    sd.instances[2] = (2, 2, scenes.IDENTITY.copy(), 78)
    osc = orc.OracleScene(sd)
    out, _ = osc.illum_mesh(S.EnvironmentProperties(), S.RayProperties(bounces=0, samples=4096), seed=3, push=1e-3,
                            front=True, back=False, tri_begin=0, tri_end=1)
    assert abs(out[0, 1] - 1.5) < 0.06 and np.allclose(out[0, :3] / out[0, 1], [0.5, 1.0, 0.25], rtol=1e-5)


def test_sample_sharding_sums_to_the_whole(orc, S, scenes):
    # partial sums over sample ranges add up to the full job: what cross-GPU sample sharding relies on
    sd = scenes.random_soup(40, seed=2, size=0.4)
    osc = orc.OracleScene(sd)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=64)
    whole, _ = osc.illum_mesh(env, rp, 9, 1e-3, normalise=False)
    a, _ = osc.illum_mesh(env, rp, 9, 1e-3, s0=0, s1=32, normalise=False)
    b, _ = osc.illum_mesh(env, rp, 9, 1e-3, s0=32, s1=64, normalise=False)
    assert np.allclose(a + b, whole, rtol=1e-5, atol=5e-5)
    # triangle-range sharding is bit-identical
    c, _ = osc.illum_mesh(env, rp, 9, 1e-3, tri_begin=10, tri_end=25, normalise=False)
    assert_bit_equal(c, whole[10:25], "triangle shard")


def test_probe_api_equals_mesh_api(orc, S, scenes, mod):
    sd = scenes.random_soup(50, seed=6, size=0.3)
    osc = orc.OracleScene(sd)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=8)
    mesh, _ = osc.illum_mesh(env, rp, 5, 1e-3)
    probes = mod.probes_from_scene(sd)
    osc.illum_probes(env, rp, probes, 5, 1e-3)
    assert_bit_equal(probes["energy"], mesh[:, :3], "energy")
    assert_bit_equal(probes["direction"], mesh[:, 3:], "direction")
    # no enabled face: outputs untouched (IlluminationEstimation.cu:125)
    probes["front_face"], probes["back_face"] = 0, 0
    probes["energy"] = 7.0
    osc.illum_probes(env, rp, probes, 5, 1e-3)
    assert np.all(probes["energy"] == 7.0)


def test_lidar_plane_exact(orc, S, scenes):
    sd = scenes.SceneDesc()
    gv, gt = scenes.ground_plane(4.0, 4.0, y=0.25)
    sd.geometries[1], sd.materials[1] = (gv, gt), S.Material()
    sd.instances[3] = (1, 1, scenes.IDENTITY.copy(), 4242)
    osc = orc.OracleScene(sd)
    grid = scenes.lidar_grid(center=(0.0, 2.25, 0.0), size=(2.0, 1.0), distance=(0.125, 0.125))
    t, prim, inst, st = osc.scan_grid(grid)
    assert len(t) == 16 * 8 and st.rays == 128
    assert np.all(t == 2.0) and np.all(prim < 2) and np.all(inst == 0)
    # AoS entry point: handle = instance component handle, position = hit point, normal faces the ray
    rays = orc.scan_grid_rays(grid, 0, 128)
    smp = np.zeros(130, dtype=S.PC_SAMPLE_DTYPE)
    smp["start"][:128], smp["direction"][:128] = rays[:, 0:3], rays[:, 4:7]
    smp["start"][128:], smp["direction"][128:] = (0, 1, 0), (0, 1, 0)  # two misses
    p = osc.scan_point_cloud(smp)
    assert np.all(smp["hit"][:128] == 1) and np.all(smp["handle"][:128] == 4242)
    assert np.allclose(smp["hit_info"]["position"][:128, 1], 0.25) and np.all(smp["hit_info"]["normal"][:128, 1] == 1.0)
    assert np.all(smp["hit"][128:] == 0) and np.all(smp["handle"][128:] == 0) and np.all(p[128:] == 0xFFFFFFFF)


def test_camera_sky_only_and_gamma(orc, S, scenes):
    from tests.golden.make_golden import camera_props
    sd = _single_triangle(scenes, S, y=-50.0, scale=0.01)  # nothing in view
    osc = orc.OracleScene(sd)
    cp = camera_props(8, 6)
    env = S.EnvironmentProperties(color=(0.25, 0.5, 1.0))
    col, nrm, alb, st = osc.camera(env, S.RayProperties(bounces=1, samples=4), cp)
    expect = np.power(np.array([0.25, 0.5, 1.0]), 1 / 2.2)
    assert np.allclose(col[:, :3], expect, rtol=1e-5) and np.all(col[:, 3] == 1.0)
    assert np.allclose(alb[:, :3], [0.25, 0.5, 1.0]) and np.all(nrm[:, :3] == 0)  # MissFunc: albedo = energy = env
    assert st.rays == 8 * 6 * 4


def test_single_light_source_shadow_ray(orc, S, scenes):
    # RayFunctions.cuh:61-92: ambient + one shadow ray to the sun; emission added once; no bounce
    sd = _single_triangle(scenes, S, y=0.0, scale=3.0, material=S.Material(albedo=(0.5, 0.5, 0.5)))
    osc = orc.OracleScene(sd)
    from tests.golden.make_golden import camera_props
    cp = camera_props(4, 4)
    cp.from_[:] = (0.0, 2.0, 0.0)
    env = S.EnvironmentProperties(lighting_type=S.ENV_SINGLE_LIGHT_SOURCE, ambient_light_intensity=0.1, light_size=0.0,
                                  sun_direction=(0.0, 1.0, 0.0), color=(1.0, 1.0, 1.0), gamma=1.0)
    # probe above the triangle looking straight down: use illumination API with a probe whose rays all hit
    probes = np.zeros(1, dtype=S.ILLUM_SAMPLER_DTYPE)
    for k, p in zip("abc", [(-0.1, 1.0, -0.1), (0.1, 1.0, -0.1), (0.0, 1.0, 0.1)]):
        probes[k]["position"], probes[k]["normal"] = p, (0.0, -1.0, 0.0)
    probes["front_face"], probes["back_face"] = 1, 0
    osc.illum_probes(env, S.RayProperties(bounces=4, samples=1024), probes, 2, 1e-3)
    # hit radiance = ambient*albedo + L_sun*(n.l = 1)*albedo = 0.05 + 0.5 for rays that hit (cos-uniform hemisphere,
    # triangle subtends most of it); misses see the sun colour 1.0.  Energy within (0.55*0.5*frac, ...] -> loose bounds
    e = probes["energy"][0, 0]
    assert 0.2 < e < 0.5


def test_lambert_ggx_closed_forms(orc):
    """The Lambert+GGX sampler (option brdf = 1; not in the reference, BASELINE config 2 names it): closed forms of its
    two lobes."""
    n, v = np.array([0, 1, 0], np.float32), np.array([0.6, 0.8, 0.0], np.float32)
    rng = np.random.default_rng(3)
    r = rng.random((200000, 3)).astype(np.float32)
    alb = np.array([0.8, 0.5, 0.3], np.float32)
    # dielectric: half the samples go to each lobe; the diffuse weight is the constant albedo * 0.96 / (1 - p_spec)
    w, d = orc.sample_lambert_ggx(alb, 0.0, 0.6, n, v, r)
    diff = r[:, 2] >= 0.5
    assert np.allclose(w[diff], alb * np.float32(1.92), rtol=1e-6)
    assert np.all(d[diff][:, 1] >= 0) and abs(d[diff][:, 1].mean() - 2.0 / 3.0) < 5e-3  # cosine-weighted: E[cos] = 2/3
    assert np.allclose(np.linalg.norm(d[diff], axis=1), 1.0, atol=1e-5)
    spec = ~diff
    assert np.all(w[spec] >= 0) and w[spec].mean() < 0.3  # F0 = 0.04: a dielectric's specular albedo is small
    # perfect conductor, white: the visible-normal sampler's weight is F * G1(l) <= 1 — no sample ever creates energy —
    # and single-scatter GGX keeps most of the energy at low roughness
    for rough, lo in ((0.15, 0.93), (0.5, 0.75), (1.0, 0.3)):
        w, d = orc.sample_lambert_ggx(np.ones(3, np.float32), 1.0, rough, n, v, r)
        assert w.max() <= 1.0 + 1e-6 and w.min() >= 0.0
        mean = w[:, 0].mean()
        assert lo < mean <= 1.0, (rough, mean)
        assert np.all(d[w[:, 0] > 0][:, 1] > 0)  # reflected directions stay above the surface
    # roughness -> 0: the sampled direction is the mirror direction
    w, d = orc.sample_lambert_ggx(np.ones(3, np.float32), 1.0, 0.0, n, v, r[:1000])
    mirror = 2 * (v @ n) * n - v
    assert np.percentile(np.abs(d - mirror).max(1), 95) < 1e-2 and np.allclose(w, 1.0, atol=1e-2)
    # grazing view below the surface is absorbed, not reflected
    w, _ = orc.sample_lambert_ggx(np.ones(3, np.float32), 1.0, 0.5, n, np.array([0.8, -0.6, 0.0], np.float32), r[:100])
    assert np.all(w == 0)
