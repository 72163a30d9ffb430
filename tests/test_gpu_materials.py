"""GPU vs oracle for the table-shaped material inputs, all through the C ABI: linear-memory textures (tex2D restated),
cube-map environment (RayTracerUtilities.cuh:54-80), compressed-BTF evaluation (MLVQ/*.cuh) and loader
(CompressedBTF.cpp:74-390), and the three entry points on a scene that mixes textured, vertex-colour, BTF and
alpha-textured materials (RayFunctions.cuh:5-258).

Tolerances: lookups and angle arithmetic are bit-exact (portable polynomials, explicit association); the only
tolerance-class operations are powf/expf (BTF gamma, sRGB curve, sky gamma): 1e-5 relative."""
import numpy as np
import pytest

from .conftest import assert_bit_equal

pytestmark = pytest.mark.gpu

RTOL = 1e-5


def _close(a, b, what, rtol=RTOL, atol=1e-7):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, what
    both_nan = np.isnan(a) & np.isnan(b)
    err = np.abs(a - b) - (atol + rtol * np.abs(b))
    err[both_nan] = -1
    assert np.all(err <= 0), f"{what}: max excess {err.max():.3e} at {np.unravel_index(np.argmax(err), err.shape)}"


def _rand_dirs(rng, n, upper=False):
    d = rng.normal(size=(n, 3))
    if upper:
        d[:, 1] = np.abs(d[:, 1])
    return (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)


def test_texture_sampling_bit_exact(rt, orc, scenes):
    sd = scenes.materials_scene()
    rt.load_scene(sd)
    osc = orc.OracleScene(sd)
    rng = np.random.default_rng(2)
    uv = rng.uniform(-3, 4, size=(50000, 2)).astype(np.float32)
    uv[:64] = np.array([[0, 0], [1, 1], [0.5, 0.5], [-1, 2]], np.float32).repeat(16, 0)
    for tex in (1, 2, 5, 11):
        assert_bit_equal(rt.sample_texture(tex, uv), osc.sample_texture(tex, uv), f"texture {tex}")


def test_cube_map_environment_bit_exact(rt, orc, scenes, S):
    sd = scenes.materials_scene()
    rt.load_scene(sd)
    osc = orc.OracleScene(sd)
    rng = np.random.default_rng(3)
    d = _rand_dirs(rng, 30000)
    d[:6] = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], np.float32)
    d[6:9] = np.array([[1, 1, 1], [1, -1, 1], [-1, 1, -1]], np.float32) / np.sqrt(3)  # face ties (>= rules, RTU.cuh:62-72)
    pd = np.concatenate([rng.normal(size=(len(d), 3)).astype(np.float32), d], axis=1)
    env = S.EnvironmentProperties()
    env.environmental_map_id = 3
    env.skylight_intensity = 1.7
    assert_bit_equal(rt.sample_environment(env, pd), osc.sample_environment(env, pd), "cube map")
    env.gamma = 2.2  # pow(c, 1/gamma): tolerance class
    _close(rt.sample_environment(env, pd), osc.sample_environment(env, pd), "cube map with gamma")
    env.environmental_map_id = 0
    env.gamma = 1.0
    assert_bit_equal(rt.sample_environment(env, pd), osc.sample_environment(env, pd), "constant sky")


def test_skydome_environment_against_the_reference_code(rt, scenes, S):
    """The device's environment lookup against tests/golden/env_ref.npz = outputs of the REFERENCE's own
    CalculateEnvironmentalLight / NishitaSkyIncidentLight (Environment.cuh compiled on the host,
    oracle/ref_env_driver.cpp) — no oracle in between.  The sky integrates 16 x 8 expf terms and ends in a powf:
    CUDA's and glibc's differ by ulps, so 1e-5 relative (+1e-7 of the brightest channel for the dark ones); unlit
    directions are exactly black in both; the Scene / SingleLightSource colours are bit-equal at gamma 1."""
    import os
    from .golden.make_golden import env_ref_cases, sky_env
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "env_ref.npz"))
    sky = {n[len("in_sky_"):]: g[n] for n in g.files if n.startswith("in_sky_")}
    ref = g["sky"].view(np.float32)
    rt.load_scene(scenes.materials_scene())
    n = len(ref)
    got = np.concatenate([rt.sample_environment(sky_env(sky, k), np.concatenate([sky["pos"][k], sky["dir"][k]])[None])
                          for k in range(n)])
    assert np.array_equal(got == 0, ref == 0)
    _close(got, ref, "skydome vs reference code", atol=1e-7 * float(ref.max()))
    flat = env_ref_cases()[2]
    got = np.concatenate([rt.sample_environment(S.EnvironmentProperties(**kw), np.float32([[0, 0, 0, 0, 1, 0]]))
                          for kw in flat])
    ref = g["flat"].view(np.float32)
    for k, kw in enumerate(flat):
        if kw["gamma"] == 1.0:
            assert_bit_equal(got[k], ref[k], "flat environment")
    _close(got, ref, "flat environment with gamma")


@pytest.mark.parametrize("params", [dict(), dict(use_cos_beta=False), dict(code_btf_flag=1),
                                    dict(tex_coord_multiplier=2.5, rows_offset=3, cols_offset=7, gamma=1.0),
                                    dict(hdr=True, hdr_value=1.5, multiplier=1.5)])
def test_btf_eval_matches_oracle(rt, orc, scenes, params):
    btf = scenes.synthetic_btf(seed=4, texels=(64, 48), n_pdf1d=512, **params)
    if params.get("hdr"):  # LogLuv tables: L in the linear segment, u'v' near white
        btf.t["pdf1d_basis"][:] *= 0.4
        btf.t["colors"][:, 0] = 0.19 + 0.02 * (btf.t["colors"][:, 0] - 90.0) / 70.0
        btf.t["colors"][:, 1] = 0.46 + 0.02 * (btf.t["colors"][:, 1] - 100.0) / 60.0
    sd = scenes.SceneDesc()
    sd.btfs[1] = btf
    rt.set_btf(1, btf)
    osc = orc.OracleScene(sd)
    rng = np.random.default_rng(5)
    n = 100000
    normal = _rand_dirs(rng, n)
    t0 = np.cross(normal, _rand_dirs(rng, n))
    tangent = (t0 / np.linalg.norm(t0, axis=1, keepdims=True)).astype(np.float32)
    view, illum = _rand_dirs(rng, n), _rand_dirs(rng, n)  # both hemispheres: horizon rules included
    tc = rng.uniform(-2, 3, size=(n, 2)).astype(np.float32)
    # exact special directions: along the normal, in the tangent plane, light == mirror
    view[:100], illum[:100] = -normal[:100], normal[:100]
    illum[100:200] = tangent[100:200]
    got, want = rt.btf_eval(1, tc, view, illum, normal, tangent), osc.btf_eval(1, tc, view, illum, normal, tangent)
    assert np.isfinite(want).all() and want.max() > 0
    _close(got, want, "btf value")
    if params.get("gamma") == 1.0:  # no pow anywhere on the LDR path: bit-exact
        assert_bit_equal(got, want, "btf value, gamma 1")


def test_btf_folder_loader_round_trip(rt, scenes, tmp_path):
    """CompressedBTF::ImportFromFolder format written by scenes.write_btf_folder, parsed by the C loader: evaluation
    equals the directly uploaded tables bit for bit (%.9g round-trips float32)."""
    btf = scenes.synthetic_btf(seed=6, texels=(12, 10), n_pdf1d=64, n_colors=32, n_index=32, n_pdf2d_color=32, n_pdf2d_lum=32,
                               n_pdf2d=64, n_pdf3d=32, n_pdf4d=16, use_cos_beta=True, rows_offset=2, cols_offset=5)
    scenes.write_btf_folder(str(tmp_path / "db"), btf, material="fabric01")
    rt.set_btf(1, btf)
    rt.btf_import_folder(2, tmp_path / "db")
    rng = np.random.default_rng(7)
    n = 20000
    normal = np.tile(np.array([0, 1, 0], np.float32), (n, 1))
    tangent = np.tile(np.array([1, 0, 0], np.float32), (n, 1))
    view, illum = -_rand_dirs(rng, n, upper=True), _rand_dirs(rng, n, upper=True)
    tc = rng.uniform(0, 1, size=(n, 2)).astype(np.float32)
    assert_bit_equal(rt.btf_eval(2, tc, view, illum, normal, tangent), rt.btf_eval(1, tc, view, illum, normal, tangent), "loader")
    with pytest.raises(Exception):
        rt.btf_import_folder(3, tmp_path / "missing")


def test_btf_validation_and_missing_links(rt, mod, scenes, S):
    btf = scenes.synthetic_btf(seed=4, texels=(8, 8), n_pdf1d=32)
    btf.t["pdf3d_slices"][3, 2] = btf.t["pdf2d_indices"].shape[0]  # one past the end
    with pytest.raises(mod.RtfError) as e:
        rt.set_btf(1, btf)
    assert e.value.code == mod.ERR_INVALID
    sd = scenes.materials_scene()
    del sd.material_btf[3]  # BTF material without tables
    with pytest.raises(mod.RtfError) as e:
        rt.load_scene(sd)
    assert e.value.code == mod.ERR_NOT_FOUND
    sd = scenes.materials_scene()
    sd.material_textures[1] = (99, 0, 0, 0)  # unknown texture
    rt2 = mod.RayTracer(0)
    with pytest.raises(mod.RtfError) as e:
        rt2.load_scene(sd)
    assert e.value.code == mod.ERR_NOT_FOUND
    rt2.destroy()


def _envs(S):
    sky = S.EnvironmentProperties()
    cube = S.EnvironmentProperties()
    cube.environmental_map_id = 1
    cube.skylight_intensity = 1.3
    sun = S.EnvironmentProperties()
    sun.lighting_type = S.ENV_SINGLE_LIGHT_SOURCE
    sun.sun_direction[:] = (0.3, 0.9, 0.2)
    sun.light_size = 0.05
    sun.ambient_light_intensity = 0.2
    return [("sky", sky), ("cube", cube), ("sun", sun)]


def test_illumination_with_all_material_kinds(rt, orc, scenes, S):
    sd = scenes.materials_scene()
    rt.load_scene(sd)
    osc = orc.OracleScene(sd)
    for name, env in _envs(S):
        rp = S.RayProperties(bounces=3, samples=24)
        gpu = rt.estimate_illumination_mesh(env, rp, seed=0x5EED, push=1e-3)
        cpu, st = osc.illum_mesh(env, rp, seed=0x5EED, push=1e-3)
        assert np.isfinite(cpu[:, :3]).all() and cpu[:, :3].max() > 0
        _close(gpu[:, :3], cpu[:, :3], f"irradiance ({name})", atol=1e-6)
        rays, paths = rt.last_job_stats()
        assert (rays, paths) == (st.rays, st.paths), name  # same any-hit decisions, same bounce counts


def test_camera_with_all_material_kinds(rt, orc, scenes, S):
    sd = scenes.materials_scene()
    rt.load_scene(sd)
    osc = orc.OracleScene(sd)
    for name, env in _envs(S):
        cam = rt.create_camera()
        cam.resize(48, 36)
        cam.props.accumulate, cam.props.gamma, cam.props.fov = 0, 2.2, 120.0
        cam.set((0.0, 1.7, 2.4), (0.9689124, -0.2474040, 0.0, 0.0))  # pitched down at the scene
        rp = S.RayProperties(bounces=2, samples=4)
        assert rt.render_to_camera(env, cam, rp)
        cp = S.CameraProps.from_buffer_copy(cam.props)
        cp.frame_id = 0
        col, nrm, alb, st = osc.camera(env, rp, cp)
        g = cam.download(S.OUT_COLOR).reshape(-1, 4)
        assert g[:, :3].std() > 0.02, name
        _close(g, col, f"colour ({name})", atol=1e-6)
        _close(cam.download(S.OUT_NORMAL).reshape(-1, 4), nrm, f"normal ({name})", atol=1e-6)
        _close(cam.download(S.OUT_ALBEDO).reshape(-1, 4), alb, f"albedo ({name})", atol=1e-6)
        assert rt.last_job_stats() == (st.rays, st.paths), name
        cam.destroy()


def test_lambert_ggx_mode_matches_oracle(mod, orc, scenes, S):
    """BASELINE config 2's BRDF: same sampler on both sides (sqrt / portable sincos only), so irradiance, frame buffers and
    ray counts agree like the reference model's do."""
    for sd in (scenes.materials_scene(), scenes.sorghum_scene(1)):
        sd.materials[1].roughness, sd.materials[1].metallic = 0.5, 0.3
        rt = mod.RayTracer(0)  # a context keeps its scene maps: one per scene
        rt.set_option("brdf", S.BRDF_LAMBERT_GGX)
        rt.load_scene(sd)
        osc = orc.OracleScene(sd)
        osc.set_brdf(S.BRDF_LAMBERT_GGX)
        env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=4, samples=16)
        n = min(rt.triangle_count, 600)
        gpu = rt.estimate_illumination_mesh(env, rp, seed=7, push=1e-3, tri_begin=0, tri_end=n)
        cpu, st = osc.illum_mesh(env, rp, seed=7, push=1e-3, tri_begin=0, tri_end=n)
        _close(gpu[:, :3], cpu[:, :3], "irradiance (lambert_ggx)", atol=1e-6)
        assert rt.last_job_stats() == (st.rays, st.paths)
        ref = orc.OracleScene(sd)  # the reference model gives a different answer: the option does something
        cpu_ref, _ = ref.illum_mesh(env, rp, seed=7, push=1e-3, tri_begin=0, tri_end=n)
        assert np.abs(cpu_ref[:, :3] - cpu[:, :3]).max() > 1e-3
        cam = rt.create_camera()
        cam.resize(40, 30)
        cam.props.accumulate, cam.props.gamma, cam.props.fov = 0, 2.2, 120.0
        cam.set((0.0, 1.7, 2.4), (0.9689124, -0.2474040, 0.0, 0.0))
        assert rt.render_to_camera(env, cam, S.RayProperties(bounces=4, samples=4))
        cp = S.CameraProps.from_buffer_copy(cam.props)
        cp.frame_id = 0
        col, _, _, st = osc.camera(env, S.RayProperties(bounces=4, samples=4), cp)
        _close(cam.download(S.OUT_COLOR).reshape(-1, 4), col, "colour (lambert_ggx)", atol=1e-6)
        assert rt.last_job_stats() == (st.rays, st.paths)
        cam.destroy()
        rt.destroy()


def test_subsurface_with_all_material_kinds(rt, orc, scenes, S):
    """BSSRDF branch (RayFunctions.cuh:97-133) on textured (normal-mapped records, BSSDF.cuh:31-32), vertex-colour and
    alpha-textured materials, under every lighting type, illumination and camera."""
    sd = scenes.materials_scene()
    for mh in (1, 2, 4):
        m = sd.materials[mh]
        m.subsurface_factor = 0.3 + 0.1 * mh
        m.subsurface_color[:] = (0.9, 0.5 + 0.1 * mh, 0.3)
        m.subsurface_radius[:] = (0.25, 0.0, 0.0)
    rt.load_scene(sd)
    osc = orc.OracleScene(sd)
    base = scenes.materials_scene()
    rt0_scene = orc.OracleScene(base)
    for name, env in _envs(S):
        rp = S.RayProperties(bounces=3, samples=16)
        gpu = rt.estimate_illumination_mesh(env, rp, seed=0xC0DE, push=1e-3)
        cpu, st = osc.illum_mesh(env, rp, seed=0xC0DE, push=1e-3)
        _close(gpu[:, :3], cpu[:, :3], f"irradiance ({name})", atol=1e-6)
        assert rt.last_job_stats() == (st.rays, st.paths), name
        if env.lighting_type != S.ENV_SINGLE_LIGHT_SOURCE:  # the branch is live (the single-light branch never enters it)
            off, _ = rt0_scene.illum_mesh(env, rp, seed=0xC0DE, push=1e-3)
            assert np.abs(off[:, :3] - cpu[:, :3]).max() > 1e-3, name
        cam = rt.create_camera()
        cam.resize(40, 30)
        cam.props.accumulate, cam.props.gamma, cam.props.fov = 0, 2.2, 120.0
        cam.set((0.0, 1.7, 2.4), (0.9689124, -0.2474040, 0.0, 0.0))
        crp = S.RayProperties(bounces=2, samples=4)
        assert rt.render_to_camera(env, cam, crp)
        cp = S.CameraProps.from_buffer_copy(cam.props)
        cp.frame_id = 0
        col, nrm, alb, cst = osc.camera(env, crp, cp)
        _close(cam.download(S.OUT_COLOR).reshape(-1, 4), col, f"colour ({name})", atol=1e-6)
        _close(cam.download(S.OUT_NORMAL).reshape(-1, 4), nrm, f"normal ({name})", atol=1e-6)
        assert rt.last_job_stats() == (cst.rays, cst.paths), name
        cam.destroy()
    # small batches: the waiting-bounce state is per batch
    rt.set_option("batch_paths", 2048)
    name, env = _envs(S)[0]
    rp = S.RayProperties(bounces=3, samples=16)
    small = rt.estimate_illumination_mesh(env, rp, seed=0xC0DE, push=1e-3)
    cpu, _ = osc.illum_mesh(env, rp, seed=0xC0DE, push=1e-3)
    _close(small[:, :3], cpu[:, :3], "irradiance (2048-path batches)", atol=1e-6)


def test_subsurface_switched_on_by_a_material_only_commit_and_ray_sort(rt, orc, scenes, S):
    """has_sss flips in a material-only commit (BVH kept); the ray-sort option leaves the BSSRDF + any-hit result alone."""
    sd = scenes.materials_scene()
    rt.load_scene(sd)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=3, samples=8)
    plain = rt.estimate_illumination_mesh(env, rp, seed=11, push=1e-3)
    m = sd.materials[1]
    m.subsurface_factor, m.subsurface_radius[0] = 0.6, 0.3
    rt.set_material(1, m)
    rt.build_ias(63)
    st = rt.commit_stats()
    assert st["materials"] == 1 and st["full"] == 1
    osc = orc.OracleScene(sd)
    cpu, cst = osc.illum_mesh(env, rp, seed=11, push=1e-3)
    on = rt.estimate_illumination_mesh(env, rp, seed=11, push=1e-3)
    _close(on[:, :3], cpu[:, :3], "irradiance after the material-only commit", atol=1e-6)
    assert rt.last_job_stats() == (cst.rays, cst.paths)
    assert np.abs(on[:, :3] - plain[:, :3]).max() > 1e-3
    rt.set_option("sort_rays", 1)
    srt = rt.estimate_illumination_mesh(env, rp, seed=11, push=1e-3)
    assert np.array_equal(srt.view(np.uint32), on.view(np.uint32)) and rt.last_job_stats() == (cst.rays, cst.paths)
