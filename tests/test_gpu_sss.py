"""GPU parity of the BSSRDF branch (BSSDF.cuh:18-89, RayFunctions.cuh:97-133) through the C ABI: the spatial probe bit for
bit against the oracle (which equals brute force, tests/test_oracle_sss.py), per-triangle irradiance and camera buffers
<= 1e-5 relative, ray counts equal, and the slab closed form."""
import numpy as np
import pytest

from .test_gpu_entrypoints import close
from .test_oracle_sss import down_probe, probe_rays, slab_scene, sss_field

pytestmark = pytest.mark.gpu


def test_probe_parity(rt, orc, S, scenes):
    sd = sss_field(scenes, S)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    handles = sorted({c for (_, _, _, c) in sd.instances.values()})
    rays = probe_rays(scenes, sd, 20000, seed=33, length=4.0)
    hs = np.array([handles[i % len(handles)] for i in range(len(rays))], dtype=np.uint64)
    gpu, cpu = rt.sss_probe(rays, hs), osc.sss_probe(rays, hs)
    assert np.array_equal(gpu, cpu)
    assert gpu[:, 0].max() >= 3
    short = probe_rays(scenes, sd, 20000, seed=34, length=0.05)
    assert np.array_equal(rt.sss_probe(short, hs), osc.sss_probe(short, hs))


def test_slab_closed_form(rt, orc, S, scenes):
    sd = slab_scene(scenes, S)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    env = S.EnvironmentProperties(color=(1.0, 0.5, 0.25))
    n = 1 << 16
    rp = S.RayProperties(bounces=1, samples=n)
    pr, ref = down_probe(S), down_probe(S)
    rt.estimate_illumination(env, rp, pr, seed=77, push_normal_distance=0.0)
    st = osc.illum_probes(env, rp, ref, seed=77, push=0.0)
    close(pr["energy"], ref["energy"], "energy")
    assert rt.last_job_stats()[0] == st.rays
    expect = np.array([1.0, 0.5, 0.25]) * 17.0 / 64.0
    assert np.all(np.abs(pr["energy"][0] - expect) < 4 * 0.33 / np.sqrt(n) + 2e-4)


@pytest.mark.parametrize("brdf", ["reference", "lambert_ggx"])
def test_illumination_and_camera_parity(rt, orc, S, scenes, brdf):
    sd = sss_field(scenes, S)
    if brdf == "lambert_ggx":
        rt.set_option("brdf", S.BRDF_LAMBERT_GGX)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    if brdf == "lambert_ggx":
        osc.set_brdf(S.BRDF_LAMBERT_GGX)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=3, samples=8)
    gpu = rt.estimate_illumination_mesh(env, rp, seed=0xBEEF, push=1e-3, tri_begin=0, tri_end=3000)
    cpu, st = osc.illum_mesh(env, rp, seed=0xBEEF, push=1e-3, tri_begin=0, tri_end=3000)
    close(gpu[:, :3], cpu[:, :3], "energy")
    close(gpu[:, 3:], cpu[:, 3:], "direction", atol=1e-5)
    assert rt.last_job_stats() == (st.rays, st.paths)
    # the branch is live: switching it off changes the result
    for m in sd.materials.values():
        m.subsurface_factor = 0.0
    rt.load_scene(sd, 63)
    off = rt.estimate_illumination_mesh(env, rp, seed=0xBEEF, push=1e-3, tri_begin=0, tri_end=3000)
    assert np.abs(off[:, :3] - gpu[:, :3]).max() > 1e-3
    for m in sd.materials.values():
        m.subsurface_factor = 0.4
    rt.load_scene(sd, 63)
    # camera
    cam = rt.create_camera()
    cam.resize(96, 54)
    cam.props.accumulate, cam.props.gamma, cam.props.fov = 1, 2.2, 120.0
    cam.set((0.0, 1.5, 4.0), (0.9950042, -0.0998334, 0.0, 0.0))
    crp = S.RayProperties(bounces=3, samples=4)
    assert rt.render_to_camera(env, cam, crp)
    cp = S.CameraProps.from_buffer_copy(cam.props)
    cp.frame_id = 0
    col, nrm, alb, cst = osc.camera(env, crp, cp)
    close(cam.download(S.OUT_COLOR).reshape(-1, 4), col, "colour")
    close(cam.download(S.OUT_NORMAL).reshape(-1, 4), nrm, "normal", atol=1e-6)
    close(cam.download(S.OUT_ALBEDO).reshape(-1, 4), alb, "albedo")
    assert rt.last_job_stats() == (cst.rays, cst.paths)
    cam.destroy()


def test_extra_golden_fixtures(rt, S):
    """CUDA path vs the committed fixtures (tests/golden/oracle_extra.npz): probe records and pre-kernels bit for bit,
    irradiance <= 1e-5, flux bit for bit on the fixture's probes."""
    import os

    from tests.golden.make_golden import extra_cases

    from .conftest import GOLDEN
    g = np.load(os.path.join(GOLDEN, "oracle_extra.npz"))
    x = extra_cases()
    rt.load_scene(x["sss_scene"], 63)
    hs = np.full(len(x["probe_rays"]), x["handle"], dtype=np.uint64)
    assert np.array_equal(rt.sss_probe(x["probe_rays"], hs), g["sss_records"])
    irr = rt.estimate_illumination_mesh(x["env"], x["rp"], seed=0xABCD, push=1e-3)
    gold = g["sss_irradiance"].view(np.float32)
    close(irr[:, :3], gold[:, :3], "energy")
    close(irr[:, 3:], gold[:, 3:], "direction", atol=1e-5)
    assert list(rt.last_job_stats()) == g["sss_rays"].tolist()
    pr = np.ascontiguousarray(g["probes"]).view(S.ILLUM_SAMPLER_DTYPE)
    assert rt.illumination_flux(pr).view(np.uint32).tolist() == g["flux"].tolist()
    got = rt.prepare_light_probes()
    got["energy"] = pr["energy"]
    assert got.tobytes() == pr.tobytes()
    rt.set_geometry_instanced(50, x["verts"], x["tris"], x["mats"], x["cols"])
    bv, bt = rt.read_baked_geometry(50, len(x["verts"]) * 5, len(x["tris"]) * 5)
    assert np.array_equal(bv.view(np.uint8), g["bake_instanced"]) and np.array_equal(bt, g["bake_tris"])
    rt.set_geometry_skinned(51, x["skinned"], x["tris"], x["mats"])
    sk, _ = rt.read_baked_geometry(51, len(x["verts"]), len(x["tris"]))
    assert np.array_equal(sk.view(np.uint8), g["bake_skinned"])
