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
