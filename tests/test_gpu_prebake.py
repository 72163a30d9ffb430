"""GPU parity (through the C ABI) of the geometry pre-kernels, light-probe construction and flux integration against
the oracle: bit-exact (glm operation order, -fmad=false vs -ffp-contract=off)."""
import numpy as np
import pytest
import torch

from .conftest import assert_bit_equal
from .test_oracle_prebake import _trs, small_mesh

pytestmark = pytest.mark.gpu


def test_instanced_geometry_bake_and_trace(rt, orc, S, scenes):
    sd1 = scenes.sorghum_scene(1, ground=False)
    v, t = sd1.geometries[1]
    v = v.copy()
    v["tangent"] = np.roll(v["normal"], 1, axis=1) + np.float32(0.25)
    n_inst = 7
    mats = np.stack([_trs(0.9 * i, 0.8 + 0.07 * i, (0.8 * (i % 3), 0.0, 0.9 * (i // 3))) for i in range(n_inst)])
    cols = scenes.hash_uniform(3, np.arange(n_inst * 4), 1).reshape(-1, 4).astype(np.float32)
    rt.set_geometry_instanced(5, v, t, mats, cols)
    gv, gt = rt.read_baked_geometry(5, len(v) * n_inst, len(t) * n_inst)
    ov, ot = orc.bake_instanced(v, t, mats, cols)
    assert gv.tobytes() == ov.tobytes(), "baked vertices differ from the oracle"
    assert np.array_equal(gt, ot)
    # the baked geometry traces like any other: scene = baked mesh under the identity
    rt.set_material(1, S.Material(albedo=(0.5, 0.6, 0.4)))
    rt.set_instance(1, 5, 1, scenes.IDENTITY.copy(), 77)
    rt.build_ias(63)
    assert rt.triangle_count == len(t) * n_inst
    sd = scenes.SceneDesc()
    sd.geometries[5] = (ov, ot)
    sd.materials[1] = S.Material(albedo=(0.5, 0.6, 0.4))
    sd.instances[1] = (5, 1, scenes.IDENTITY.copy(), 77)
    osc = orc.OracleScene(sd, 63)
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(1 << 15, lo, hi, seed=8)
    assert_bit_equal(rt.trace_closest(rays).view(np.uint32), osc.trace(rays).view(np.uint32), "hits")
    # LiDAR hit records carry the instance colour in HitInfo::m_data (RDD.hpp:150-159)
    smp = np.zeros(4096, dtype=S.PC_SAMPLE_DTYPE)
    smp["start"] = rays[:4096, 0:3]
    smp["direction"] = rays[:4096, 4:7]
    ref = smp.copy()
    rt.sample_point_cloud(S.EnvironmentProperties(), smp)
    osc.scan_point_cloud(ref)
    assert smp.tobytes() == ref.tobytes()
    hit = smp["hit"] != 0
    assert hit.sum() > 100
    assert np.isin(smp["hit_info"]["data"][hit][:, 3], cols[:, 3]).all()
    # re-baking an existing handle invalidates the committed scene until the next commit
    rt.set_geometry_instanced(5, v, t, mats[:2], None)
    with pytest.raises(Exception):
        rt.trace_closest(rays[:16])
    rt.build_ias(63)
    assert rt.triangle_count == len(t) * 2


def test_skinned_geometry_bake(rt, orc, S, scenes):
    v, t = small_mesh(scenes)
    sv = np.zeros(len(v), dtype=S.SKINNED_VERTEX_DTYPE)
    sv["v"] = v
    idx = np.arange(len(v))
    n_bones = 6
    sv["bone_id"] = -1
    sv["bone_id2"] = -1
    sv["bone_id"][:, 0] = idx % n_bones
    sv["bone_id"][:, 1] = np.where(idx % 2 == 0, (idx + 1) % n_bones, -1)
    sv["bone_id"][:, 3] = np.where(idx % 3 == 0, (idx + 2) % n_bones, -1)
    sv["bone_id2"][:, 2] = np.where(idx % 5 == 0, (idx + 3) % n_bones, -1)
    w = scenes.hash_uniform(9, np.arange(len(v) * 8), 2).reshape(-1, 8).astype(np.float32)
    sv["weight"], sv["weight2"] = w[:, :4], w[:, 4:]
    bones = np.stack([_trs(0.5 * i, 1.0 + 0.1 * i, (0.1 * i, 0.2, -0.1 * i)) for i in range(n_bones)])
    rt.set_geometry_skinned(3, sv, t, bones)
    gv, gt = rt.read_baked_geometry(3, len(v), len(t))
    assert gv.tobytes() == orc.bake_skinned(sv, bones).tobytes()
    assert np.array_equal(gt, t)
    sv["bone_id"][5, 0] = 99
    with pytest.raises(Exception, match="bone index"):
        rt.set_geometry_skinned(3, sv, t, bones)


def test_pre_kernels_equal_the_reference_kernels(rt):
    """T20 on the GPU: k_bake_instanced / k_bake_skinned against tests/golden/prebake_ref.npz, the output of the reference's
    own CopyVerticesInstancedKernel / CopySkinnedVerticesKernel (compiled unmodified, oracle/Makefile) — every byte."""
    import os

    from tests.golden import make_golden as G
    g = np.load(os.path.join(G.HERE, "prebake_ref.npz"))
    c = G.prebake_ref_cases()
    v, t = c["vertices"], c["triangles"]
    ni = len(c["matrices"])
    rt.set_geometry_instanced(5, v, t, c["matrices"], c["colors"])
    gv, gt = rt.read_baked_geometry(5, len(v) * ni, len(t) * ni)
    assert gv.view(np.uint8).tobytes() == g["instanced_vertices"].tobytes(), "instanced bake differs from the reference kernel"
    assert np.array_equal(gt, np.concatenate([t + k * len(v) for k in range(ni)]))
    rt.set_geometry_skinned(6, c["skinned"], t, c["bones"])
    sv, st = rt.read_baked_geometry(6, len(v), len(t))
    assert sv.view(np.uint8).tobytes() == g["skinned_vertices"].tobytes(), "skinned bake differs from the reference kernel"
    assert np.array_equal(st, t)


def test_light_probes_and_flux(rt, orc, mod, S, scenes):
    sd = scenes.sorghum_scene(2)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    n = rt.triangle_count
    pr = rt.prepare_light_probes(front=True, back=True)
    assert pr.tobytes() == osc.make_probes().tobytes()
    part = rt.prepare_light_probes(1000, 3000, front=False, back=True)
    assert part.tobytes() == osc.make_probes(1000, 3000, False, True).tobytes()
    # flux over the AoS entry point's output
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=8)
    rt.estimate_illumination(env, rp, pr, seed=7)
    flux = rt.illumination_flux(pr)
    assert_bit_equal(flux, orc.flux_probes(pr), "probe flux")
    area = mod.probe_area(pr).astype(np.float64)
    exact = (pr["energy"].astype(np.float64) * area[:, None]).sum(0)
    assert np.allclose(flux[:3], exact, rtol=1e-6) and np.isclose(flux[3], 2 * area.sum(), rtol=1e-6)
    assert np.allclose(flux[4:7], flux[:3] / flux[3], rtol=1e-6)
    # mesh form on a device array: same probes -> same bits
    e6 = np.zeros((n, 6), dtype=np.float32)
    e6[:, :3] = pr["energy"]
    d = torch.from_numpy(e6).cuda()
    assert_bit_equal(rt.illumination_flux_mesh_device(d.data_ptr(), 0, n), flux, "mesh flux")
    sub = rt.illumination_flux_mesh_device(d.data_ptr(), 1234, 9999, front=True, back=False)
    assert_bit_equal(sub, osc.flux_mesh(e6, 1234, 9999, True, False), "mesh flux, sub-range")
    assert rt.illumination_flux(pr[:0])[3] == 0


def test_prebake_edge_cases(rt, mod, S, scenes):
    v, t = small_mesh(scenes)
    # not committed yet: probes need an acceleration structure (reference: silent return, RayTracer.cu:610-611)
    with pytest.raises(mod.RtfError) as ei:
        rt.prepare_light_probes(0, 0)
    assert ei.value.code == mod.ERR_NOT_READY
    # zero instances -> a geometry with no triangles: its scene instance contributes nothing
    rt.set_geometry_instanced(9, v, t, np.zeros((0, 16), dtype=np.float32))
    rt.set_geometry(1, v, t)
    rt.set_material(1, S.Material())
    rt.set_instance(1, 9, 1, scenes.IDENTITY.copy(), 5)
    rt.set_instance(2, 1, 1, scenes.IDENTITY.copy(), 6)
    rt.build_ias(63)
    assert rt.triangle_count == len(t)
    assert len(rt.prepare_light_probes(3, 3)) == 0
    with pytest.raises(mod.RtfError):
        rt.prepare_light_probes(0, len(t) + 1)
    # bad index, bad stride
    bad = t.copy()
    bad[0, 0] = len(v)
    with pytest.raises(mod.RtfError, match="out of range"):
        rt.set_geometry_instanced(9, v, bad, scenes.IDENTITY.copy().reshape(1, 16))
    # single-term flux = energy * area (two faces count the area twice)
    pr = rt.prepare_light_probes(0, 1)
    pr["energy"] = (2.0, 0.0, 1.0)
    out = rt.illumination_flux(pr)
    a = mod.probe_area(pr)[0]
    assert out[3] == np.float32(2) * a and out[0] == np.float32(2) * a and out[4] == np.float32(1.0)
