"""CPU: the oracle's restatement of the geometry pre-kernels (CopyVerticesInstancedKernel RayTracer.cu:1148-1175,
CopySkinnedVerticesKernel :1194-1247), of PrepareLightProbeGroup and of the flux integration
(TriangleIlluminationEstimator.cpp:84-139) against closed forms and float64 numpy."""
import numpy as np

from .conftest import assert_bit_equal


def _trs(yaw, s, t):
    c, sn = np.cos(yaw), np.sin(yaw)
    m = np.array([[c * s, 0, sn * s, t[0]], [0, s, 0, t[1]], [-sn * s, 0, c * s, t[2]], [0, 0, 0, 1]], dtype=np.float64)
    return m.T.astype(np.float32).reshape(16)  # column-major


def small_mesh(scenes):
    sd = scenes.random_soup(40, seed=11, size=0.4)
    v, t = sd.geometries[1]
    v = v.copy()
    v["tangent"] = np.roll(v["normal"], 1, axis=1) + np.float32(0.25)
    v["tex_coord"] = v["position"][:, :2]
    v["color"][:, 0] = np.linspace(0, 1, len(v), dtype=np.float32)
    return v, t


def test_bake_instanced_closed_forms(orc, scenes):
    v, t = small_mesh(scenes)
    mats = np.stack([_trs(0.0, 1.0, (0, 0, 0)), _trs(0.7, 2.0, (1, 2, 3)), _trs(-1.9, 0.5, (-4, 0, 1))])
    cols = np.array([[1, 0, 0, 1], [0, 1, 0, 0.5], [0.2, 0.3, 0.4, 0.9]], dtype=np.float32)
    ov, ot = orc.bake_instanced(v, t, mats, cols)
    nv, nt = len(v), len(t)
    assert len(ov) == 3 * nv and len(ot) == 3 * nt
    # identity instance: position unchanged, normal/tangent orthonormalised
    assert np.allclose(ov["position"][:nv], v["position"], atol=0, rtol=0)
    for i in range(3):
        sl = slice(i * nv, (i + 1) * nv)
        m = mats[i].reshape(4, 4).T.astype(np.float64)
        p = v["position"].astype(np.float64) @ m[:3, :3].T + m[:3, 3]
        assert np.allclose(ov["position"][sl], p, rtol=1e-6, atol=1e-6)
        n = v["normal"].astype(np.float64) @ m[:3, :3].T
        n /= np.linalg.norm(n, axis=1, keepdims=True)
        assert np.allclose(ov["normal"][sl], n, atol=2e-6)
        tg = ov["tangent"][sl].astype(np.float64)
        assert np.allclose(np.linalg.norm(tg, axis=1), 1, atol=1e-5)
        assert np.abs(np.sum(tg * ov["normal"][sl], axis=1)).max() < 1e-5  # Gram-Schmidt
        assert np.all(ov["position_padding"][sl] == cols[i, 0]) and np.all(ov["normal_padding"][sl] == cols[i, 1])
        assert np.all(ov["tangent_padding"][sl] == cols[i, 2])
        assert np.all(ov["tex_coord_padding"][sl, 0] == cols[i, 3]) and np.all(ov["tex_coord_padding"][sl, 1] == i)
        assert np.array_equal(ov["tex_coord"][sl], v["tex_coord"]) and np.array_equal(ov["color"][sl], v["color"])
        assert np.array_equal(ot[i * nt:(i + 1) * nt], t + i * nv)
    ov2, _ = orc.bake_instanced(v, t, mats, None)
    assert np.all(ov2["position_padding"] == 1) and np.all(ov2["tex_coord_padding"][:, 0] == 1)


def test_bake_skinned_closed_forms(orc, S, scenes):
    v, t = small_mesh(scenes)
    sv = np.zeros(len(v), dtype=S.SKINNED_VERTEX_DTYPE)
    sv["v"] = v
    sv["bone_id"][:] = -1
    sv["bone_id2"][:] = -1
    bones = np.stack([_trs(0.0, 1.0, (0, 0, 0)), _trs(0.0, 1.0, (2, 0, 0)), _trs(1.1, 1.5, (0, 1, 0))])
    # one bone, weight 1 -> that bone's transform
    sv["bone_id"][:, 0] = 2
    sv["weight"][:, 0] = 1.0
    ov = orc.bake_skinned(sv, bones)
    ref, _ = orc.bake_instanced(v, t, bones[2:3], None)
    assert_bit_equal(ov["position"], ref["position"], "single bone position")
    assert_bit_equal(ov["normal"], ref["normal"], "single bone normal")
    assert np.array_equal(ov["position_padding"], v["position_padding"])
    # two translations blended half/half -> translation by (1,0,0); influences in the second quadruple count too
    sv["bone_id"][:, 0] = 0
    sv["weight"][:, 0] = 0.5
    sv["bone_id2"][:, 3] = 1
    sv["weight2"][:, 3] = 0.5
    ov = orc.bake_skinned(sv, bones)
    assert np.allclose(ov["position"], v["position"] + np.array([1, 0, 0], dtype=np.float32), atol=1e-6)
    sv["bone_id"][3, 1] = 7
    try:
        orc.bake_skinned(sv, bones)
        assert False, "bone index out of range accepted"
    except ValueError:
        pass


def test_probes_and_flux(orc, mod, S, scenes):
    sd = scenes.sorghum_scene(1)
    osc = orc.OracleScene(sd, 63)
    pr = osc.make_probes(front=True, back=False)
    host = mod.probes_from_scene(sd, front=True, back=False)  # numpy restatement of PrepareLightProbeGroup
    assert pr.tobytes() == host.tobytes()
    # flux: energy = (1, 2, 0.5) everywhere -> totalFlux = energy * sum(area); averageFlux = energy (one-sided)
    pr["energy"] = np.array([1.0, 2.0, 0.5], dtype=np.float32)
    out = orc.flux_probes(pr)
    area = mod.probe_area(pr).astype(np.float64)
    assert np.isclose(out[3], area.sum(), rtol=1e-6)
    assert np.allclose(out[:3], np.array([1.0, 2.0, 0.5]) * area.sum(), rtol=1e-6)
    assert np.allclose(out[4:7], [1.0, 2.0, 0.5], rtol=1e-6)
    # two-sided probes count their area twice (TriangleIlluminationEstimator.cpp:124)
    pr["back_face"] = 1
    out2 = orc.flux_probes(pr)
    assert np.isclose(out2[3], 2 * area.sum(), rtol=1e-6) and np.allclose(out2[4:7], [0.5, 1.0, 0.25], rtol=1e-6)
    # random energies vs float64, and vs the reference's serial float32 loop (TriangleIlluminationEstimator.cpp:87-90)
    e = (scenes.hash_uniform(5, np.arange(len(pr) * 3), 0).reshape(-1, 3)).astype(np.float32)
    pr["energy"] = e
    out3 = orc.flux_probes(pr)
    a32 = mod.probe_area(pr)
    exact = (e.astype(np.float64) * a32[:, None].astype(np.float64)).sum(0)
    assert np.allclose(out3[:3], exact, rtol=2e-7)
    serial = np.zeros(3, dtype=np.float32)
    for k in range(len(pr)):
        serial += e[k] * a32[k]
    assert np.allclose(out3[:3], serial, rtol=1e-4)
    # mesh form = probe form
    e6 = np.zeros((osc.n_tris, 6), dtype=np.float32)
    e6[:, :3] = e
    assert_bit_equal(osc.flux_mesh(e6, front=True, back=True), out3, "mesh flux")
    # empty and single-tile inputs
    assert orc.flux_probes(pr[:0])[3] == 0
    one = orc.flux_probes(pr[:1])
    assert one[3] == np.float32(2) * a32[0]
