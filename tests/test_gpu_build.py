"""GPU LBVH builder vs the CPU oracle, stage by stage, bit for bit (through the C ABI's parity read-back).
Replaces optixAccelBuild (RayTracer.cu:1528-1584, 1654-1707) — closed source, so the oracle is the pin."""
import numpy as np
import pytest

from .conftest import assert_bit_equal

pytestmark = pytest.mark.gpu


def _compare_all(rt, osc, S):
    assert rt.triangle_count == osc.n_tris
    for what, name in [(S.DBG_FLAT_POS, "flat pos"), (S.DBG_FLAT_NRM, "flat nrm"), (S.DBG_FLAT_INST, "flat inst"),
                       (S.DBG_SCENE_BOUNDS, "scene bounds"), (S.DBG_REF_TRI, "reference triangles"),
                       (S.DBG_REF_BOX, "reference boxes"), (S.DBG_MORTON, "morton"), (S.DBG_SORTED_IDS, "sorted ids"),
                       (S.DBG_BIN_CHILDREN, "children"), (S.DBG_BIN_BOXES, "boxes"),
                       (S.DBG_WIDE_NODES, "wide nodes"), (S.DBG_WIDE_TRIS, "wide tris")]:
        assert_bit_equal(rt.debug(what), osc.debug(what), name)


@pytest.mark.parametrize("n,bits", [(1, 63), (2, 63), (3, 30), (4, 63), (5, 30), (9, 63), (100, 30), (257, 63),
                                    (5000, 63), (5000, 30), (100000, 63)])
@pytest.mark.parametrize("presplit", [1, 8])
def test_soup_build_bit_exact(rt, orc, S, scenes, n, bits, presplit):
    sd = scenes.random_soup(n, seed=n, size=0.08 if n < 50000 else 0.02)
    rt.load_scene(sd, bits, presplit=presplit)
    _compare_all(rt, orc.OracleScene(sd, bits, presplit=presplit), S)


def test_duplicates_and_degenerates_bit_exact(rt, orc, S, scenes):
    sd = scenes.random_soup(64, seed=5, size=0.6)
    v, t = sd.geometries[1]
    v = np.concatenate([v, v, v[:3]])
    t = np.concatenate([t, t + 192, [[384, 384, 384]], [[384, 385, 385]]]).astype(np.uint32)
    sd.geometries[1] = (v, t)
    for bits in (30, 63):
        rt.load_scene(sd, bits)
        _compare_all(rt, orc.OracleScene(sd, bits), S)


@pytest.mark.parametrize("bits,presplit", [(30, None), (63, None), (63, 1), (63, 3), (63, 32)])
def test_sorghum_c1_build_bit_exact(rt, orc, S, scenes, bits, presplit):
    """Sliver-heavy scene: the pre-split references (count, order, clipped boxes) and everything built on them."""
    sd = scenes.sorghum_scene(1)
    rt.load_scene(sd, bits, presplit=presplit)
    assert rt.triangle_count == 20002
    osc = orc.OracleScene(sd, bits, presplit=S.DEFAULT_PRESPLIT if presplit is None else presplit)
    n_ref = len(rt.debug(S.DBG_REF_TRI))
    assert n_ref == 20002 if presplit == 1 else n_ref > 2 * 20002
    _compare_all(rt, osc, S)
    tm = rt.build_timings()
    assert tm["total"] > 0


def test_instanced_field_build_bit_exact(rt, orc, S, scenes):
    sd = scenes.sorghum_scene(6)  # 6 transformed instances of one geometry + ground = 120 002 triangles
    rt.load_scene(sd, 63)
    assert rt.triangle_count == 120002
    _compare_all(rt, orc.OracleScene(sd, 63), S)


def test_rebuild_after_scene_edit(rt, orc, S, scenes):
    sd = scenes.sorghum_scene(3)
    rt.load_scene(sd, 63)
    rt.remove_instance(11)
    del sd.instances[11]
    rt.build_ias(63)
    _compare_all(rt, orc.OracleScene(sd, 63), S)
    with pytest.raises(Exception):
        rt.remove_instance(11)


def test_not_ready_and_bad_arguments(rt, mod, S, scenes):
    env, rp = S.EnvironmentProperties(), S.RayProperties()
    rays = np.zeros((4, 8), dtype=np.float32)
    with pytest.raises(mod.RtfError) as e:  # reference: silent return when there is no AS (RayTracer.cu:610-611)
        rt.trace_closest(rays)
    assert e.value.code == mod.ERR_NOT_READY
    cam = rt.create_camera()
    cam.resize(8, 8)
    cam.set((0, 0, 3))
    assert rt.render_to_camera(env, cam, rp) is False  # RayTracer.cu:248-249
    cam.destroy()
    with pytest.raises(mod.RtfError):
        rt.build_ias(17)
    sd = scenes.random_soup(4, seed=1)
    v, t = sd.geometries[1]
    with pytest.raises(mod.RtfError):
        rt.set_geometry(1, v, t + 100)  # index out of range
    rt.set_geometry(1, v, t)
    rt.set_material(1, sd.materials[1])
    rt.set_instance(1, 1, 99, scenes.IDENTITY)  # unknown material
    with pytest.raises(mod.RtfError) as e:
        rt.build_ias()
    assert e.value.code == mod.ERR_NOT_FOUND
    # empty scene: commit succeeds, entry points stay "not ready"
    rt.remove_instance(1)
    rt.build_ias()
    with pytest.raises(mod.RtfError):
        rt.trace_closest(rays)


def test_scene_sync_diffing(rt, orc, S, scenes):
    """rtf_commit rebuilds only what changed (RayTracerLayer.cpp:47-92 update flags): identical re-submission -> skipped,
    material-side change -> tables only (BVH bits untouched), geometry / instance / alpha-class change -> full build."""
    sd = scenes.sorghum_scene(1)
    for gh, (v, t) in sd.geometries.items():
        rt.set_geometry(gh, v, t, version=7)
    for mh, m in sd.materials.items():
        rt.set_material(mh, m)
    for ih, (g, m, x, comp) in sd.instances.items():
        rt.set_instance(ih, g, m, x, comp)
    rt.build_ias(63)
    nodes0 = rt.debug(S.DBG_WIDE_NODES).copy()
    assert rt.commit_stats() == {"full": 1, "materials": 0, "skipped": 0, "uploads_skipped": 0, "refits": 0}
    # the same scene again, same versions
    for gh, (v, t) in sd.geometries.items():
        rt.set_geometry(gh, v, t, version=7)
    for mh, m in sd.materials.items():
        rt.set_material(mh, m)
    for ih, (g, m, x, comp) in sd.instances.items():
        rt.set_instance(ih, g, m, x, comp)
    rt.build_ias(63)
    assert rt.commit_stats() == {"full": 1, "materials": 0, "skipped": 1, "uploads_skipped": len(sd.geometries), "refits": 0}
    # albedo change: tables only, and the new albedo is what the shading sees
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=8)
    before = rt.estimate_illumination_mesh(env, rp, seed=3, tri_begin=0, tri_end=500)
    mh = sorted(sd.materials)[0]
    sd.materials[mh].albedo_color[:] = (0.05, 0.05, 0.05)
    for h, m in sd.materials.items():
        rt.set_material(h, m)
    rt.build_ias(63)
    assert rt.commit_stats()["materials"] == 1 and rt.commit_stats()["full"] == 1
    assert np.array_equal(rt.debug(S.DBG_WIDE_NODES), nodes0)
    after = rt.estimate_illumination_mesh(env, rp, seed=3, tri_begin=0, tri_end=500)
    osc = orc.OracleScene(sd, 63)
    cpu, _ = osc.illum_mesh(env, rp, seed=3, push=1e-3, tri_begin=0, tri_end=500)
    assert np.abs(after[:, :3] - before[:, :3]).max() > 1e-3
    assert np.allclose(after[:, :3], cpu[:, :3], rtol=1e-5, atol=0)
    # transmission > 0 changes the any-hit class baked into the triangle records -> full build
    sd.materials[mh].transmission = 0.5
    rt.set_material(mh, sd.materials[mh])
    rt.build_ias(63)
    assert rt.commit_stats()["full"] == 2
    gt = rt.debug(S.DBG_WIDE_TRIS)
    mat_of_inst = [sd.instances[ih][1] for ih in sorted(sd.instances)]
    want = np.array([1 if mat_of_inst[i] == mh else 0 for i in gt["inst"]], dtype=np.uint32)
    assert np.array_equal(gt["flags"], want) and want.any()  # TRI_FLAG_ALPHA follows the new class (the oracle keeps 0)
    # a moved instance -> full build, same bits as a fresh oracle build
    ih = sorted(sd.instances)[-1]
    g, m, x, comp = sd.instances[ih]
    x = x.copy()
    x[12] += 0.25
    sd.instances[ih] = (g, m, x, comp)
    rt.set_instance(ih, g, m, x, comp)
    rt.build_ias(63)
    assert rt.commit_stats()["full"] == 3
    osc = orc.OracleScene(sd, 63)
    assert np.array_equal(rt.debug(S.DBG_WIDE_NODES).view(np.uint8), osc.debug(S.DBG_WIDE_NODES).view(np.uint8))
    # new geometry version -> upload + full build
    for gh, (v, t) in sd.geometries.items():
        rt.set_geometry(gh, v, t, version=8)
    rt.build_ias(63)
    assert rt.commit_stats()["full"] == 4


def test_refit_of_unmoved_geometry_reproduces_the_build(rt, S, scenes):
    """The in-place refit (segment-tree node boxes + re-encoded wide-node frames + triangle records) of geometry that did
    not move gives the bytes of the full build: same boxes (min/max are exact), same frames, same records."""
    rt.set_option("refit", 1)
    for presplit in (1, 8):
        sd = scenes.sorghum_scene(2)
        rt.load_scene(sd, 63, presplit=presplit)
        before = [rt.debug(w).copy() for w in (S.DBG_BIN_BOXES, S.DBG_WIDE_NODES, S.DBG_WIDE_TRIS)]
        full = rt.commit_stats()["full"]
        refits = rt.commit_stats()["refits"]
        for gh, (v, t) in sd.geometries.items():
            rt.set_geometry(gh, v, t, version=100 + presplit)
        rt.build_ias(63)
        st = rt.commit_stats()
        assert (st["full"], st["refits"]) == (full, refits + 1)
        for a, w in zip(before, ("boxes", "wide nodes", "wide tris")):
            b = rt.debug({"boxes": S.DBG_BIN_BOXES, "wide nodes": S.DBG_WIDE_NODES, "wide tris": S.DBG_WIDE_TRIS}[w])
            assert_bit_equal(b, a, f"{w} after refit (presplit {presplit})")
    rt.set_option("refit", 0)


def test_refit_path(rt, orc, S, scenes):
    """Option "refit": moved instances / vertices with unchanged topology reuse the sorted order and the radix tree of the
    last full build.  Hits, LiDAR and irradiance equal a fresh oracle build of the moved scene (they do not depend on
    the tree); a topology change falls back to the full build."""
    rt.set_option("refit", 1)
    sd = scenes.sorghum_scene(2)
    rt.load_scene(sd, 63)
    assert rt.commit_stats()["full"] == 1
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(1 << 15, lo - 0.3, hi + 0.3, seed=12)

    def check(tag):
        osc = orc.OracleScene(sd, 63)
        assert_bit_equal(rt.trace_closest(rays).view(np.uint32), osc.trace(rays).view(np.uint32), f"hits ({tag})")
        env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=3, samples=8)
        gpu = rt.estimate_illumination_mesh(env, rp, seed=5, push=1e-3, tri_begin=0, tri_end=2000)
        cpu, st = osc.illum_mesh(env, rp, seed=5, push=1e-3, tri_begin=0, tri_end=2000)
        assert np.allclose(gpu[:, :3], cpu[:, :3], rtol=1e-5, atol=0), tag
        assert rt.last_job_stats() == (st.rays, st.paths), tag

    # 1. an instance is moved, turned and scaled (the planned slab counts of its triangles would differ now)
    ih = max(sd.instances)
    g, m, x, comp = sd.instances[ih]
    c, s_ = np.cos(0.9), np.sin(0.9)
    x2 = np.array([[1.3 * c, 0, 1.3 * s_, 0.4], [0, 1.3, 0, 0.05], [-1.3 * s_, 0, 1.3 * c, -0.3], [0, 0, 0, 1]], dtype=np.float32)
    sd.instances[ih] = (g, m, np.ascontiguousarray(x2.T).reshape(16), comp)
    rt.set_instance(ih, *sd.instances[ih])
    rt.build_ias(63)
    st = rt.commit_stats()
    assert (st["full"], st["refits"]) == (1, 1)
    check("moved instance")
    # 2. vertex animation: same triangle count, new positions
    gh = sd.instances[ih][0]
    v, t = sd.geometries[gh]
    v = v.copy()
    v["position"][:, 0] += (0.05 * np.sin(7.0 * v["position"][:, 1])).astype(np.float32)
    sd.geometries[gh] = (v, t)
    rt.set_geometry(gh, v, t, version=2)
    rt.build_ias(63)
    st = rt.commit_stats()
    assert (st["full"], st["refits"]) == (1, 2)
    check("moved vertices")
    # 3. refit off -> full build of the same scene gives the oracle's nodes again
    rt.set_option("refit", 0)
    rt.set_instance(ih, g, m, x, comp)
    sd.instances[ih] = (g, m, x, comp)
    rt.build_ias(63)
    assert rt.commit_stats()["full"] == 2
    osc = orc.OracleScene(sd, 63)
    assert np.array_equal(rt.debug(S.DBG_WIDE_NODES).view(np.uint8), osc.debug(S.DBG_WIDE_NODES).view(np.uint8))
    # 4. topology change with refit on -> full build
    rt.set_option("refit", 1)
    sd.instances[ih + 1] = (g, m, scenes.IDENTITY.copy(), 4711)
    rt.set_instance(ih + 1, *sd.instances[ih + 1])
    rt.build_ias(63)
    st = rt.commit_stats()
    assert (st["full"], st["refits"]) == (3, 2)
    check("added instance")
