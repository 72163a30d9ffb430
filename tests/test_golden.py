"""The oracle reproduces the committed fixtures of tests/golden/oracle_small.npz bit for bit (CPU), so a change to
the oracle that alters any hit, code or irradiance is caught even without a GPU."""
import os

import numpy as np

from .conftest import GOLDEN, assert_bit_equal


def test_oracle_reproduces_golden(orc, S):
    from tests.golden.make_golden import camera_props, small_cases
    g = np.load(os.path.join(GOLDEN, "oracle_small.npz"))
    c = small_cases()
    osc = orc.OracleScene(c["soup"], 63, presplit=1)
    hits, nv, tv = osc.trace(c["rays"], counts=True)
    assert_bit_equal(hits.view(np.uint32).reshape(-1, 4), g["hits"], "hits")
    assert np.array_equal(nv, g["node_visits"]) and np.array_equal(tv, g["tri_visits"])
    assert np.array_equal(osc.debug(S.DBG_MORTON), g["morton63"])
    assert np.array_equal(osc.debug(S.DBG_SORTED_IDS), g["order63"])
    assert [osc.n_nodes, osc.depth] == g["n_nodes"].tolist()
    # pre-split references change the BVH, never the hits: the fixtures were made without splitting
    split = orc.OracleScene(c["soup"], 63, presplit=8)
    assert len(split.debug(S.DBG_REF_TRI)) > split.n_tris
    assert_bit_equal(split.trace(c["rays"]).view(np.uint32).reshape(-1, 4), g["hits"], "hits (presplit 8)")
    o30 = orc.OracleScene(c["soup"], 30, presplit=1)
    assert np.array_equal(o30.debug(S.DBG_MORTON), g["morton30"]) and np.array_equal(o30.debug(S.DBG_SORTED_IDS), g["order30"])
    ill = orc.OracleScene(c["illum_scene"], 63)
    irr, st = ill.illum_mesh(c["env"], c["rp"], seed=0x5EED, push=1e-3)
    assert_bit_equal(irr.view(np.uint32), g["irradiance"], "irradiance")
    assert [st.rays, st.paths] == g["illum_rays"].tolist()
    t, prim, _, _ = ill.scan_grid(c["grid"])
    assert_bit_equal(t.view(np.uint32), g["lidar_t"], "lidar t")
    assert np.array_equal(prim, g["lidar_prim"])
    col, nrm, alb, _ = ill.camera(c["env"], S.RayProperties(bounces=2, samples=4), camera_props())
    assert_bit_equal(col.view(np.uint32), g["camera_color"], "camera colour")
    assert_bit_equal(nrm.view(np.uint32), g["camera_normal"], "camera normal")
    assert_bit_equal(alb.view(np.uint32), g["camera_albedo"], "camera albedo")


def test_oracle_reproduces_extra_golden(orc, S):
    """BSSRDF probe / irradiance, instanced and skinned pre-kernels, light probes and flux vs tests/golden/oracle_extra.npz."""
    from tests.golden.make_golden import extra_cases
    g = np.load(os.path.join(GOLDEN, "oracle_extra.npz"))
    x = extra_cases()
    osc = orc.OracleScene(x["sss_scene"], 63)
    hs = np.full(len(x["probe_rays"]), x["handle"], dtype=np.uint64)
    assert np.array_equal(osc.sss_probe(x["probe_rays"], hs), g["sss_records"])
    assert np.array_equal(osc.sss_probe(x["probe_rays"], hs, brute=True), g["sss_records"])
    irr, st = osc.illum_mesh(x["env"], x["rp"], seed=0xABCD, push=1e-3)
    assert_bit_equal(irr.view(np.uint32), g["sss_irradiance"], "BSSRDF irradiance")
    assert [st.rays, st.paths] == g["sss_rays"].tolist()
    bv, bt = orc.bake_instanced(x["verts"], x["tris"], x["mats"], x["cols"])
    assert np.array_equal(bv.view(np.uint8), g["bake_instanced"]) and np.array_equal(bt, g["bake_tris"])
    assert np.array_equal(orc.bake_skinned(x["skinned"], x["mats"]).view(np.uint8), g["bake_skinned"])
    pr = osc.make_probes(front=True, back=True)
    pr["energy"] = irr[:, :3]
    assert np.array_equal(pr.view(np.uint8), g["probes"])
    assert_bit_equal(orc.flux_probes(pr).view(np.uint32), g["flux"], "flux")
