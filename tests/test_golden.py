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
