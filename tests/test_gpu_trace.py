"""GPU closest-hit traversal vs the CPU oracle: {t,u,v,primId} and node/triangle visit counts bit for bit."""
import os

import numpy as np
import pytest

from .conftest import GOLDEN, assert_bit_equal

pytestmark = pytest.mark.gpu


def test_golden_hits(rt, S):
    from tests.golden.make_golden import small_cases
    g = np.load(os.path.join(GOLDEN, "oracle_small.npz"))
    c = small_cases()
    rt.load_scene(c["soup"], 63, presplit=1)  # the fixture's visit counts belong to the unsplit BVH
    hits, nv, tv = rt.trace_closest(c["rays"], counts=True)
    assert_bit_equal(hits.view(np.uint32).reshape(-1, 4), g["hits"], "hits vs golden")
    assert np.array_equal(nv, g["node_visits"]) and np.array_equal(tv, g["tri_visits"])
    assert_bit_equal(rt.trace_closest(c["rays"]), hits, "stats-free kernel")
    rt.load_scene(c["soup"], 63, presplit=8)  # pre-split references change the BVH, never the hits
    assert_bit_equal(rt.trace_closest(c["rays"]).view(np.uint32).reshape(-1, 4), g["hits"], "hits vs golden (presplit 8)")


def test_oracle_bvh_uploaded(rt, orc, S, scenes):
    """Traversal in isolation: BVH built by the oracle, uploaded, traced on the GPU."""
    sd = scenes.random_soup(3000, seed=21, size=0.25)
    osc = orc.OracleScene(sd, 63)
    rt.debug_upload_bvh(osc.debug(S.DBG_WIDE_NODES), osc.debug(S.DBG_WIDE_TRIS), osc.depth)
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(50000, lo, hi, seed=4)
    assert_bit_equal(rt.trace_closest(rays), osc.trace(rays), "hits")


@pytest.mark.parametrize("bits", [30, 63])
def test_sorghum_incoherent_hits(rt, orc, S, scenes, bits):
    sd = scenes.sorghum_scene(1)
    rt.load_scene(sd, bits)
    osc = orc.OracleScene(sd, bits)
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(1 << 17, lo, hi, seed=3)
    hits, nv, tv = rt.trace_closest(rays, counts=True)
    oh, onv, otv = osc.trace(rays, counts=True)
    assert_bit_equal(hits, oh, "hits")
    assert np.array_equal(nv, onv) and np.array_equal(tv, otv)
    # ground truth independent of any BVH on a prefix
    assert_bit_equal(hits[:2000], osc.trace_brute(rays[:2000]), "brute force")


def test_edge_rays(rt, orc, scenes):
    sd = scenes.sorghum_scene(1)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    r = []
    for k in range(6):  # axis-aligned directions: zero components go through safe_dir
        d = [0.0, 0.0, 0.0]
        d[k % 3] = 1.0 if k < 3 else -1.0
        r.append([1.9, 1.0, 1.9, 0.0, *d, 1e20])
    r += [[0.5, 2.0, 0.5, 0.0, 0, -1, 0, 1e20], [0.5, 2.0, 0.5, 0.0, 0, -1, 0, 1.5],
          [0.5, 2.0, 0.5, 2.5, 0, -1, 0, 1e20], [0.5, 2.0, 0.5, 2.0, 0, -1, 0, 2.0],
          [0.0, 5.0, 0.0, 0.0, 0, -1, 0, 1e20],       # down the stem axis / panicle pole
          [2.0, 1.0, 2.0, 0.0, 0, -1, 0, 1e20],       # exactly on the ground-plane corner
          [0.0, 1.0, 0.0, 0.0, 1e-30, -1, 1e-30, 1e20],  # denormal-ish components
          [100.0, 100.0, 100.0, 0.0, 1, 1, 1, 1e20]]  # pointing away
    for x in (-2.0, 2.0):  # axis-parallel rays running exactly inside a face of the scene box (+0 and -0 components)
        for z in (-2.0, 0.3, 2.0):
            r += [[x, 1.0, z, 0.0, 0, -1, 0, 1e20], [x, 1.0, z, 0.0, -0.0, -1, -0.0, 1e20], [z, 1.0, x, 0.0, 0, -1, 0, 1e20]]
    r = np.array(r, dtype=np.float32)
    assert_bit_equal(rt.trace_closest(r), osc.trace(r), "edge rays")
    assert_bit_equal(rt.trace_closest(r), osc.trace_brute(r), "edge rays vs brute force")
    assert len(rt.trace_closest(np.zeros((0, 8), dtype=np.float32))) == 0  # empty input


def test_ragged_counts(rt, orc, scenes):
    sd = scenes.random_soup(500, seed=8, size=0.3)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    lo, hi = scenes.scene_bounds(sd)
    for n in (1, 31, 32, 33, 127, 4097):
        rays = scenes.incoherent_rays(n, lo, hi, seed=n)
        assert_bit_equal(rt.trace_closest(rays), osc.trace(rays), f"n={n}")


def test_field_1m_triangles_properties(rt, orc, S, scenes):
    """Full-size scene (C2: ~1 M triangles): size-independent properties + oracle on a sub-sample."""
    sd = scenes.sorghum_scene(50)
    rt.load_scene(sd, 63)
    assert rt.triangle_count == 1000002
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(1 << 20, lo, hi, seed=3)
    hits = rt.trace_closest(rays)
    hit = hits["prim"] != 0xFFFFFFFF
    assert 0.2 < hit.mean() < 1.0
    assert np.all(hits["t"][hit] >= 0) and np.all(hits["prim"][hit] < 1000002)
    uv = hits["u"][hit] + hits["v"][hit]
    assert np.all(hits["u"][hit] >= 0) and np.all(hits["v"][hit] >= 0) and np.all(uv <= 1.0 + 1e-6)
    # idempotence: re-tracing with tmax just beyond t returns the same hit.  (With tmax == t exactly the property
    # holds only up to the rounding of the watertight t of millimetre-sized triangles seen from metres away, whose
    # error can exceed the slack of their leaf box: assert a 99 % floor there, not equality.)
    r2 = rays.copy()
    r2[hit, 7] = hits["t"][hit] * np.float32(1.001)
    h2 = rt.trace_closest(r2)
    assert_bit_equal(h2[hit], hits[hit], "closest hit is stable under tmax = 1.001 t")
    r2[hit, 7] = hits["t"][hit]
    h3 = rt.trace_closest(r2)
    same = (h3.view(np.uint32).reshape(-1, 4) == hits.view(np.uint32).reshape(-1, 4)).all(1)
    assert same[hit].mean() > 0.99 and same[~hit].all()
    # the hit point lies on the triangle's plane: reconstruct from barycentrics
    pos = rt.debug(S.DBG_FLAT_POS).reshape(-1, 3, 3)
    idx = np.nonzero(hit)[0][:20000]
    p = pos[hits["prim"][idx]]
    bary = (1 - hits["u"][idx] - hits["v"][idx])[:, None] * p[:, 0] + hits["u"][idx][:, None] * p[:, 1] + hits["v"][idx][:, None] * p[:, 2]
    along = rays[idx, 0:3] + hits["t"][idx][:, None] * rays[idx, 4:7]
    assert np.max(np.abs(bary - along)) < 2e-3
    # oracle on every 64th ray (the oracle builds the same 1 M-triangle BVH on the CPU)
    osc = orc.OracleScene(sd, 63)
    sub = rays[::64]
    assert_bit_equal(hits[::64], osc.trace(sub), "sub-sample vs oracle")


def test_host_pipeline_chunks(rt, orc, scenes):
    """rtf_trace_closest with host buffers larger than one pipeline chunk: H2D / trace / D2H overlap over chunks,
    ragged last chunk; result identical to the single-shot path and to the oracle."""
    sd = scenes.sorghum_scene(1)
    rt.load_scene(sd, 63)
    osc = orc.OracleScene(sd, 63)
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(4096 * 5 + 777, lo, hi, seed=21)
    rt.set_option("pipe_chunk_rays", 4096)
    piped = rt.trace_closest(rays)
    rt.set_option("pipe_chunk_rays", 1 << 22)
    single = rt.trace_closest(rays)
    assert_bit_equal(piped, single, "pipelined vs single shot")
    assert_bit_equal(piped, osc.trace(rays), "pipelined vs oracle")
