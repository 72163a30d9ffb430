"""Pins the oracle's acceleration structure (CPU only).  The reference's BVH lives inside closed-source OptiX
(RayTracer.cu:1528-1584, 1654-1707; optixTrace call sites IlluminationEstimation.cu:66 ...), so the pin is a
brute-force all-triangles intersector (same watertight test, no BVH): every BVH hit must equal it bit for bit."""
import numpy as np
import pytest

from .conftest import assert_bit_equal


def _check_refs(osc, S):
    """Triangle references: every triangle owns >= 1 consecutive references whose boxes tile it."""
    ref_tri = osc.debug(S.DBG_REF_TRI)
    ref_box = osc.debug(S.DBG_REF_BOX).reshape(-1, 6)
    pos = osc.debug(S.DBG_FLAT_POS).reshape(-1, 3, 3).astype(np.float64)
    assert np.all(np.diff(ref_tri.astype(np.int64)) >= 0) and set(ref_tri.tolist()) == set(range(osc.n_tris))
    tlo, thi = pos.min(1), pos.max(1)
    assert np.all(ref_box[:, :3] >= tlo[ref_tri]) and np.all(ref_box[:, 3:] <= thi[ref_tri])
    assert np.all(ref_box[:, :3] <= ref_box[:, 3:])
    # random points of each triangle lie in at least one of its reference boxes
    rng = np.random.default_rng(1)
    first = np.searchsorted(ref_tri, np.arange(osc.n_tris))
    last = np.searchsorted(ref_tri, np.arange(osc.n_tris), side="right")
    for t in rng.integers(0, osc.n_tris, size=min(osc.n_tris, 400)):
        a, b = rng.random(24), rng.random(24)
        flip = a + b > 1
        a[flip], b[flip] = 1 - a[flip], 1 - b[flip]
        pts = (1 - a - b)[:, None] * pos[t, 0] + a[:, None] * pos[t, 1] + b[:, None] * pos[t, 2]
        bx = ref_box[first[t]:last[t]].astype(np.float64)
        tol = 1e-6 * np.maximum(1.0, np.abs(pts).max())
        inside = np.all(pts[:, None, :] >= bx[None, :, :3] - tol, axis=2) & np.all(pts[:, None, :] <= bx[None, :, 3:] + tol, axis=2)
        assert inside.any(axis=1).all(), f"triangle {t}: point outside all {len(bx)} reference boxes"
    return ref_tri, ref_box


def _check_tree(osc, S):
    ref_tri, ref_box = _check_refs(osc, S)
    n = len(ref_tri)
    morton, order = osc.debug(S.DBG_MORTON), osc.debug(S.DBG_SORTED_IDS)
    keys = morton[order]
    assert sorted(order.tolist()) == list(range(n))
    assert np.all(keys[1:] >= keys[:-1])
    same = keys[1:] == keys[:-1]
    assert np.all(order[1:][same] > order[:-1][same])  # stable: ties keep primitive order
    ch = osc.debug(S.DBG_BIN_CHILDREN).reshape(-1, 2)
    par = osc.debug(S.DBG_BIN_PARENT)
    boxes = osc.debug(S.DBG_BIN_BOXES).reshape(-1, 6)
    ni = n - 1
    if n > 1:
        assert par[0] == -1
        seen = np.zeros(2 * n - 1, dtype=np.int32)
        for i in range(ni):
            for c in ch[i]:
                idx = c if c >= 0 else ni + (~c)
                assert par[idx] == i
                seen[idx] += 1
                assert np.all(boxes[idx, :3] >= boxes[i, :3]) and np.all(boxes[idx, 3:] <= boxes[i, 3:])
        assert seen[0] == 0 and np.all(seen[1:] == 1)
    nodes, tris = osc.debug(S.DBG_WIDE_NODES), osc.debug(S.DBG_WIDE_TRIS)
    assert sorted(tris["prim"].tolist()) == ref_tri.tolist()  # every reference exactly once
    pos = osc.debug(S.DBG_FLAT_POS).reshape(-1, 3, 3)
    assert_bit_equal(tris["v0"], pos[tris["prim"], 0], "tri v0")
    # quantised child boxes contain the triangles they lead to (leaf slots checked directly)
    for nd in nodes:
        scale = (nd["e"].astype(np.uint32) << 23).view(np.float32)
        for s, off, cnt in S.node_leaf_slots(nd):
            t0 = int(nd["tri_base"]) + off
            lo = nd["p"] + nd["qlo"][:, s].astype(np.float32) * scale
            hi = nd["p"] + nd["qhi"][:, s].astype(np.float32) * scale
            for t in tris[t0:t0 + cnt]:  # the slot holds (at least) one reference box of each of its triangles
                rb = ref_box[ref_tri == t["prim"]]
                assert (np.all(rb[:, :3] >= lo, axis=1) & np.all(rb[:, 3:] <= hi, axis=1)).any()
    # child_base / imask consistency: internal children are contiguous, BFS order
    n_int = sum(bin(int(nd["imask"])).count("1") for nd in nodes)
    assert n_int == len(nodes) - 1


@pytest.mark.parametrize("n,bits", [(1, 63), (2, 63), (3, 30), (4, 63), (9, 63), (257, 30), (1500, 63)])
@pytest.mark.parametrize("presplit", [1, 8])
def test_tree_invariants(orc, S, scenes, n, bits, presplit):
    osc = orc.OracleScene(scenes.random_soup(n, seed=n), bits, presplit=presplit)
    assert osc.n_tris == n
    assert (len(osc.debug(S.DBG_REF_TRI)) == n) if presplit == 1 else (len(osc.debug(S.DBG_REF_TRI)) >= n)
    _check_tree(osc, S)


def test_presplit_cuts_slivers_and_keeps_hits(orc, S, scenes):
    """Sorghum leaves are diagonal slivers: pre-splitting multiplies their references, divides the triangle tests per
    ray, and leaves every hit bit-identical (to the unsplit BVH and to brute force)."""
    sd = scenes.sorghum_scene(1)
    plain, split = orc.OracleScene(sd, 63, presplit=1), orc.OracleScene(sd, 63, presplit=8)
    _check_refs(split, S)
    n_ref = len(split.debug(S.DBG_REF_TRI))
    assert 2 * plain.n_tris < n_ref <= 8 * plain.n_tris
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(20000, lo, hi, seed=5)
    h0, _, tv0 = plain.trace(rays, counts=True)
    h1, _, tv1 = split.trace(rays, counts=True)
    assert_bit_equal(h0, h1, "hits with and without pre-splitting")
    assert_bit_equal(h1[:2000], split.trace_brute(rays[:2000]), "hits vs brute force")
    assert tv1.mean() < 0.8 * tv0.mean()


def test_duplicate_triangles_and_degenerates(orc, S, scenes):
    sd = scenes.random_soup(64, seed=5, size=0.6)
    v, t = sd.geometries[1]
    v = np.concatenate([v, v, v[:3]])  # duplicates -> equal Morton keys (Karras index tie-break)
    t = np.concatenate([t, t + 192, [[384, 384, 384]], [[384, 385, 385]]]).astype(np.uint32)  # + degenerate
    sd.geometries[1] = (v, t)
    osc = orc.OracleScene(sd, 30)
    _check_tree(osc, S)
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(3000, lo - 0.1, hi + 0.1, seed=9)
    assert_bit_equal(osc.trace(rays), osc.trace_brute(rays), "dup hits")
    # equal-t ties resolve to the lower primitive id
    h = osc.trace(rays)
    hit = h["prim"] != 0xFFFFFFFF
    assert hit.sum() > 100 and np.all(h["prim"][hit] < 64)


@pytest.mark.parametrize("bits", [30, 63])
def test_bvh_hits_equal_brute_force_soup(orc, scenes, bits):
    sd = scenes.random_soup(3000, seed=21, size=0.25)
    osc = orc.OracleScene(sd, bits)
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(20000, lo, hi, seed=4)
    a, b = osc.trace(rays), osc.trace_brute(rays)
    assert_bit_equal(a, b, "hits")
    assert (a["prim"] != 0xFFFFFFFF).mean() > 0.3


def test_bvh_hits_equal_brute_force_sorghum(orc, scenes):
    sd = scenes.sorghum_scene(1)
    osc = orc.OracleScene(sd, 63)
    assert osc.n_tris == 20002
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(3000, lo, hi, seed=3)
    a, b = osc.trace(rays), osc.trace_brute(rays)
    assert_bit_equal(a, b, "hits")
    # axis-aligned and zero-component directions (safe_dir path), rays starting on the ground plane
    ax = np.zeros((6, 8), dtype=np.float32)
    ax[:, 0:3] = (1.9, 1.0, 1.9)
    ax[:, 7] = 1e20
    for k in range(6):
        ax[k, 4 + k % 3] = 1.0 if k < 3 else -1.0
    assert_bit_equal(osc.trace(ax), osc.trace_brute(ax), "axis rays")
    h = osc.trace(ax)
    assert h["prim"][4] in (0, 1) and h["t"][4] == 1.0  # straight down onto the ground (primitives 0,1)
    # axis-parallel rays running exactly inside a face of the scene box (the ground's edges and corners):
    # the brute-force intersector hits the boundary triangles, so the box test must not cull them
    face = []
    for x in (-2.0, 2.0):
        for z in (-2.0, 0.3, 2.0):
            face += [[x, 1.0, z, 0.0, 0, -1, 0, 1e20], [x, 1.0, z, 0.0, -0.0, -1, -0.0, 1e20], [z, 1.0, x, 0.0, 0, -1, 0, 1e20]]
    face = np.array(face, dtype=np.float32)
    hf = osc.trace(face)
    assert_bit_equal(hf, osc.trace_brute(face), "on-face rays")
    assert (hf["prim"] != 0xFFFFFFFF).all()


def test_tmin_tmax_window(orc, scenes):
    sd = scenes.sorghum_scene(1)
    osc = orc.OracleScene(sd, 63)
    r = np.array([[0.5, 2.0, 0.5, 0.0, 0, -1, 0, 1e20], [0.5, 2.0, 0.5, 0.0, 0, -1, 0, 1.5],
                  [0.5, 2.0, 0.5, 2.5, 0, -1, 0, 1e20], [0.5, 2.0, 0.5, 2.0, 0, -1, 0, 2.0]], dtype=np.float32)
    h = osc.trace(r)
    assert h["t"][0] == 2.0 and h["prim"][0] in (0, 1)
    assert h["prim"][1] == 0xFFFFFFFF and h["prim"][2] == 0xFFFFFFFF
    assert h["prim"][3] in (0, 1)  # t == tmin == tmax is accepted
    assert_bit_equal(h, osc.trace_brute(r), "window")


def test_visit_counts_are_deterministic(orc, scenes):
    sd = scenes.random_soup(2000, seed=7)
    osc = orc.OracleScene(sd, 63)
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(2000, lo, hi, seed=3)
    _, nv1, tv1 = osc.trace(rays, counts=True, nthreads=1)
    _, nv2, tv2 = osc.trace(rays, counts=True, nthreads=4)
    assert np.array_equal(nv1, nv2) and np.array_equal(tv1, tv2)
    assert nv1.min() >= 1 and tv1.mean() < 40
