"""CPU: the oracle's restatement of the BSSRDF spatial probe (BSSDF.cuh:18-89, RayFunctions.cuh:97-133): BVH probe vs brute
force, glm::rotate closed forms, and a closed-form acceptance probability on a slab."""
import numpy as np

from .conftest import assert_bit_equal


def slab_scene(scenes, S, half=50.0, thickness=0.05, handle=4242):
    """Box [-half,half] x [-thickness,0] x [-half,half]: per-face vertices (flat normals), 12 triangles, one instance."""
    lo, hi = np.array([-half, -thickness, -half]), np.array([half, 0.0, half])
    pos, nrm, tris = [], [], []
    for axis in range(3):
        for sign in (-1, 1):
            a, b = (axis + 1) % 3, (axis + 2) % 3
            n = np.zeros(3)
            n[axis] = sign
            corners = []
            for sa, sb in ((0, 0), (1, 0), (1, 1), (0, 1)):
                c = np.zeros(3)
                c[axis] = hi[axis] if sign > 0 else lo[axis]
                c[a] = hi[a] if sa else lo[a]
                c[b] = hi[b] if sb else lo[b]
                corners.append(c)
            base = len(pos)
            pos += corners
            nrm += [n] * 4
            tris += [[base, base + 1, base + 2], [base, base + 2, base + 3]]
    sd = scenes.SceneDesc()
    sd.geometries[1] = (scenes.make_vertices(np.array(pos), np.array(nrm)), np.array(tris, dtype=np.uint32))
    m = S.Material(albedo=(0.8, 0.7, 0.6), metallic=1.0, roughness=0.0)
    m.subsurface_factor = 1.0
    m.subsurface_color[:] = (1.0, 1.0, 1.0)
    m.subsurface_radius[:] = (4 * thickness, 0.0, 0.0)  # thickness = radius / 4
    sd.materials[1] = m
    sd.instances[1] = (1, 1, scenes.IDENTITY.copy(), handle)
    return sd


def down_probe(S, height=0.01, size=0.01):
    """One light probe just above the slab centre, facing down (front face only)."""
    pr = np.zeros(1, dtype=S.ILLUM_SAMPLER_DTYPE)
    for name, p in (("a", (0, height, 0)), ("b", (size, height, 0)), ("c", (0, height, size))):
        pr[name]["position"] = p
        pr[name]["normal"] = (0, -1, 0)
    pr["front_face"], pr["back_face"] = 1, 0
    return pr


def sss_field(scenes, S):
    """Sorghum plant (subsurface leaves) + ground: several handles, pre-split slivers."""
    sd = scenes.sorghum_scene(1)
    for mh, m in sd.materials.items():
        m.subsurface_factor = 0.4
        m.subsurface_color[:] = (0.3, 0.9, 0.2)
        m.subsurface_radius[:] = (0.03, 0.0, 0.0)
    return sd


def probe_rays(scenes, sd, n, seed, length=0.05):
    lo, hi = scenes.scene_bounds(sd)
    rays = scenes.incoherent_rays(n, lo, hi, seed=seed)
    rays[:, 3] = scenes.hash_uniform(seed, np.arange(n), 7) * length * 0.5
    rays[:, 7] = rays[:, 3] + length
    return rays


def test_rotate_about_closed_forms(orc):
    assert np.allclose(orc.rotate_about((1, 0, 0), 0.25, (0, 0, 1)), (0, 1, 0), atol=1e-6)
    assert np.allclose(orc.rotate_about((1, 0, 0), 0.25, (0, 2, 0)), (0, 0, -1), atol=1e-6)  # axis is normalised
    v = np.array([0.3, -0.7, 0.2])
    assert np.allclose(orc.rotate_about(v, 0.0, (0.1, 0.9, 0.3)), v, atol=1e-6)
    ax = np.array([0.2, 0.5, -0.8])
    ax /= np.linalg.norm(ax)
    for u in (0.1, 0.37, 0.93):
        a = 2 * np.pi * u
        rod = v * np.cos(a) + np.cross(ax, v) * np.sin(a) + ax * np.dot(ax, v) * (1 - np.cos(a))
        assert np.allclose(orc.rotate_about(v, u, ax), rod, atol=2e-6)


def test_probe_bvh_equals_brute_force(orc, S, scenes):
    sd = sss_field(scenes, S)
    osc = orc.OracleScene(sd, 63)  # default pre-split: one triangle through several references
    handles = sorted({c for (_, _, _, c) in sd.instances.values()})
    rays = probe_rays(scenes, sd, 6000, seed=21, length=4.0)
    most = 0
    for k, h in enumerate(handles + [999]):
        hs = np.full(len(rays), h, dtype=np.uint64)
        a, b = osc.sss_probe(rays, hs), osc.sss_probe(rays, hs, brute=True)
        assert np.array_equal(a, b)
        n = a[:, 0]
        assert (n <= 4).all()
        if h == 999:
            assert n.sum() == 0
            continue
        assert (n > 0).sum() > 20
        most = max(most, int(n.max()))
        t = a[:, 5:9].view(np.float32)
        for j in range(3):  # ascending (t, prim), no duplicates
            both = n > j + 1
            assert ((t[both, j] < t[both, j + 1]) | ((t[both, j] == t[both, j + 1]) & (a[both, 1 + j] < a[both, 2 + j]))).all()
        inside = n > 0
        assert (t[inside, 0] >= rays[inside, 3]).all() and (t[inside, 0] <= rays[inside, 7]).all()
    assert most >= 3  # multi-hit records do occur
    # the nearest record is the closest hit of the same interval when only one handle is in reach
    sd1 = slab_scene(scenes, S)
    o1 = orc.OracleScene(sd1, 63)
    r1 = probe_rays(scenes, sd1, 2000, seed=5, length=1.0)
    rec = o1.sss_probe(r1, np.full(len(r1), 4242, dtype=np.uint64))
    hit = o1.trace(r1)
    got = rec[:, 0] > 0
    assert np.array_equal(got, hit["prim"] != 0xFFFFFFFF)
    assert np.array_equal(rec[got, 1], hit["prim"][got]) and np.array_equal(rec[got, 5], hit["t"][got].view(np.uint32))


def test_slab_acceptance_probability(orc, S, scenes):
    """factor 1, colour 1, metallic 1, roughness 0: every accepted subsurface sample leaves along the record's normal
    into the sky with weight 1, the bounce path has weight 0.  Thickness = radius/4 gives
    P(accept) = 5/16 + 1/8 + 3/32 = 17/32 (derivation in DESIGN.md), so E[energy] = sky * 17/32 * E[cos] = sky*17/64."""
    sd = slab_scene(scenes, S)
    osc = orc.OracleScene(sd, 63)
    pr = down_probe(S)
    env = S.EnvironmentProperties(color=(1.0, 0.5, 0.25))
    n = 1 << 16
    st = osc.illum_probes(env, S.RayProperties(bounces=1, samples=n), pr, seed=77, push=0.0)
    e = pr["energy"][0]
    expect = np.array([1.0, 0.5, 0.25]) * 17.0 / 64.0
    sd_mean = 0.33 / np.sqrt(n)
    assert np.all(np.abs(e - expect) < 4 * sd_mean * np.array([1.0, 0.5, 0.25]) + 2e-4), (e, expect)
    # rays: primary + subsurface ray for the accepted samples (the bounce ray has weight 0 and is not traced)
    assert st.paths == n and abs(st.rays - n * (1 + 17 / 32)) < 0.01 * n
    # factor 0 switches the branch off: plain mirror bounce (metallic 1: straight back up) -> sky * albedo * E[cos]
    sd.materials[1].subsurface_factor = 0.0
    osc0 = orc.OracleScene(sd, 63)
    pr0 = down_probe(S)
    osc0.illum_probes(env, S.RayProperties(bounces=1, samples=4096), pr0, seed=77, push=0.0)
    assert np.allclose(pr0["energy"][0], np.array([1.0, 0.5, 0.25]) * np.array([0.8, 0.7, 0.6]) * 0.5, rtol=0.03)
