"""Pins the oracle's table-shaped material inputs (CPU only): portable inverse trigonometry, the tex2D restatement,
the cube-map lookup (RayTracerUtilities.cuh:54-80) and the compressed-BTF cascade (MLVQ/*.cuh).  The reference ships no
textures, BTF database or tests, so the pins are closed forms: texel centres, wrap periodicity, face axes, constant
tables -> constant colour, Hermite end points, horizon rules."""
import math

import numpy as np
import pytest


def test_portable_inverse_trig_matches_libm(orc):
    rng = np.random.default_rng(0)
    y = np.concatenate([rng.uniform(-1, 1, 20000), [0.0, 1.0, -1.0, 0.5, -0.5]]).astype(np.float32)
    x = np.concatenate([rng.uniform(-1, 1, 20000), [1.0, 0.0, 0.0, -0.5, 0.5]]).astype(np.float32)
    a, s, c = orc.inverse_trig(y, x)
    assert np.max(np.abs(a - np.arctan2(y.astype(np.float64), x.astype(np.float64)))) < 6e-7
    # asin/acos lose relative accuracy only where sqrt(1 - y^2) does (|y| -> 1): absolute error stays tiny
    assert np.max(np.abs(s - np.arcsin(y.astype(np.float64)))) < 4e-4
    inner = np.abs(y) < 0.999
    assert np.max(np.abs(s[inner] - np.arcsin(y[inner].astype(np.float64)))) < 2e-6
    assert np.max(np.abs(c[inner] - np.arccos(y[inner].astype(np.float64)))) < 2e-6
    a2, s2, c2 = orc.inverse_trig(np.array([np.nan, 2.0], np.float32), np.array([1.0, 1.0], np.float32))
    assert np.isnan(a2[0]) and np.isnan(s2[1]) and np.isnan(c2[1])  # out-of-range acos stays NaN (RDD.hpp:319 relies on it)


def _scene_with_textures(orc, scenes):
    sd = scenes.materials_scene()
    return sd, orc.OracleScene(sd)


def test_tex2d_texel_centres_wrap_and_midpoints(orc, scenes):
    sd, osc = _scene_with_textures(orc, scenes)
    img = sd.textures[1]  # 16 rows x 32 columns
    h, w = img.shape[:2]
    ys, xs = np.mgrid[0:h, 0:w]
    uv = np.stack([(xs.ravel() + 0.5) / w, (ys.ravel() + 0.5) / h], axis=1).astype(np.float32)
    out = osc.sample_texture(1, uv)
    assert np.array_equal(out, img.reshape(-1, 4))  # texel centres return the texel, exactly
    assert np.array_equal(osc.sample_texture(1, uv + np.float32(3.0)), osc.sample_texture(1, uv + np.float32(-2.0)))  # wrap
    mid = np.array([[1.0 / w, 0.5 / h]], dtype=np.float32)  # halfway between texel (0,0) and (1,0)
    assert np.allclose(osc.sample_texture(1, mid)[0], 0.5 * (img[0, 0] + img[0, 1]), atol=1e-7)
    edge = np.array([[0.0, 0.5 / h]], dtype=np.float32)  # u = 0: halfway between the last and the first column (wrap)
    assert np.allclose(osc.sample_texture(1, edge)[0], 0.5 * (img[0, -1] + img[0, 0]), atol=1e-7)


def test_cube_map_faces_and_environment(orc, scenes, S):
    sd, osc = _scene_with_textures(orc, scenes)
    env = S.EnvironmentProperties()
    env.environmental_map_id = 7
    env.skylight_intensity = 2.0
    axes = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], dtype=np.float32)
    pd = np.concatenate([np.zeros((6, 3), np.float32), axes], axis=1)
    got = osc.sample_environment(env, pd)
    for k in range(6):  # an axis direction lands on uv = (0.5, 0.5) of its face: the mean of the four centre texels
        img = sd.textures[11 + k]
        centre = img[3:5, 3:5, :3].reshape(-1, 3).mean(0)
        assert np.allclose(got[k], 2.0 * centre, atol=1e-6), k
    env.environmental_map_id = 0  # falls back to the constant colour (Env.cuh:159-161)
    assert np.allclose(osc.sample_environment(env, pd), 2.0)


def _const_btf(scenes, S, y=120.0, cb=110.0, cr=140.0, **params):
    b = scenes.synthetic_btf(texels=(4, 4), n_pdf1d=8, n_colors=4, n_index=4, n_pdf2d_color=4, n_pdf2d_lum=4, n_pdf2d=4,
                             n_pdf3d=4, n_pdf4d=4, **params)
    b.t["pdf1d_basis"][:] = y
    b.t["colors"][:, 0], b.t["colors"][:, 1] = cb, cr
    for k in ("pdf2d_scales", "pdf3d_scales", "pdf4d_scales", "pdf6_scales"):
        b.t[k][:] = 1.0
    return b


def _ycbcr(y, cb, cr):
    r = y * 1.1643828 + cr * 1.5960273 - 222.921
    g = y * 1.1643828 - cb * 0.39176172 - cr * 0.81296875 + 135.576
    b = y * 1.1643828 + cb * 2.0172344 - 276.836
    return np.maximum(np.array([r, g, b]), 0.0)


def test_btf_constant_tables_give_constant_colour(orc, scenes, S):
    """Every interpolation weight pair sums to 1 and the Hermite spline of a constant is the constant, so constant
    tables must give YCbCr->RGB(y, cb, cr)^gamma / 256 for every direction pair above the horizon
    (BTFBase.cuh:49-81, CIELab.cuh:8-28, RDD.hpp:349)."""
    sd = scenes.SceneDesc()
    sd.btfs[1] = _const_btf(scenes, S)
    osc = orc.OracleScene(sd)
    rng = np.random.default_rng(1)
    n = 4000

    def hemi():
        d = rng.normal(size=(n, 3))
        d[:, 1] = np.abs(d[:, 1]) + 0.05
        return (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    illum, out_dir = hemi(), hemi()
    normal = np.tile(np.array([0, 1, 0], np.float32), (n, 1))
    tangent = np.tile(np.array([1, 0, 0], np.float32), (n, 1))
    tc = rng.uniform(-2, 3, size=(n, 2)).astype(np.float32)
    got = osc.btf_eval(1, tc, -out_dir, illum, normal, tangent)  # view_dir = direction of the incoming ray
    want = _ycbcr(120.0, 110.0, 140.0) ** 2.2 / 256.0
    assert np.allclose(got, want[None, :], rtol=2e-5)
    # light below the horizon -> theta = 90, phi = 0 (RDD.hpp:310-314): still a valid lookup, same constant
    below = illum.copy()
    below[:, 1] *= -1
    assert np.allclose(osc.btf_eval(1, tc, -out_dir, below, normal, tangent), want[None, :], rtol=2e-5)


def test_btf_hermite_interpolates_the_beta_samples(orc, scenes, S):
    """At an illumination direction whose beta equals a tabulated beta angle, the 1-D Hermite spline returns the sample
    itself (PDF1D.cuh:57-66: w = 0 -> h00 = 1): luminance varies with beta exactly as the table says."""
    b = _const_btf(scenes, S, use_cos_beta=False)
    nb = b.t["pdf1d_basis"].shape[1]
    b.t["pdf1d_basis"][:] = (60.0 + 10.0 * np.arange(nb))[None, :]
    sd = scenes.SceneDesc()
    sd.btfs[1] = b
    osc = orc.OracleScene(sd)
    angles = osc.btf_beta_angles(1, nb)
    assert angles[0] == -90 and angles[nb // 2] == 0 and angles[-1] == 90 and np.all(np.diff(angles) > 0)
    normal, tangent = np.array([[0, 1, 0]], np.float32), np.array([[1, 0, 0]], np.float32)
    view = np.array([[0, -1, 0]], np.float32)  # looking straight down: view theta = 0, view phi = 0
    for i in (2, nb // 2, nb - 3):
        # Tangent frame (T, B, N) = (x, -z, y) (RDD.hpp:297-305): phi_i = atan2(-d.z, d.x).  View phi = 0 -> illumination
        # phi' = (360 - phi_i) - 90 (PDF6D.cuh:49-52) and beta = asin(cos(phi') sin(theta_i)) (SharedCoordinates.cuh:149-157):
        # light in the +z half plane has phi_i = 270, phi' = 0, beta = +theta_i; in the -z half plane beta = -theta_i.
        be = math.radians(float(angles[i]))
        illum = np.array([[0.0, math.cos(abs(be)), math.sin(be)]], np.float32)
        got = osc.btf_eval(1, np.zeros((1, 2), np.float32), view, illum, normal, tangent)[0]
        want = _ycbcr(60.0 + 10.0 * i, 110.0, 140.0) ** 2.2 / 256.0
        assert np.allclose(got, want, rtol=2e-3), (i, got, want)


def test_btf_scene_runs_and_hdr_branch(orc, scenes, S):
    sd = scenes.materials_scene()
    osc = orc.OracleScene(sd)
    env, rp = S.EnvironmentProperties(), S.RayProperties(bounces=2, samples=8)
    env.environmental_map_id = 1
    irr, st = osc.illum_mesh(env, rp, seed=0x5EED, push=1e-3)
    assert np.isfinite(irr[:, :3]).all() and irr[:, :3].max() > 0 and st.rays > st.paths
    # HDR (LogLuv) colour model branch stays finite and positive
    sd2 = scenes.SceneDesc()
    sd2.btfs[1] = _const_btf(scenes, S, y=60.0, cb=0.2, cr=0.45, hdr=True, hdr_value=2.0, multiplier=2.0)
    o2 = orc.OracleScene(sd2)
    v = o2.btf_eval(1, np.zeros((1, 2), np.float32), np.array([[0, -1, 0]], np.float32), np.array([[0, 1, 0]], np.float32),
                    np.array([[0, 1, 0]], np.float32), np.array([[1, 0, 0]], np.float32))
    assert np.isfinite(v).all() and (v >= 0).all() and v.max() > 0
