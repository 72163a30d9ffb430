"""Pins the oracle's arithmetic primitives (CPU only).
* orc::LCG vs the reference's own header (golden vectors from oracle/_ref, tests/golden/lcg_kat.json).
* Philox-4x32-10 vs the Random123 known-answer vectors (kat_vectors of Random123 v1.14).
* sincos2pi vs double-precision libm; samplers (RayTracerUtilities.cuh:110-145) moments and frames.
"""
import json
import math
import os

import numpy as np

from .conftest import GOLDEN


def test_lcg_matches_reference_header(orc):
    kat = json.load(open(os.path.join(GOLDEN, "lcg_kat.json")))["kat"]
    assert len(kat) >= 5
    for k in kat:
        out, state = orc.lcg(k["v0"], k["v1"], 8)
        assert state == k["state"]
        assert [int(x) for x in out.view(np.uint32)] == k["out_bits"]


def test_lcg_output_range(orc):
    out, _ = orc.lcg(42, 7, 4096)
    assert out.min() >= 0.0 and out.max() < 1.0
    assert abs(out.mean() - 0.5) < 0.03


def test_philox_known_answers(orc):
    # Random123 kat_vectors: philox4x32 10 rounds
    assert orc.philox([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert orc.philox([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert orc.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_path_rng_is_counter_based(orc):
    a = orc.path_rng(0x5EED, 0, 123456789012, 2, 0)
    b = orc.path_rng(0x5EED, 0, 123456789012, 2, 0)
    c = orc.path_rng(0x5EED, 0, 123456789013, 2, 0)
    assert a == b and a != c
    assert all(0.0 <= x < 1.0 for x in a)
    # 24-bit granularity like LinearCongruenceGenerator.hpp:32
    assert all(float(x) * 2 ** 24 == int(float(x) * 2 ** 24) for x in a)


def test_sincos2pi_accuracy(orc):
    u = np.concatenate([np.linspace(0, 1, 20001, dtype=np.float64), [0.0, 0.25, 0.5, 0.75, 1.0]]).astype(np.float32)
    s, c = orc.sincos2pi(u)
    assert np.max(np.abs(s - np.sin(2 * np.pi * u.astype(np.float64)))) < 3e-7
    assert np.max(np.abs(c - np.cos(2 * np.pi * u.astype(np.float64)))) < 3e-7
    s0, c0 = orc.sincos2pi(np.array([0.0, 0.25, 0.5, 0.75], dtype=np.float32))
    assert list(s0) == [0.0, 1.0, 0.0, -1.0] and list(c0) == [1.0, 0.0, -1.0, 0.0]


def test_hemisphere_sampler_frame_and_moments(orc):
    n = np.array([0.2, 0.9, -0.3], dtype=np.float32)
    n /= np.linalg.norm(n)
    rng = np.random.default_rng(1)
    cos_sum = 0.0
    N = 4000
    for r0, r1 in rng.random((N, 2)).astype(np.float32):
        d = orc.sample_hemisphere(float(r0), float(r1), n)
        assert abs(np.linalg.norm(d) - 1.0) < 1e-5
        ct = float(d @ n)
        assert ct >= -1e-6
        assert abs(ct - float(r0)) < 1e-5  # cosTheta = rnd (uniform in cos), RayTracerUtilities.cuh:138
        cos_sum += ct
    assert abs(cos_sum / N - 0.5) < 0.02


def test_lobe_sampler(orc):
    n = np.array([0.0, 1.0, 0.0], dtype=np.float32)
    d = orc.sample_lobe(0.7, 0.3, n, 1.0)  # alpha = 1 -> cosTheta = 1 -> the axis itself
    assert np.allclose(d, n, atol=1e-6)
    d = orc.sample_lobe(0.5, 0.1, n, 0.0)  # alpha = 0 -> cosTheta = 1 - rnd
    assert abs(float(d @ n) - 0.5) < 1e-6
    # helper axis switches to Z when |n.x| > 0.99 (RayTracerUtilities.cuh:112-114)
    d = orc.sample_hemisphere(0.25, 0.0, np.array([1.0, 0.0, 0.0], dtype=np.float32))
    assert abs(float(d[0]) - 0.25) < 1e-6


def test_unnormalised_normal_scales_direction(orc):
    # IlluminationEstimation.cu:60-62 passes an un-normalised normal: the n-component scales, T and B do not.
    n = np.array([0.0, 2.0, 0.0], dtype=np.float32)
    d = orc.sample_hemisphere(0.5, 0.0, n)
    assert abs(float(d[1]) - 1.0) < 1e-6 and abs(np.linalg.norm(d[[0, 2]]) - math.sqrt(0.75)) < 1e-6


def test_env_light_modes(orc, S):
    e = S.EnvironmentProperties(color=(0.5, 0.25, 1.0), skylight_intensity=2.0)
    assert np.allclose(orc.env_light(e, (0, 0, 0), (0, 1, 0)), [1.0, 0.5, 2.0])
    e.gamma = 2.0
    assert np.allclose(orc.env_light(e, (0, 0, 0), (0, 1, 0)), np.sqrt([1.0, 0.5, 2.0]), rtol=1e-6)
    sky = S.EnvironmentProperties(lighting_type=S.ENV_SKYDOME)
    up = orc.env_light(sky, (0, 0, 0), (0, 1, 0))
    horizon = orc.env_light(sky, (0, 0, 0), (1, 0.02, 0))
    assert np.all(up > 0) and np.all(np.isfinite(up)) and up[2] > up[0]  # bluish zenith
    assert horizon[0] / horizon[2] > up[0] / up[2]  # longer optical path at the horizon scatters the blue out
