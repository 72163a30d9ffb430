"""Generates the committed fixtures in tests/golden/.  Run in the authoring container:
    make -C oracle ref && python tests/golden/make_golden.py

* lcg_kat.json   — outputs of the REFERENCE's own RNG header
                   (/root/reference/include/RayTracerFacility/RayTracer/LinearCongruenceGenerator.hpp),
                   compiled verbatim into oracle/_ref/liblcg_ref.so.  Pins orc::LCG (oracle_math.h).
* oracle_small.npz — outputs of the CPU oracle on small seeded inputs (hits, visit counts, Morton codes,
                   irradiance, LiDAR, camera).  The reference has no tests/fixtures for these, so this
                   file guards the oracle against drift and gives the GPU tests stored vectors to hit as well.
* env_ref.npz    — outputs of the REFERENCE's own sampling helpers, sky model and environment lookup (RayTracer/
                   {RayTracerUtilities,Environment,BSDF}.cuh compiled unmodified into oracle/_ref/libenv_ref.so).  Pins
                   orc::reflect / tangent_space / sample_hemisphere / sample_lobe / cube_map_lookup / nishita_sky / env_light.
* programs_ref.npz — outputs of the REFERENCE's own device programs of the three entry points (ptx/{IlluminationEstimation,
                   PointCloudScanning,CameraRendering}.cu + RayFunctions.cuh + RayDataDefinations.hpp, compiled unmodified over a
                   host emulation of the OptiX intrinsics: oracle/ref_optix_driver.cpp), tracing through the oracle's BVH
                   and drawing from the reference's own LCG.  Pins orc::illum_probe / trace_path / camera_pixel /
                   camera_resolve / full_hit_info (run with the same LCG stream: orc_*_lcg).
* c1_ref.npz     — BASELINE config 1 (20k-triangle sorghum + ground, sky light, 64 samples per triangle and face, 4 bounces)
                   run through the REFERENCE's own illumination programs on the host (as programs_ref.npz) with two seeds:
                   the per-triangle energies of both runs.  Their difference is the Monte-Carlo noise of the reference's
                   own estimator, against which the GPU's (keyed Philox draws) is tested statistically.
* prebake_ref.npz — outputs of the REFERENCE's own geometry pre-kernels (CopyVerticesInstancedKernel RayTracer.cu:1148-1175,
                   CopySkinnedVerticesKernel :1194-1249, cut out of the reference file at build time and compiled unmodified
                   into oracle/_ref/libprebake_ref.so over the glm stand-in) for the seeded inputs of prebake_ref_cases().
                   Pins orc::bake_instanced / bake_skinned and, on the GPU, k_bake_instanced / k_bake_skinned.
* btf_ref.npz    — outputs of the REFERENCE's own compressed-BTF evaluator (include/RayTracerFacility/MLVQ/*.cuh compiled
                   unmodified into oracle/_ref/libbtf_ref.so against oracle/ref_stubs).  Pins orc::btf_value_deg.
* oracle_extra.npz — the same for the BSSRDF probe / irradiance, the instanced and skinned pre-kernels, light probes, flux.
"""
import ctypes as C
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import pyoracle as orc  # noqa: E402
from raytracerfacility_b200 import scenes, structs as S  # noqa: E402

LCG_SEEDS = [(0, 0), (1, 0), (0, 1), (12345, 678), (20001, 0x5EED), (0xFFFFFFFF, 0xFFFFFFFF), (1920 * 1080 - 1, 7)]


def small_cases():
    """The seeded inputs shared by make_golden.py and the tests."""
    soup = scenes.random_soup(2000, seed=7)
    lo, hi = scenes.scene_bounds(soup)
    rays = scenes.incoherent_rays(4096, lo - 0.2, hi + 0.2, seed=3)
    illum_scene = scenes.random_soup(300, seed=11, extent=1.0, size=0.35)
    env = S.EnvironmentProperties()
    rp = S.RayProperties(bounces=3, samples=16)
    grid = scenes.lidar_grid(center=(0.5, 3.0, 0.5), size=(1.6, 0.8), distance=(0.04, 0.04))
    return dict(soup=soup, rays=rays, illum_scene=illum_scene, env=env, rp=rp, grid=grid)


def extra_cases():
    """Seeded inputs of oracle_extra.npz: BSSRDF scene + probe rays, instanced / skinned bake, flux."""
    sss_scene = scenes.random_soup(300, seed=11, extent=1.0, size=0.35)
    m = sss_scene.materials[1]
    m.subsurface_factor = 0.5
    m.subsurface_color[:] = (0.9, 0.4, 0.2)
    m.subsurface_radius[:] = (0.08, 0.0, 0.0)
    lo, hi = scenes.scene_bounds(sss_scene)
    probe_rays = scenes.incoherent_rays(512, lo, hi, seed=13)
    probe_rays[:, 3] = 0.01
    probe_rays[:, 7] = 0.6
    mesh = scenes.random_soup(24, seed=17, size=0.4)
    v, t = mesh.geometries[1]
    v = v.copy()
    v["tangent"] = np.roll(v["normal"], 1, axis=1) + np.float32(0.25)
    k = np.arange(5, dtype=np.float64)
    mats = np.zeros((5, 4, 4))
    c, sn, sc = np.cos(0.7 * k), np.sin(0.7 * k), 0.6 + 0.2 * k
    mats[:, 0, 0], mats[:, 0, 2], mats[:, 2, 0], mats[:, 2, 2] = c * sc, sn * sc, -sn * sc, c * sc
    mats[:, 1, 1], mats[:, 3, 3] = sc, 1.0
    mats[:, 0, 3], mats[:, 1, 3], mats[:, 2, 3] = 0.3 * k, -0.1 * k, 0.05 * k * k
    mats = np.ascontiguousarray(mats.transpose(0, 2, 1)).astype(np.float32).reshape(5, 16)  # column-major
    cols = scenes.hash_uniform(19, np.arange(20), 1).reshape(5, 4).astype(np.float32)
    sv = np.zeros(len(v), dtype=S.SKINNED_VERTEX_DTYPE)
    sv["v"] = v
    i = np.arange(len(v))
    sv["bone_id"], sv["bone_id2"] = -1, -1
    sv["bone_id"][:, 0] = i % 5
    sv["bone_id"][:, 2] = np.where(i % 2 == 0, (i + 1) % 5, -1)
    sv["bone_id2"][:, 1] = np.where(i % 3 == 0, (i + 2) % 5, -1)
    w = scenes.hash_uniform(23, np.arange(len(v) * 8), 2).reshape(-1, 8).astype(np.float32)
    sv["weight"], sv["weight2"] = w[:, :4], w[:, 4:]
    return dict(sss_scene=sss_scene, probe_rays=probe_rays, handle=101, verts=v, tris=t, mats=mats, cols=cols, skinned=sv,
                env=S.EnvironmentProperties(color=(0.9, 1.0, 0.8)), rp=S.RayProperties(bounces=3, samples=16))


BTF_REF_VARIANTS = [dict(), dict(use_cos_beta=0), dict(hdr=1, hdr_value=2.5, multiplier=2.5), dict(code_btf_flag=1),
                    dict(gamma=1.0, tex_coord_multiplier=2.5, rows_offset=3, cols_offset=7)]


def _rigid(seed, k, scale=True):
    """Column-major 4x4: rotation about a hashed axis, hashed (non-uniform when `scale`) scale and translation."""
    h = lambda d: float(scenes.hash_uniform(seed, np.array([k]), d)[0])  # noqa: E731
    ax = np.array([h(0) - 0.5, h(1) - 0.5, h(2) - 0.5])
    ax /= np.linalg.norm(ax)
    a = 6.0 * h(3)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    R = np.eye(3) + np.sin(a) * K + (1 - np.cos(a)) * (K @ K)
    sc = np.diag([0.5 + h(4), 0.5 + h(5), 0.5 + h(6)]) if scale else np.eye(3)
    m = np.eye(4)
    m[:3, :3] = R @ sc
    m[:3, 3] = [4 * h(7) - 2, 4 * h(8) - 2, 4 * h(9) - 2]
    return m.T.astype(np.float32).reshape(16)


def prebake_ref_cases():
    """Seeded inputs of prebake_ref.npz: a 120-vertex soup mesh with non-trivial tangents, colours and paddings; 9 instance
    matrices (identity first, then rotations with non-uniform scale) with colours; skinned vertices with 1..8 influences
    (-1 = unused, in both quadruples), weights that sum to 1, and 6 bone matrices."""
    sd = scenes.random_soup(40, seed=23, size=0.5)
    v, t = sd.geometries[1]
    v = v.copy()
    n = len(v)
    idx = np.arange(n)
    v["tangent"] = np.stack([scenes.hash_uniform(31, idx, d) - 0.5 for d in range(3)], 1).astype(np.float32)
    v["color"] = np.stack([scenes.hash_uniform(32, idx, d) for d in range(4)], 1).astype(np.float32)
    v["tex_coord"] = np.stack([scenes.hash_uniform(33, idx, d) * 3 - 1 for d in range(2)], 1).astype(np.float32)
    v["position_padding"] = scenes.hash_uniform(34, idx, 0).astype(np.float32)
    v["normal_padding"] = scenes.hash_uniform(34, idx, 1).astype(np.float32)
    v["tangent_padding"] = scenes.hash_uniform(34, idx, 2).astype(np.float32)
    v["tex_coord_padding"] = np.stack([scenes.hash_uniform(34, idx, 3 + d) for d in range(2)], 1).astype(np.float32)
    mats = np.stack([np.eye(4, dtype=np.float32).reshape(16)] + [_rigid(41, k) for k in range(8)])
    cols = np.stack([scenes.hash_uniform(42, np.arange(9), d) for d in range(4)], 1).astype(np.float32)
    bones = np.stack([_rigid(43, k, scale=(k % 2 == 1)) for k in range(6)])
    sv = np.zeros(n, dtype=S.SKINNED_VERTEX_DTYPE)
    sv["v"] = v
    ids = np.full((n, 8), -1, dtype=np.int32)
    w = np.zeros((n, 8), dtype=np.float32)
    for i in range(n):
        k = 1 + i % 8                                  # number of influences
        slots = [0] + sorted(((np.arange(1, 8) * (i + 3)) % 7 + 1)[:k - 1].tolist())  # slot 0 always (the kernel reads it unconditionally)
        slots = sorted(set(slots))
        raw = np.array([scenes.hash_uniform(44, np.array([i]), s)[0] + 0.05 for s in slots])
        raw /= raw.sum()
        for s_, r_ in zip(slots, raw):
            ids[i, s_] = int(scenes.hash_uniform(45, np.array([i]), s_)[0] * 6) % 6
            w[i, s_] = np.float32(r_)
    sv["bone_id"], sv["bone_id2"] = ids[:, :4], ids[:, 4:]
    sv["weight"], sv["weight2"] = w[:, :4], w[:, 4:]
    return dict(vertices=v, triangles=t, matrices=mats, colors=cols, skinned=sv, bones=bones)


def prebake_ref_run(c):
    """The reference's two kernels on prebake_ref_cases() through oracle/_ref/libprebake_ref.so."""
    ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libprebake_ref.so"))
    v, mats, cols, sv, bones = c["vertices"], c["matrices"], c["colors"], c["skinned"], c["bones"]
    ptr = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    n, ni = len(v), len(mats)
    inst_v = np.zeros(n * ni, dtype=S.VERTEX_DTYPE)
    inst_p = np.zeros((n * ni, 3), dtype=np.float32)
    ref.ref_copy_vertices_instanced(C.c_int(ni), C.c_int(n), ptr(mats), ptr(cols), ptr(v), ptr(inst_p), ptr(inst_v))
    skin_v = np.zeros(n, dtype=S.VERTEX_DTYPE)
    skin_p = np.zeros((n, 3), dtype=np.float32)
    ref.ref_copy_skinned_vertices(C.c_int(n), ptr(sv), ptr(bones), ptr(skin_p), ptr(skin_v))
    return dict(instanced_vertices=inst_v.view(np.uint8), instanced_positions=inst_p.view(np.uint32),
                skinned_vertices=skin_v.view(np.uint8), skinned_positions=skin_p.view(np.uint32))


def write_prebake_ref():
    np.savez_compressed(os.path.join(HERE, "prebake_ref.npz"), **prebake_ref_run(prebake_ref_cases()))


def btf_ref_cases():
    """Seeded inputs of btf_ref.npz: small synthetic MLVQ databases (one per parameter variant) and GetValueDeg queries
    {u, v, illumTheta, illumPhi, viewTheta, viewPhi}; the first half stays below 80 degrees, the second half covers the
    whole range including theta > 90 (black) and grazing angles."""
    n = 4096
    idx = np.arange(n)
    h = lambda d: scenes.hash_uniform(31, idx, d)  # noqa: E731
    tmax = np.where(idx < n // 2, 80.0, 95.0)
    q = np.stack([h(0) * 3 - 1, h(1) * 3 - 1, h(2) * tmax, h(3) * 360, h(4) * tmax, h(5) * 360], 1).astype(np.float32)
    small = dict(seed=6, texels=(32, 32), n_pdf1d=256, n_colors=64, n_index=128, n_pdf2d_color=64, n_pdf2d_lum=128,
                 n_pdf2d=256, n_pdf3d=128, n_pdf4d=64)
    return [scenes.synthetic_btf(**small, **kw) for kw in BTF_REF_VARIANTS], q


def env_ref_cases():
    """Seeded inputs of env_ref.npz.  `vec`: n unit normals (the first rows hit GetTangentSpace's |n.x| > 0.99 switch from
    both sides), unit incident directions, unnormalised cube-map directions (with exact ties between the axes), LCG seeds
    and lobe widths; `sky`: Skydome environments (sun, intensity, gamma, sample counts) with positions and directions;
    `flat`: Scene / SingleLightSource environments."""
    n = 1024
    idx = np.arange(n)
    h = lambda s, d: scenes.hash_uniform(s, idx, d).astype(np.float64)  # noqa: E731

    def sphere(s):
        z, ph = 2 * h(s, 0) - 1, 2 * np.pi * h(s, 1)
        r = np.sqrt(1 - z * z)
        return np.stack([r * np.cos(ph), r * np.sin(ph), z], 1).astype(np.float32)

    normal, incident = sphere(41), sphere(42)
    normal[:6] = [[1, 0, 0], [-1, 0, 0], [0.9900001, 0.14106695, 0], [0.99, 0.14106736, 0], [0, 1, 0], [0, 0, -1]]
    cube = ((np.stack([h(43, 0), h(43, 1), h(43, 2)], 1) * 2 - 1) * 3).astype(np.float32)
    cube[:8] = [[1, 1, 1], [-1, 1, 1], [1, -1, -1], [2, 2, 1], [2, 1, 2], [1, 2, 2], [-2, 0, 2], [0, -3, 0]]
    seeds = np.stack([(h(44, 0) * 2 ** 31).astype(np.uint32), (h(44, 1) * 2 ** 31).astype(np.uint32)], 1)
    alpha = h(45, 0).astype(np.float32)
    alpha[:3] = [0.0, 1.0, 0.5]
    vec = dict(normal=normal, incident=incident, cube=cube, seeds=seeds, alpha=alpha)

    sun = sphere(46) * np.float32([1, 0.5, 1]) + np.float32([0, 0.6, 0])
    sun /= np.linalg.norm(sun, axis=1, keepdims=True)
    sun[0] = [0, 1, 0]
    pos = np.stack([(h(47, 0) - 0.5) * 400, h(47, 1) * 120, (h(47, 2) - 0.5) * 400], 1).astype(np.float32)
    sky = dict(sun=sun.astype(np.float32), dir=sphere(48), pos=pos, intensity=(0.5 + 2.5 * h(49, 0)).astype(np.float32),
               gamma=np.float32([1.0, 2.2, 0.7])[idx % 3], small=(idx % 7 == 0))
    flat = [dict(lighting_type=lt, color=(0.3, 0.9, 1.7), skylight_intensity=1.3, gamma=g)
            for lt in (S.ENV_SCENE, S.ENV_SINGLE_LIGHT_SOURCE) for g in (1.0, 2.2)]
    return vec, sky, flat


def sky_env(sky, k):
    env = S.EnvironmentProperties(lighting_type=S.ENV_SKYDOME, sun_direction=tuple(float(x) for x in sky["sun"][k]),
                                  skylight_intensity=float(sky["intensity"][k]), gamma=float(sky["gamma"][k]))
    if sky["small"][k]:
        env.atmosphere.num_samples, env.atmosphere.num_samples_light = 8, 4
    return env


def programs_ref_scenes():
    """Scenes of programs_ref.npz: `plants` (Default materials, every instance under its own transform), `mixed`
    (materials_scene without the BTF quad: textures, normal / metallic / roughness maps, vertex colours, alpha-textured
    canopy, emission, a transmissive material, cube map; identity transforms), `btf` (materials_scene as it is) and `sss`
    (below)."""
    mixed = scenes.materials_scene()
    mixed.materials[3] = scenes.Material(albedo=(0.5, 0.4, 0.9), metallic=0.6, roughness=0.3, transmission=0.4)
    del mixed.material_btf[3]
    # `sss`: `mixed` with a subsurface ground (one flat sheet per handle, so the BSSRDF's spatial probe can only ever
    # find the sheet itself — at most one record, the case in which the repo's 4-nearest probe and the reference's
    # traversal-order collection coincide): BSSRDF sampling arithmetic, normal map at the exit point, draw order
    sss = scenes.materials_scene()
    sss.materials[3] = mixed.materials[3]
    del sss.material_btf[3]
    g = sss.materials[1]
    g.subsurface_factor = 0.5
    g.subsurface_color[:] = (0.9, 0.5, 0.3)
    g.subsurface_radius[:] = (0.05, 0.0, 0.0)
    return dict(plants=scenes.sorghum_scene(n_plants=2, seed=2), mixed=mixed, btf=scenes.materials_scene(), sss=sss)


PROGRAMS_ENVS = dict(
    scene=dict(),
    skydome=dict(lighting_type=S.ENV_SKYDOME, sun_direction=(0.3, 0.8, 0.52)),
    sun=dict(lighting_type=S.ENV_SINGLE_LIGHT_SOURCE, sun_direction=(0.3, 0.8, 0.52), light_size=0.1, color=(1.0, 0.9, 0.7)),
    cube=dict(environmental_map_id=3, skylight_intensity=1.7, gamma=2.2))
# (scene, environment, bounces, samples) of the illumination-estimation cases; seed 7, push 1e-3, first 400 probes
PROGRAMS_IE = [(sc, e, b, n) for sc, envs in (("plants", ("scene", "skydome", "sun")), ("mixed", ("scene", "skydome", "sun", "cube")),
                                               ("btf", ("scene", "sun")), ("sss", ("scene", "skydome", "cube")))
               for e in envs for b, n in ((1, 1), (4, 8))]
# (environment, bounces, samples, output type, accumulate) of the camera cases on `mixed`, 32 x 24, aperture 0.02
PROGRAMS_CR = [("scene", 1, 1, S.OUT_COLOR, 0), ("scene", 3, 4, S.OUT_DEPTH, 0), ("skydome", 1, 1, S.OUT_COLOR, 0),
               ("skydome", 3, 4, S.OUT_COLOR, 1), ("cube", 2, 2, S.OUT_COLOR, 0)]


def programs_camera(accumulate):
    cam = camera_props(32, 24)
    cam.aperture, cam.focal_length = 0.02, 2.5
    if accumulate:
        cam.accumulate, cam.frame_id = 1, 3
    return cam


def programs_prev_color():
    return scenes.hash_uniform(77, np.arange(32 * 24 * 4), 0).astype(np.float32).reshape(-1, 4)


def programs_pc_samples(sd, n=2000):
    lo, hi = scenes.scene_bounds(sd)
    idx = np.arange(n)
    h = lambda d: scenes.hash_uniform(55, idx, d)  # noqa: E731
    o = np.stack([lo[k] - 0.2 + (hi[k] - lo[k] + 0.4) * h(k) for k in range(3)], 1)
    t = np.stack([lo[k] + (hi[k] - lo[k]) * h(3 + k) for k in range(3)], 1)
    d = t - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    smp = np.zeros(n, dtype=S.PC_SAMPLE_DTYPE)
    smp["start"], smp["direction"] = o.astype(np.float32), d.astype(np.float32)
    return smp


def camera_props(width=16, height=12):
    """CameraProperties::Set (RayTracer.cu:28-46) for position (0.5,0.5,3), identity rotation, fov 120 (60 effective)."""
    import math
    cp = S.CameraProps()
    cp.accumulate, cp.fov, cp.max_distance, cp.gamma = 0, 120.0, 50.0, 2.2
    cp.width, cp.height, cp.frame_id, cp.aperture, cp.focal_length = width, height, 0, 0.0, 1.0
    cp.output_type = S.OUT_COLOR
    pos = np.array([0.5, 0.5, 3.0])
    front, up = np.array([0.0, 0, -1]), np.array([0.0, 1, 0])
    aspect = width / height
    half = math.radians(cp.fov * 0.5)
    th = math.tan(half / 2)
    zn, zf = 0.1, 100.0
    proj = np.zeros((4, 4))
    proj[0, 0], proj[1, 1], proj[2, 2], proj[3, 2], proj[2, 3] = 1 / (aspect * th), 1 / th, -(zf + zn) / (zf - zn), -1, -2 * zf * zn / (zf - zn)
    s = np.cross(front, up)
    s /= np.linalg.norm(s)
    u = np.cross(s, front)
    view = np.eye(4)
    view[0, :3], view[1, :3], view[2, :3] = s, u, -front
    view[0, 3], view[1, 3], view[2, 3] = -s @ pos, -u @ pos, front @ pos
    inv = np.linalg.inv(proj @ view)
    cp.inverse_projection_view[:] = [float(x) for x in inv.T.reshape(16)]  # column-major
    cp.from_[:] = [float(x) for x in pos]
    cp.horizontal[:] = [float(x) for x in (half * aspect * s)]
    cp.vertical[:] = [float(x) for x in (half * up)]
    return cp


def main():
    ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "liblcg_ref.so"))
    kat = []
    for v0, v1 in LCG_SEEDS:
        out = (C.c_float * 8)()
        st = C.c_uint32()
        ref.ref_lcg(C.c_uint32(v0), C.c_uint32(v1), 8, out, C.byref(st))
        kat.append(dict(v0=v0, v1=v1, state=st.value,
                        out_bits=[int(np.float32(x).view(np.uint32)) for x in out]))
    with open(os.path.join(HERE, "lcg_kat.json"), "w") as f:
        json.dump(dict(source="reference LinearCongruenceGenerator.hpp via oracle/_ref/liblcg_ref.so", kat=kat), f, indent=1)

    c = small_cases()
    osc = orc.OracleScene(c["soup"], 63, presplit=1)
    hits, nv, tv = osc.trace(c["rays"], counts=True)
    morton = osc.debug(S.DBG_MORTON)
    order = osc.debug(S.DBG_SORTED_IDS)
    osc30 = orc.OracleScene(c["soup"], 30, presplit=1)
    ill = orc.OracleScene(c["illum_scene"], 63, presplit=1)
    irr, st = ill.illum_mesh(c["env"], c["rp"], seed=0x5EED, push=1e-3)
    t, prim, inst, _ = ill.scan_grid(c["grid"])
    cp = camera_props()
    col, nrm, alb, _ = ill.camera(c["env"], S.RayProperties(bounces=2, samples=4), cp)
    np.savez_compressed(os.path.join(HERE, "oracle_small.npz"), hits=hits.view(np.uint32).reshape(-1, 4), node_visits=nv,
                        tri_visits=tv, morton63=morton, order63=order, morton30=osc30.debug(S.DBG_MORTON),
                        order30=osc30.debug(S.DBG_SORTED_IDS), n_nodes=np.array([osc.n_nodes, osc.depth]),
                        irradiance=irr.view(np.uint32), illum_rays=np.array([st.rays, st.paths]),
                        lidar_t=t.view(np.uint32), lidar_prim=prim, camera_color=col.view(np.uint32),
                        camera_normal=nrm.view(np.uint32), camera_albedo=alb.view(np.uint32))
    write_extra()
    write_btf_ref()
    write_prebake_ref()
    write_env_ref()
    write_programs_ref()
    write_c1_ref()
    print("wrote", os.listdir(HERE))


def write_env_ref():
    """env_ref.npz — outputs of the REFERENCE's own helpers (oracle/ref_env_driver.cpp -> oracle/_ref/libenv_ref.so,
    `make -C oracle ref`) for env_ref_cases().  Floats are stored as their bit patterns."""
    R = orc.RefEnv()
    vec, sky, flat = env_ref_cases()
    n = len(vec["normal"])
    u32 = lambda rows: np.asarray(rows, dtype=np.float32).view(np.uint32)  # noqa: E731
    out = dict(reflect=u32([R.reflect(vec["incident"][k], vec["normal"][k]) for k in range(n)]),
               tangent=u32([R.tangent_space(vec["normal"][k]).reshape(9) for k in range(n)]))
    hemi = [R.sample_hemisphere(int(a), int(b), vec["normal"][k]) for k, (a, b) in enumerate(vec["seeds"])]
    lobe = [R.sample_lobe(int(a), int(b), vec["normal"][k], float(vec["alpha"][k])) for k, (a, b) in enumerate(vec["seeds"])]
    brdf = [R.brdf(float(vec["alpha"][k]), int(a), int(b), vec["incident"][k], vec["normal"][k])
            for k, (a, b) in enumerate(vec["seeds"])]
    out["uniforms"] = u32([r for r, _ in hemi])
    assert all(np.array_equal(r, r2) and np.array_equal(r, r3) for (r, _), (r2, _), (r3, _) in zip(hemi, lobe, brdf))
    out["hemisphere"], out["lobe"], out["brdf"] = (u32([d for _, d in x]) for x in (hemi, lobe, brdf))
    cube = [R.cube_map_lookup(d) for d in vec["cube"]]
    out["cube_face"] = np.array([f for f, _ in cube], dtype=np.int32)
    out["cube_uv"] = u32([uv for _, uv in cube])
    out["sky"] = u32([R.env_light(sky_env(sky, k), sky["pos"][k], sky["dir"][k]) for k in range(n)])
    out["sky_raw"] = u32([R.nishita_sky(sky_env(sky, k), sky["pos"][k], sky["dir"][k]) for k in range(n)])
    out["flat"] = u32([R.env_light(S.EnvironmentProperties(**kw), (0, 0, 0), (0, 1, 0)) for kw in flat])
    # the inputs travel with the outputs (the tests read them from here, not from env_ref_cases())
    out.update({f"in_{k}": v for k, v in vec.items()})
    out.update({f"in_sky_{k}": v for k, v in sky.items()})
    np.savez_compressed(os.path.join(HERE, "env_ref.npz"), **out)


def write_programs_ref():
    """programs_ref.npz — see the module docstring; needs oracle/_ref/lib{ie,pcs,cr}_ref.so (`make -C oracle ref`)."""
    out = {}
    oscs = {k: orc.OracleScene(sd) for k, sd in programs_ref_scenes().items()}
    for sc, e, b, n in PROGRAMS_IE:
        osc = oscs[sc]
        pr = osc.make_probes(0, min(400, osc.n_tris))
        rays = osc.ref_illum_probes(S.EnvironmentProperties(**PROGRAMS_ENVS[e]), S.RayProperties(b, n), pr, 7, 1e-3)
        key = f"ie_{sc}_{e}_{b}_{n}"
        out[key] = np.concatenate([pr["energy"], pr["direction"]], 1).view(np.uint32)
        out[key + "_rays"] = np.array([rays])
    for k, (e, b, n, ot, acc) in enumerate(PROGRAMS_CR):
        cam = programs_camera(acc)
        cam.output_type = ot
        c, nm, al, rays = oscs["mixed"].ref_camera(S.EnvironmentProperties(**PROGRAMS_ENVS[e]), S.RayProperties(b, n), cam,
                                                   programs_prev_color() if acc else None)
        out[f"cr_{k}"] = np.concatenate([c, nm, al], 1).view(np.uint32)
        out[f"cr_{k}_rays"] = np.array([rays])
    for sc in ("plants", "mixed"):
        smp = programs_pc_samples(programs_ref_scenes()[sc])
        out[f"pcs_{sc}_in"] = smp.copy().view(np.uint8)
        oscs[sc].ref_scan_point_cloud(smp)
        out[f"pcs_{sc}"] = smp.view(np.uint8)
    np.savez_compressed(os.path.join(HERE, "programs_ref.npz"), **out)


C1_SEEDS = (11, 12)


def write_c1_ref():
    """c1_ref.npz — see the module docstring; half a minute of single-thread CPU."""
    osc = orc.OracleScene(scenes.sorghum_scene(1))
    env, rp = S.EnvironmentProperties(), S.RayProperties(4, 64)
    out = {}
    for name, seed in zip("ab", C1_SEEDS):
        pr = osc.make_probes()
        out["rays_" + name] = np.array([osc.ref_illum_probes(env, rp, pr, seed, 1e-3)])
        out["energy_" + name] = pr["energy"].copy()
    np.savez_compressed(os.path.join(HERE, "c1_ref.npz"), **out)


def write_btf_ref():
    """btf_ref.npz — outputs of the REFERENCE's own compressed-BTF evaluator (MLVQ/*.cuh compiled unmodified into
    oracle/_ref/libbtf_ref.so, `make -C oracle ref`) for the queries of btf_ref_cases()."""
    btfs, q = btf_ref_cases()
    out = {}
    for k, btf in enumerate(btfs):
        osc = orc.OracleScene(scenes.materials_scene(btf=btf))
        beta = osc.btf_beta_angles(1, btf.desc().num_beta)
        out[f"rgb{k}"] = orc.ref_btf_value_deg(btf, beta, q).view(np.uint32)
    np.savez_compressed(os.path.join(HERE, "btf_ref.npz"), **out)


def write_extra():
    """oracle_extra.npz — BSSRDF (probe records, irradiance), pre-kernels, light probes and flux from the CPU oracle."""
    x = extra_cases()
    osc = orc.OracleScene(x["sss_scene"], 63)
    rec = osc.sss_probe(x["probe_rays"], np.full(len(x["probe_rays"]), x["handle"], dtype=np.uint64))
    irr, st = osc.illum_mesh(x["env"], x["rp"], seed=0xABCD, push=1e-3)
    bv, bt = orc.bake_instanced(x["verts"], x["tris"], x["mats"], x["cols"])
    sk = orc.bake_skinned(x["skinned"], x["mats"])
    pr = osc.make_probes(front=True, back=True)
    pr["energy"] = irr[:, :3]
    flux = orc.flux_probes(pr)
    np.savez_compressed(os.path.join(HERE, "oracle_extra.npz"), sss_records=rec, sss_irradiance=irr.view(np.uint32),
                        sss_rays=np.array([st.rays, st.paths]), bake_instanced=bv.view(np.uint8), bake_tris=bt,
                        bake_skinned=sk.view(np.uint8), probes=pr.view(np.uint8), flux=flux.view(np.uint32))


if __name__ == "__main__":
    main()
