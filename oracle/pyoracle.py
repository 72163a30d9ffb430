"""ctypes binding of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Nothing under raytracerfacility_b200/ imports this module.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
from raytracerfacility_b200 import structs as S  # noqa: E402  (POD layouts only)


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("oracle_api.cpp", "oracle_sss.h", "oracle_prebake.h", "oracle_shade.h", "oracle_bvh.h", "oracle_btf.h", "oracle_math.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        L = _lib
        L.orc_version.restype = C.c_char_p
        L.orc_scene_create.restype = C.c_void_p
        L.orc_scene_ntris.restype = C.c_uint64
        L.orc_scene_nnodes.restype = C.c_uint64
        L.orc_debug_size.restype = C.c_uint64
        L.orc_scan_grid_count.restype = C.c_uint64
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def philox(ctr, key):
    out = (C.c_uint32 * 4)()
    lib().orc_philox((C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), out)
    return list(out)


def lcg(v0, v1, n):
    out = np.zeros(n, dtype=np.float32)
    st = C.c_uint32()
    lib().orc_lcg(C.c_uint32(v0), C.c_uint32(v1), n, _p(out), C.byref(st))
    return out, st.value


def sincos2pi(u):
    u = np.ascontiguousarray(u, dtype=np.float32)
    s, c = np.empty_like(u), np.empty_like(u)
    lib().orc_sincos2pi(len(u), _p(u), _p(s), _p(c))
    return s, c


def path_rng(seed, stream, unit, bounce, extra=0):
    r = (C.c_float * 4)()
    lib().orc_path_rng(C.c_uint32(seed), C.c_uint32(stream), C.c_uint64(unit), C.c_uint32(bounce), C.c_uint32(extra), r)
    return list(r)


def sample_hemisphere(r0, r1, n):
    out = (C.c_float * 3)()
    lib().orc_sample_hemisphere(C.c_float(r0), C.c_float(r1), (C.c_float * 3)(*n), out)
    return np.array(out[:], dtype=np.float32)


def sample_lobe(r0, r1, n, alpha):
    out = (C.c_float * 3)()
    lib().orc_sample_lobe(C.c_float(r0), C.c_float(r1), (C.c_float * 3)(*n), C.c_float(alpha), out)
    return np.array(out[:], dtype=np.float32)


def sample_lambert_ggx(albedo, metallic, roughness, normal, view, r3):
    """(weights, directions) of oracle_shade.h's sample_lambert_ggx for n triples of uniforms."""
    r3 = np.ascontiguousarray(r3, dtype=np.float32).reshape(-1, 3)
    a, nrm, v = (np.ascontiguousarray(x, dtype=np.float32) for x in (albedo, normal, view))
    w, d = np.empty_like(r3), np.empty_like(r3)
    lib().orc_sample_lambert_ggx(len(r3), _p(a), C.c_float(metallic), C.c_float(roughness), _p(nrm), _p(v), _p(r3), _p(w), _p(d))
    return w, d


def inverse_trig(y, x):
    """Portable atan2(y, x), asin(y), acos(y) of oracle_math.h."""
    y, x = np.ascontiguousarray(y, dtype=np.float32), np.ascontiguousarray(x, dtype=np.float32)
    a, s, c = np.empty_like(y), np.empty_like(y), np.empty_like(y)
    lib().orc_inverse_trig(len(y), _p(y), _p(x), _p(a), _p(s), _p(c))
    return a, s, c


def env_light(env, pos, d):
    out = (C.c_float * 3)()
    lib().orc_env_light(C.byref(env), (C.c_float * 3)(*pos), (C.c_float * 3)(*d), out)
    return np.array(out[:], dtype=np.float32)


class OracleScene:
    """CPU oracle scene built from a scenes.SceneDesc (instances in ascending handle order)."""

    def __init__(self, desc, morton_bits=63, presplit=S.DEFAULT_PRESPLIT):
        L = lib()
        self.h = C.c_void_p(L.orc_scene_create())
        gmap, mmap = {}, {}
        for gh in sorted(desc.geometries):
            v, t = desc.geometries[gh]
            v = np.ascontiguousarray(v)
            t = np.ascontiguousarray(t, dtype=np.uint32)
            gmap[gh] = L.orc_scene_add_geometry(self.h, _p(v), C.c_uint64(len(v)), C.c_uint64(80), _p(t),
                                                C.c_uint64(len(t)))
        for mh in sorted(desc.materials):
            mmap[mh] = L.orc_scene_add_material(self.h, C.byref(desc.materials[mh]))
        for ih in sorted(desc.instances):
            g, m, x, comp = desc.instances[ih]
            x = np.ascontiguousarray(x, dtype=np.float32)
            L.orc_scene_add_instance(self.h, gmap[g], mmap[m], _p(x), C.c_uint64(comp))
        # images, cube map, compressed BTF (oracle_btf.h)
        self.tex_index, self.btf_index, self.mat_index = {}, {}, mmap
        for th in sorted(getattr(desc, "textures", {})):
            img = np.ascontiguousarray(desc.textures[th], dtype=np.float32)
            self.tex_index[th] = L.orc_scene_add_texture(self.h, int(img.shape[1]), int(img.shape[0]), _p(img))
        for mh, tex in getattr(desc, "material_textures", {}).items():
            ids = [self.tex_index[t] if t else -1 for t in tex]
            L.orc_scene_set_material_textures(self.h, mmap[mh], *ids)
        for bh in sorted(getattr(desc, "btfs", {})):
            d = desc.btfs[bh].desc()
            self.btf_index[bh] = L.orc_scene_add_btf(self.h, C.byref(d))
        for mh, bh in getattr(desc, "material_btf", {}).items():
            L.orc_scene_set_material_btf(self.h, mmap[mh], self.btf_index[bh])
        if getattr(desc, "env_map", None) is not None:
            faces = (C.c_int * 6)(*[self.tex_index[t] for t in desc.env_map])
            L.orc_scene_set_environment_map(self.h, faces)
        L.orc_scene_commit(self.h, int(morton_bits), int(presplit))
        self.n_tris = int(L.orc_scene_ntris(self.h))
        self.n_nodes = int(L.orc_scene_nnodes(self.h))
        self.depth = int(L.orc_scene_depth(self.h))

    def __del__(self):
        try:
            if self.h:
                lib().orc_scene_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def set_brdf(self, brdf):
        lib().orc_scene_set_brdf(self.h, int(brdf))

    def sample_texture(self, handle, uv):
        uv = np.ascontiguousarray(uv, dtype=np.float32).reshape(-1, 2)
        out = np.empty((len(uv), 4), dtype=np.float32)
        lib().orc_sample_texture(self.h, self.tex_index[handle], C.c_uint64(len(uv)), _p(uv), _p(out))
        return out

    def sample_environment(self, env, pos_dir):
        pd = np.ascontiguousarray(pos_dir, dtype=np.float32).reshape(-1, 6)
        out = np.empty((len(pd), 3), dtype=np.float32)
        lib().orc_sample_environment(self.h, C.byref(env), C.c_uint64(len(pd)), _p(pd), _p(out))
        return out

    def btf_eval(self, btf, tex_coord, view_dir, illum_dir, normal, tangent, nthreads=0):
        a = [np.ascontiguousarray(x, dtype=np.float32) for x in (tex_coord, view_dir, illum_dir, normal, tangent)]
        n = len(a[0])
        out = np.empty((n, 3), dtype=np.float32)
        lib().orc_btf_eval(self.h, self.btf_index[btf], C.c_uint64(n), _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(a[4]),
                           _p(out), int(nthreads))
        return out

    def btf_value_deg(self, btf, q6):
        """BTFBase::GetValueDeg layer: rows {u, v, illumTheta, illumPhi, viewTheta, viewPhi} in degrees -> rgb."""
        q = np.ascontiguousarray(q6, dtype=np.float32).reshape(-1, 6)
        out = np.empty((len(q), 3), dtype=np.float32)
        lib().orc_btf_value_deg(self.h, self.btf_index[btf], C.c_uint64(len(q)), _p(q), _p(out))
        return out

    def btf_beta_angles(self, btf, n):
        out = np.empty(n, dtype=np.float32)
        lib().orc_btf_beta_angles(self.h, self.btf_index[btf], _p(out))
        return out

    def debug(self, what):
        L = lib()
        nbytes = int(L.orc_debug_size(self.h, what))
        dt = S.DBG_DTYPES[what]
        out = np.empty(nbytes // dt.itemsize, dtype=dt)
        rc = L.orc_debug_read(self.h, what, _p(out), C.c_uint64(nbytes))
        assert rc == 0
        return out

    def trace(self, rays, counts=False, nthreads=0):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = len(rays)
        hits = np.empty(n, dtype=S.HIT_DTYPE)
        nv = np.empty(n, dtype=np.uint32) if counts else None
        tv = np.empty(n, dtype=np.uint32) if counts else None
        lib().orc_trace(self.h, _p(rays), C.c_uint64(n), _p(hits), _p(nv), _p(tv), nthreads)
        return (hits, nv, tv) if counts else hits

    def trace_brute(self, rays, nthreads=0):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        hits = np.empty(len(rays), dtype=S.HIT_DTYPE)
        lib().orc_trace_brute(self.h, _p(rays), C.c_uint64(len(rays)), _p(hits), nthreads)
        return hits

    def illum_mesh(self, env, rp, seed, push, front=True, back=True, tri_begin=0, tri_end=None, tri_stride=1,
                   s0=0, s1=None, normalise=True, nthreads=0):
        tri_end = self.n_tris if tri_end is None else tri_end
        s1 = rp.samples if s1 is None else s1
        cnt = max(0, (tri_end - tri_begin + tri_stride - 1) // tri_stride)
        out = np.zeros((cnt, 6), dtype=np.float32)
        st = S.JobStats()
        lib().orc_illum_mesh(self.h, C.byref(env), C.byref(rp), C.c_uint32(seed), C.c_float(push), int(front),
                             int(back), C.c_uint64(tri_begin), C.c_uint64(tri_end), C.c_uint64(tri_stride), s0, s1,
                             int(normalise), _p(out), C.byref(st), nthreads)
        return out, st

    def make_probes(self, tri_begin=0, tri_end=None, front=True, back=True):
        tri_end = self.n_tris if tri_end is None else tri_end
        assert 0 <= tri_begin and tri_end <= self.n_tris
        out = np.zeros(max(0, tri_end - tri_begin), dtype=S.ILLUM_SAMPLER_DTYPE)
        lib().orc_make_probes(self.h, C.c_uint64(tri_begin), C.c_uint64(tri_end), int(front), int(back), _p(out))
        return out

    def flux_mesh(self, energy6, tri_begin=0, tri_end=None, front=True, back=True):
        tri_end = self.n_tris if tri_end is None else tri_end
        e = np.ascontiguousarray(energy6, dtype=np.float32)
        assert e.shape == (self.n_tris, 6)
        out = np.zeros(8, dtype=np.float32)
        lib().orc_flux_mesh(self.h, _p(e), C.c_uint64(tri_begin), C.c_uint64(tri_end), int(front) + int(back), _p(out))
        return out

    def sss_probe(self, rays, handles, brute=False):
        """BSSDF.cuh:18-89 spatial probe: per ray {n, prim[4], t[4], u[4], v[4]} (uint32 view)."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        handles = np.ascontiguousarray(handles, dtype=np.uint64)
        out = np.zeros((len(rays), 17), dtype=np.uint32)
        lib().orc_sss_probe(self.h, _p(rays), _p(handles), C.c_uint64(len(rays)), _p(out), int(brute))
        return out

    def illum_probes(self, env, rp, probes, seed, push, nthreads=0):
        assert probes.dtype == S.ILLUM_SAMPLER_DTYPE and probes.flags.c_contiguous
        st = S.JobStats()
        lib().orc_illum_probes(self.h, C.byref(env), C.byref(rp), _p(probes), C.c_uint64(len(probes)),
                               C.c_uint32(seed), C.c_float(push), C.byref(st), nthreads)
        return st

    def illum_probes_lcg(self, env, rp, probes, seed, push):
        """illum_probes() drawing from the reference's LCG in the reference's order (test mode, oracle_math.h
        t_seq_rng).  Returns the number of rays traced."""
        assert probes.dtype == S.ILLUM_SAMPLER_DTYPE and probes.flags.c_contiguous
        rays = C.c_uint64()
        rc = lib().orc_illum_probes_lcg(self.h, C.byref(env), C.byref(rp), _p(probes), C.c_uint64(len(probes)),
                                        C.c_uint32(seed), C.c_float(push), C.byref(rays))
        assert rc == 0, "BSSRDF materials are not supported in this mode"
        return rays.value

    def ref_illum_probes(self, env, rp, probes, seed, push):
        """The REFERENCE's own __raygen__IE / closest-hit / any-hit / miss programs run on the host over this scene
        (oracle/ref_ie_driver.cpp -> oracle/_ref/libie_ref.so).  Returns the number of rays traced."""
        L = ref_ie_lib()
        if L is None:
            raise FileNotFoundError("oracle/_ref/libie_ref.so (make -C oracle ref)")
        assert probes.dtype == S.ILLUM_SAMPLER_DTYPE and probes.flags.c_contiguous
        rays = C.c_uint64()
        rc = L.ref_ie_run(self.h, C.byref(env), C.byref(rp), _p(probes), C.c_uint64(len(probes)),
                          C.c_uint32(seed), C.c_float(push), C.byref(rays))
        assert rc == 0, "scene holds materials the host run of the reference does not set up"
        return rays.value

    def scan_point_cloud(self, samples, nthreads=0):
        assert samples.dtype == S.PC_SAMPLE_DTYPE and samples.flags.c_contiguous
        prim = np.empty(len(samples), dtype=np.uint32)
        lib().orc_scan_point_cloud(self.h, _p(samples), C.c_uint64(len(samples)), _p(prim), nthreads)
        return prim

    def scan_grid(self, grid, ray_begin=0, ray_end=None, stride=1, nthreads=0):
        total = int(lib().orc_scan_grid_count(C.byref(grid)))
        ray_end = total if ray_end is None else ray_end
        cnt = max(0, (ray_end - ray_begin + stride - 1) // stride)
        t = np.empty(cnt, dtype=np.float32)
        prim = np.empty(cnt, dtype=np.uint32)
        inst = np.empty(cnt, dtype=np.uint32)
        st = S.JobStats()
        lib().orc_scan_grid(self.h, C.byref(grid), C.c_uint64(ray_begin), C.c_uint64(ray_end), C.c_uint64(stride),
                            _p(t), _p(prim), _p(inst), C.byref(st), nthreads)
        return t, prim, inst, st

    def camera_lcg(self, env, rp, cam, prev_color=None):
        """camera() over the whole frame drawing from the reference's LCG in the reference's order (test mode)."""
        n = cam.width * cam.height
        color = np.zeros((n, 4), dtype=np.float32) if prev_color is None else np.array(prev_color, dtype=np.float32).reshape(n, 4)
        normal, albedo = np.empty((n, 4), dtype=np.float32), np.empty((n, 4), dtype=np.float32)
        rays = C.c_uint64()
        rc = lib().orc_camera_lcg(self.h, C.byref(env), C.byref(rp), C.byref(cam), _p(color), _p(normal), _p(albedo),
                                  C.byref(rays))
        assert rc == 0
        return color, normal, albedo, rays.value

    def ref_camera(self, env, rp, cam, prev_color=None):
        """The REFERENCE's own __raygen__CR / closest-hit / any-hit / miss programs run on the host over this scene."""
        L = ref_cr_lib()
        if L is None:
            raise FileNotFoundError("oracle/_ref/libcr_ref.so (make -C oracle ref)")
        n = cam.width * cam.height
        color = np.zeros((n, 4), dtype=np.float32) if prev_color is None else np.array(prev_color, dtype=np.float32).reshape(n, 4)
        normal, albedo = np.empty((n, 4), dtype=np.float32), np.empty((n, 4), dtype=np.float32)
        rays = C.c_uint64()
        rc = L.ref_cr_run(self.h, C.byref(env), C.byref(rp), C.byref(cam), _p(color), _p(normal), _p(albedo), C.byref(rays))
        assert rc == 0
        return color, normal, albedo, rays.value

    def ref_scan_point_cloud(self, samples):
        """The REFERENCE's own __raygen__PCS / closest-hit / miss programs run on the host over this scene (in place)."""
        L = ref_pcs_lib()
        if L is None:
            raise FileNotFoundError("oracle/_ref/libpcs_ref.so (make -C oracle ref)")
        assert samples.dtype == S.PC_SAMPLE_DTYPE and samples.flags.c_contiguous
        rc = L.ref_pcs_run(self.h, _p(samples), C.c_uint64(len(samples)))
        assert rc == 0

    def camera(self, env, rp, cam, pixel_begin=0, pixel_end=None, stride=1, prev_color=None, nthreads=0):
        pixel_end = cam.width * cam.height if pixel_end is None else pixel_end
        cnt = max(0, (pixel_end - pixel_begin + stride - 1) // stride)
        color = np.empty((cnt, 4), dtype=np.float32)
        normal = np.empty((cnt, 4), dtype=np.float32)
        albedo = np.empty((cnt, 4), dtype=np.float32)
        st = S.JobStats()
        lib().orc_camera(self.h, C.byref(env), C.byref(rp), C.byref(cam), C.c_uint64(pixel_begin),
                         C.c_uint64(pixel_end), C.c_uint64(stride), _p(prev_color), _p(color), _p(normal), _p(albedo),
                         C.byref(st), nthreads)
        return color, normal, albedo, st


def _ref_lib(name):
    so = os.path.join(_HERE, "_ref", name)
    return C.CDLL(so) if os.path.exists(so) else None


def ref_pcs_lib():
    """oracle/_ref/libpcs_ref.so: the reference's PointCloudScanning.cu programs on the host (see ref_ie_lib)."""
    return _ref_lib("libpcs_ref.so")


def ref_cr_lib():
    """oracle/_ref/libcr_ref.so: the reference's CameraRendering.cu programs on the host (see ref_ie_lib)."""
    return _ref_lib("libcr_ref.so")


def bake_instanced(vertices, triangles, matrices, colors=None):
    v = np.ascontiguousarray(vertices)
    t = np.ascontiguousarray(triangles, dtype=np.uint32).reshape(-1, 3)
    m = np.ascontiguousarray(matrices, dtype=np.float32).reshape(-1, 16)
    c = None if colors is None else np.ascontiguousarray(colors, dtype=np.float32).reshape(-1, 4)
    ov = np.zeros(len(v) * len(m), dtype=S.VERTEX_DTYPE)
    ot = np.zeros((len(t) * len(m), 3), dtype=np.uint32)
    lib().orc_bake_instanced(_p(v), C.c_uint64(len(v)), C.c_uint64(80), _p(t), C.c_uint64(len(t)), _p(m), _p(c),
                             C.c_uint64(len(m)), _p(ov), _p(ot))
    return ov, ot


def bake_skinned(skinned_vertices, bone_matrices):
    v = np.ascontiguousarray(skinned_vertices)
    assert v.dtype == S.SKINNED_VERTEX_DTYPE
    b = np.ascontiguousarray(bone_matrices, dtype=np.float32).reshape(-1, 16)
    ov = np.zeros(len(v), dtype=S.VERTEX_DTYPE)
    rc = lib().orc_bake_skinned(_p(v), C.c_uint64(len(v)), C.c_uint64(144), _p(b), C.c_uint64(len(b)), _p(ov))
    if rc:
        raise ValueError("bone index out of range")
    return ov


def flux_probes(probes):
    out = np.zeros(8, dtype=np.float32)
    lib().orc_flux_probes(_p(probes), C.c_uint64(len(probes)), _p(out))
    return out


def ref_btf_lib():
    """oracle/_ref/libbtf_ref.so = the reference's own MLVQ headers compiled on the host (make -C oracle ref; exists only
    where /root/reference does).  None when absent."""
    so = os.path.join(_HERE, "_ref", "libbtf_ref.so")
    return C.CDLL(so) if os.path.exists(so) else None


def ref_btf_value_deg(tables, beta_angles, q6):
    """The REFERENCE's BTFBase::GetValueDeg on a structs.BTFTables set."""
    L = ref_btf_lib()
    if L is None:
        raise FileNotFoundError("oracle/_ref/libbtf_ref.so (make -C oracle ref)")
    q = np.ascontiguousarray(q6, dtype=np.float32).reshape(-1, 6)
    ba = np.ascontiguousarray(beta_angles, dtype=np.float32)
    d = tables.desc()
    out = np.empty((len(q), 3), dtype=np.float32)
    rc = L.ref_btf_value_deg(C.byref(d), _p(ba), C.c_uint64(len(q)), _p(q), _p(out))
    assert rc == 0
    return out


def ref_ie_lib():
    """oracle/_ref/libie_ref.so = the reference's own illumination-estimation device programs compiled on the host over
    an emulation of the OptiX intrinsics (make -C oracle ref; exists only where /root/reference does).  None when absent."""
    so = os.path.join(_HERE, "_ref", "libie_ref.so")
    return C.CDLL(so) if os.path.exists(so) else None


class libm_trig:
    """Context manager: the oracle's asin / acos / atan2 / sin / cos call libm, as the reference's code does, instead of
    the portable polynomials (test switch g_libm_trig, oracle_math.h)."""

    def __enter__(self):
        self.was = lib().orc_set_libm_trig(1)

    def __exit__(self, *a):
        lib().orc_set_libm_trig(self.was)


def ref_env_lib():
    """oracle/_ref/libenv_ref.so = the reference's own RayTracerUtilities.cuh / Environment.cuh / BSDF.cuh compiled on the
    host (make -C oracle ref; exists only where /root/reference does).  None when absent."""
    so = os.path.join(_HERE, "_ref", "libenv_ref.so")
    if not os.path.exists(so):
        return None
    L = C.CDLL(so)
    L.ref_cie_sky.restype = C.c_float
    return L


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


class RefEnv:
    """The REFERENCE's helpers (ref_env_driver.cpp).  Samplers return (uniforms consumed, result)."""

    def __init__(self):
        self.L = ref_env_lib()
        if self.L is None:
            raise FileNotFoundError("oracle/_ref/libenv_ref.so (make -C oracle ref)")

    def reflect(self, i, n):
        out = (C.c_float * 3)()
        self.L.ref_reflect(_f3(i), _f3(n), out)
        return np.array(out[:], dtype=np.float32)

    def tangent_space(self, n):
        out = (C.c_float * 9)()
        self.L.ref_tangent_space(_f3(n), out)
        return np.array(out[:], dtype=np.float32).reshape(3, 3)

    def sample_hemisphere(self, v0, v1, n):
        r, out = (C.c_float * 2)(), (C.c_float * 3)()
        self.L.ref_sample_hemisphere(C.c_uint(v0), C.c_uint(v1), _f3(n), r, out)
        return np.array(r[:], dtype=np.float32), np.array(out[:], dtype=np.float32)

    def sample_lobe(self, v0, v1, n, alpha):
        r, out = (C.c_float * 2)(), (C.c_float * 3)()
        self.L.ref_sample_lobe(C.c_uint(v0), C.c_uint(v1), _f3(n), C.c_float(alpha), r, out)
        return np.array(r[:], dtype=np.float32), np.array(out[:], dtype=np.float32)

    def brdf(self, metallic, v0, v1, in_dir, normal):
        r, out = (C.c_float * 2)(), (C.c_float * 3)()
        self.L.ref_brdf(C.c_float(metallic), C.c_uint(v0), C.c_uint(v1), _f3(in_dir), _f3(normal), r, out)
        return np.array(r[:], dtype=np.float32), np.array(out[:], dtype=np.float32)

    def cube_map_lookup(self, d):
        face, uv = C.c_int(), (C.c_float * 2)()
        self.L.ref_cube_map_lookup(_f3(d), C.byref(face), uv)
        return face.value, np.array(uv[:], dtype=np.float32)

    def ray_sphere(self, o, d, radius):
        t = (C.c_float * 2)()
        ok = self.L.ref_ray_sphere(_f3(o), _f3(d), C.c_float(radius), t)
        return bool(ok), np.array(t[:], dtype=np.float32)

    def nishita_sky(self, env, pos, d):
        out = (C.c_float * 3)()
        self.L.ref_nishita_sky(C.byref(env), _f3(pos), _f3(d), out)
        return np.array(out[:], dtype=np.float32)

    def env_light(self, env, pos, d):
        out = (C.c_float * 3)()
        self.L.ref_env_light(C.byref(env), _f3(pos), _f3(d), out)
        return np.array(out[:], dtype=np.float32)


def reflect(i, n):
    out = (C.c_float * 3)()
    lib().orc_reflect(_f3(i), _f3(n), out)
    return np.array(out[:], dtype=np.float32)


def tangent_space(n):
    out = (C.c_float * 9)()
    lib().orc_tangent_space(_f3(n), out)
    return np.array(out[:], dtype=np.float32).reshape(3, 3)


def cube_map_lookup(d):
    uv = (C.c_float * 2)()
    face = lib().orc_cube_map_lookup(_f3(d), uv)
    return int(face), np.array(uv[:], dtype=np.float32)


def ray_sphere(o, d, radius):
    t = (C.c_float * 2)()
    ok = lib().orc_ray_sphere(_f3(o), _f3(d), C.c_float(radius), t)
    return bool(ok), np.array(t[:], dtype=np.float32)


def rotate_about(v, u_turns, axis):
    out = (C.c_float * 3)()
    lib().orc_rotate_about((C.c_float * 3)(*v), C.c_float(u_turns), (C.c_float * 3)(*axis), out)
    return np.array(list(out), dtype=np.float32)


def scan_grid_count(grid):
    return int(lib().orc_scan_grid_count(C.byref(grid)))


def scan_grid_rays(grid, ray_begin, ray_end):
    rays = np.empty((ray_end - ray_begin, 8), dtype=np.float32)
    lib().orc_scan_grid_rays(C.byref(grid), C.c_uint64(ray_begin), C.c_uint64(ray_end), _p(rays))
    return rays
