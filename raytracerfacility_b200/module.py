"""Host-side mirror of the reference's module façade over the C ABI of librtf_b200.so (include/rtf/rtf.h).

Reference interface mirrored (paths relative to /root/reference):
  CudaModule::{Init, Terminate, GetRayTracer, EstimateIlluminationRayTracing, SamplePointCloud}
      include/RayTracerFacility/RayTracer/CUDAModule.hpp:9-46, src/RayTracerFacility/CUDAModule.cpp:23-64
  RayTracer::{RenderToCamera, BuildIAS, m_materials, m_geometries, m_instances}
      include/RayTracerFacility/RayTracer/RayTracer.hpp:358-394
  CameraProperties::{Resize, Set}    src/RayTracerFacility/RayTracer.cu:28-92

There is no CPU path: if the CUDA library is missing or no GPU is present every call raises.
Nothing here imports oracle/.
"""
import ctypes as C
import os

import numpy as np

from . import structs as S

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTF_B200_LIB", os.path.join(_HERE, "librtf_b200.so"))  # env override: developer A/B builds

OK, ERR_INVALID, ERR_CUDA, ERR_NOT_READY, ERR_NOT_FOUND, ERR_LIMIT = 0, -1, -2, -3, -4, -5
MISS = 0xFFFFFFFF

# every symbol include/rtf/rtf.h declares (tests check that the library exports all of them)
EXPORTS = [
    "rtf_create", "rtf_destroy", "rtf_last_error", "rtf_version", "rtf_set_stream", "rtf_launch_count",
    "rtf_set_geometry", "rtf_set_material", "rtf_set_instance", "rtf_remove_geometry", "rtf_remove_material",
    "rtf_remove_instance", "rtf_commit", "rtf_triangle_count", "rtf_estimate_illumination",
    "rtf_estimate_illumination_mesh", "rtf_estimate_illumination_mesh_device", "rtf_illumination_finalize_device",
    "rtf_last_job_stats", "rtf_scan_point_cloud", "rtf_scan_grid_ray_count", "rtf_scan_grid_dims", "rtf_scan_grid_trace",
    "rtf_scan_grid_trace_device", "rtf_trace_closest", "rtf_trace_closest_device", "rtf_trace_closest_counts",
    "rtf_camera_create", "rtf_camera_destroy", "rtf_camera_get_props", "rtf_camera_resize", "rtf_camera_set",
    "rtf_render_camera", "rtf_render_camera_rows", "rtf_camera_download", "rtf_camera_upload", "rtf_camera_device_buffer", "rtf_debug_size", "rtf_debug_read",
    "rtf_debug_upload_bvh", "rtf_microbench_fp32", "rtf_microbench_l2", "rtf_build_timings", "rtf_set_option",
    "rtf_set_texture", "rtf_remove_texture", "rtf_set_material_textures", "rtf_set_environment_map", "rtf_sample_texture",
    "rtf_sample_environment", "rtf_set_btf", "rtf_btf_import_folder", "rtf_remove_btf", "rtf_set_material_btf",
    "rtf_btf_eval", "rtf_set_geometry_instanced", "rtf_set_geometry_skinned", "rtf_geometry_read_baked",
    "rtf_prepare_light_probes", "rtf_illumination_flux", "rtf_illumination_flux_mesh_device", "rtf_sss_probe",
    "rtf_commit_stats", "rtf_illumination_flux_mesh_device_static",
    # multi-GPU (csrc/multi.cu)
    "rtf_comm_unique_id", "rtf_comm_init_rank", "rtf_comm_destroy", "rtf_comm_rank", "rtf_comm_barrier",
    "rtf_comm_reduce_f64", "rtf_shard_range", "rtf_estimate_illumination_mesh_sharded_device",
    "rtf_estimate_illumination_mesh_sharded", "rtf_scan_grid_trace_sharded_device", "rtf_scan_grid_shard_layout",
    "rtf_render_camera_sharded",
    "rtf_phase_timings", "rtf_profile_read", "rtf_create_multi", "rtf_multi_destroy", "rtf_multi_size", "rtf_multi_ctx",
    "rtf_multi_set_geometry", "rtf_multi_set_material", "rtf_multi_set_instance", "rtf_multi_set_option",
    "rtf_multi_commit", "rtf_multi_estimate_illumination_mesh", "rtf_multi_scan_grid_trace", "rtf_multi_trace_closest",
    "rtf_multi_camera_create", "rtf_multi_camera_destroy", "rtf_multi_camera_resize", "rtf_multi_camera_set",
    "rtf_multi_camera_get_props", "rtf_multi_render_camera", "rtf_multi_camera_download",
]


class RtfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"librtf_b200 error {code}: {msg}")
        self.code = code


_lib = None


def scan_grid_shard_layout(grid, nranks, rank):
    """(block_rays, n_blocks, n_local) of the interleaved LiDAR shards (host arithmetic of the library, no device)."""
    blk, nb, nl = C.c_uint64(), C.c_uint64(), C.c_uint64()
    _check(lib().rtf_scan_grid_shard_layout(C.byref(grid), int(nranks), int(rank), C.byref(blk), C.byref(nb), C.byref(nl)))
    return blk.value, nb.value, nl.value


def lib():
    """Load librtf_b200.so (built in-tree by `make -C raytracerfacility_b200/csrc` / __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with __graft_entry__.build(); there is no CPU fallback")
        try:  # the library links libnccl.so.2: let torch's bundled NCCL be the one the process loads (same soname;
            import torch  # noqa: F401  loading the system one first would leave torch.distributed with a mismatched NCCL)
        except ImportError:
            pass
        L = C.CDLL(LIB_PATH)
        L.rtf_last_error.restype = C.c_char_p
        L.rtf_version.restype = C.c_char_p
        L.rtf_launch_count.restype = C.c_uint64
        L.rtf_launch_count.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def _check(rc):
    if rc < 0:
        raise RtfError(rc, lib().rtf_last_error().decode())
    return rc


def _p(a):
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    return a.ctypes.data_as(C.c_void_p)


class Camera:
    """CameraProperties (RayTracer.hpp:30-96) with headless device float4 frame buffers."""

    def __init__(self, rt):
        self.rt = rt
        self.h = C.c_void_p()
        _check(lib().rtf_camera_create(rt.h, C.byref(self.h)))
        pp = C.POINTER(S.CameraProps)()
        _check(lib().rtf_camera_get_props(self.h, C.byref(pp)))
        self.props = pp.contents  # borrowed, mutable: same in-place semantics as the reference struct

    def resize(self, width, height):
        _check(lib().rtf_camera_resize(self.h, int(width), int(height)))

    def set(self, position, rotation_wxyz=(1.0, 0.0, 0.0, 0.0)):
        _check(lib().rtf_camera_set(self.h, (C.c_float * 3)(*position), (C.c_float * 4)(*rotation_wxyz)))

    def download(self, which=S.OUT_COLOR):
        out = np.empty((self.props.height, self.props.width, 4), dtype=np.float32)
        _check(lib().rtf_camera_download(self.h, int(which), _p(out)))
        return out

    def upload(self, data, which=S.OUT_COLOR):
        a = np.ascontiguousarray(data, dtype=np.float32)
        assert a.size == self.props.height * self.props.width * 4
        _check(lib().rtf_camera_upload(self.h, int(which), _p(a)))

    def device_buffer(self, which=S.OUT_COLOR):
        d = C.c_void_p()
        _check(lib().rtf_camera_device_buffer(self.h, int(which), C.byref(d)))
        return d.value

    def destroy(self):
        if self.h:
            lib().rtf_camera_destroy(self.h)
            self.h = C.c_void_p()


class RayTracer:
    """One context on one GPU = the reference's `RayTracer` object owned by `CudaModule`."""

    def __init__(self, device=0):
        self.h = C.c_void_p()
        _check(lib().rtf_create(int(device), C.byref(self.h)))
        self.device = device

    # ---- lifecycle ----
    def destroy(self):
        if self.h:
            lib().rtf_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    def set_stream(self, cuda_stream):
        _check(lib().rtf_set_stream(self.h, C.c_void_p(cuda_stream)))

    def set_option(self, name, value):
        _check(lib().rtf_set_option(self.h, name.encode(), C.c_int64(int(value))))

    @property
    def launch_count(self):
        return int(lib().rtf_launch_count(self.h))

    # ---- scene sync (m_geometries / m_materials / m_instances + BuildIAS) ----
    def set_geometry(self, handle, vertices, triangles, version=0):
        v = np.ascontiguousarray(vertices)
        assert v.dtype == S.VERTEX_DTYPE
        t = np.ascontiguousarray(triangles, dtype=np.uint32).reshape(-1, 3)
        _check(lib().rtf_set_geometry(self.h, C.c_uint64(handle), _p(v), C.c_size_t(len(v)), C.c_size_t(80), _p(t),
                                      C.c_size_t(len(t)), C.c_uint64(version)))

    def set_geometry_instanced(self, handle, vertices, triangles, matrices, colors=None, version=0):
        """RendererType::Instanced geometry: device bake of CopyVerticesInstancedKernel (RayTracer.cu:1148-1175)."""
        v = np.ascontiguousarray(vertices)
        assert v.dtype == S.VERTEX_DTYPE
        t = np.ascontiguousarray(triangles, dtype=np.uint32).reshape(-1, 3)
        m = np.ascontiguousarray(matrices, dtype=np.float32).reshape(-1, 16)
        c = None if colors is None else np.ascontiguousarray(colors, dtype=np.float32).reshape(-1, 4)
        assert c is None or len(c) == len(m)
        _check(lib().rtf_set_geometry_instanced(self.h, C.c_uint64(handle), _p(v), C.c_size_t(len(v)), C.c_size_t(80),
                                                _p(t), C.c_size_t(len(t)), _p(m), _p(c), C.c_size_t(len(m)),
                                                C.c_uint64(version)))

    def set_geometry_skinned(self, handle, skinned_vertices, triangles, bone_matrices, version=0):
        """RendererType::Skinned geometry: device bake of CopySkinnedVerticesKernel (RayTracer.cu:1194-1247)."""
        v = np.ascontiguousarray(skinned_vertices)
        assert v.dtype == S.SKINNED_VERTEX_DTYPE
        t = np.ascontiguousarray(triangles, dtype=np.uint32).reshape(-1, 3)
        b = np.ascontiguousarray(bone_matrices, dtype=np.float32).reshape(-1, 16)
        _check(lib().rtf_set_geometry_skinned(self.h, C.c_uint64(handle), _p(v), C.c_size_t(len(v)), C.c_size_t(144),
                                              _p(t), C.c_size_t(len(t)), _p(b), C.c_size_t(len(b)),
                                              C.c_uint64(version)))

    def read_baked_geometry(self, handle, n_vertices, n_triangles):
        v = np.zeros(n_vertices, dtype=S.VERTEX_DTYPE)
        t = np.zeros((n_triangles, 3), dtype=np.uint32)
        _check(lib().rtf_geometry_read_baked(self.h, C.c_uint64(handle), _p(v), C.c_size_t(n_vertices), _p(t),
                                             C.c_size_t(n_triangles)))
        return v, t

    def set_material(self, handle, material):
        _check(lib().rtf_set_material(self.h, C.c_uint64(handle), C.byref(material)))

    # ---- images in linear memory + compressed BTF (the GL texture ids / CompressedBTF assets of the reference) ----
    def set_texture(self, handle, rgba):
        """rgba: (height, width, 4) float32."""
        t = np.ascontiguousarray(rgba, dtype=np.float32)
        assert t.ndim == 3 and t.shape[2] == 4
        _check(lib().rtf_set_texture(self.h, C.c_uint64(handle), int(t.shape[1]), int(t.shape[0]), _p(t)))

    def remove_texture(self, handle):
        _check(lib().rtf_remove_texture(self.h, C.c_uint64(handle)))

    def set_material_textures(self, material, albedo=0, normal=0, metallic=0, roughness=0):
        _check(lib().rtf_set_material_textures(self.h, C.c_uint64(material), C.c_uint64(albedo), C.c_uint64(normal),
                                               C.c_uint64(metallic), C.c_uint64(roughness)))

    def set_environment_map(self, faces):
        """faces: six texture handles (+X -X +Y -Y +Z -Z) or None."""
        arr = (C.c_uint64 * 6)(*faces) if faces is not None else None
        _check(lib().rtf_set_environment_map(self.h, arr))

    def sample_texture(self, handle, uv):
        uv = np.ascontiguousarray(uv, dtype=np.float32).reshape(-1, 2)
        out = np.empty((len(uv), 4), dtype=np.float32)
        _check(lib().rtf_sample_texture(self.h, C.c_uint64(handle), C.c_size_t(len(uv)), _p(uv), _p(out)))
        return out

    def sample_environment(self, env, pos_dir):
        pd = np.ascontiguousarray(pos_dir, dtype=np.float32).reshape(-1, 6)
        out = np.empty((len(pd), 3), dtype=np.float32)
        _check(lib().rtf_sample_environment(self.h, C.byref(env), C.c_size_t(len(pd)), _p(pd), _p(out)))
        return out

    def set_btf(self, handle, tables):
        """tables: structs.BTFTables."""
        d = tables.desc()
        _check(lib().rtf_set_btf(self.h, C.c_uint64(handle), C.byref(d)))

    def btf_import_folder(self, handle, folder):
        """CompressedBTF::ImportFromFolder (CompressedBTF.cpp:74-390)."""
        _check(lib().rtf_btf_import_folder(self.h, C.c_uint64(handle), str(folder).encode()))

    def remove_btf(self, handle):
        _check(lib().rtf_remove_btf(self.h, C.c_uint64(handle)))

    def set_material_btf(self, material, btf):
        _check(lib().rtf_set_material_btf(self.h, C.c_uint64(material), C.c_uint64(btf)))

    def btf_eval(self, btf, tex_coord, view_dir, illum_dir, normal, tangent):
        """SurfaceCompressedBTF::GetValue (RayDataDefinations.hpp:329-354) for n queries -> (n, 3) rgb."""
        a = [np.ascontiguousarray(x, dtype=np.float32) for x in (tex_coord, view_dir, illum_dir, normal, tangent)]
        n = len(a[0])
        out = np.empty((n, 3), dtype=np.float32)
        _check(lib().rtf_btf_eval(self.h, C.c_uint64(btf), C.c_size_t(n), _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(a[4]),
                                  _p(out)))
        return out

    def set_instance(self, handle, geometry, material, transform, component_handle=0):
        x = np.ascontiguousarray(transform, dtype=np.float32).reshape(16)
        _check(lib().rtf_set_instance(self.h, C.c_uint64(handle), C.c_uint64(geometry), C.c_uint64(material), _p(x),
                                      C.c_uint64(component_handle)))

    def remove_geometry(self, handle):
        _check(lib().rtf_remove_geometry(self.h, C.c_uint64(handle)))

    def remove_material(self, handle):
        _check(lib().rtf_remove_material(self.h, C.c_uint64(handle)))

    def remove_instance(self, handle):
        _check(lib().rtf_remove_instance(self.h, C.c_uint64(handle)))

    def build_ias(self, morton_bits=63):
        """RayTracer::BuildIAS: flatten + LBVH build on the device."""
        _check(lib().rtf_commit(self.h, int(morton_bits)))

    commit = build_ias

    def load_scene(self, desc, morton_bits=63, presplit=None):
        for gh, (v, t) in desc.geometries.items():
            self.set_geometry(gh, v, t)
        for mh, m in desc.materials.items():
            self.set_material(mh, m)
        for ih, (g, m, x, comp) in desc.instances.items():
            self.set_instance(ih, g, m, x, comp)
        for th, img in getattr(desc, "textures", {}).items():
            self.set_texture(th, img)
        for mh, tex in getattr(desc, "material_textures", {}).items():
            self.set_material_textures(mh, *tex)
        for bh, tables in getattr(desc, "btfs", {}).items():
            self.set_btf(bh, tables)
        for mh, bh in getattr(desc, "material_btf", {}).items():
            self.set_material_btf(mh, bh)
        self.set_environment_map(getattr(desc, "env_map", None))
        if presplit is not None:
            self.set_option("presplit", presplit)
        self.build_ias(morton_bits)

    def commit_stats(self):
        """{full builds, material-only commits, skipped commits, skipped geometry uploads} (scene-sync diffing)."""
        out = (C.c_uint64 * 5)()
        _check(lib().rtf_commit_stats(self.h, out))
        return dict(zip(("full", "materials", "skipped", "uploads_skipped", "refits"), [int(x) for x in out]))

    @property
    def triangle_count(self):
        n = C.c_uint64()
        _check(lib().rtf_triangle_count(self.h, C.byref(n)))
        return n.value

    def build_timings(self):
        """ms per build stage of the last commit: flatten, bounds+morton, sort, karras+refit, collapse."""
        out = (C.c_float * 8)()
        _check(lib().rtf_build_timings(self.h, out, 8))
        names = ["flatten", "bounds_morton", "sort", "karras_refit", "collapse", "total"]
        return {k: float(out[i]) for i, k in enumerate(names)}

    # ---- illumination (CudaModule::EstimateIlluminationRayTracing) ----
    def estimate_illumination(self, env, ray_props, probes, seed, push_normal_distance=0.001):
        assert probes.dtype == S.ILLUM_SAMPLER_DTYPE and probes.flags.c_contiguous
        _check(lib().rtf_estimate_illumination(self.h, C.byref(env), C.byref(ray_props), _p(probes),
                                               C.c_size_t(len(probes)), C.c_uint32(seed),
                                               C.c_float(push_normal_distance)))

    def estimate_illumination_mesh(self, env, ray_props, seed, push=0.001, front=True, back=True, tri_begin=0,
                                   tri_end=None, s0=0, s1=None, normalise=True, out=None):
        tri_end = self.triangle_count if tri_end is None else tri_end
        s1 = ray_props.samples if s1 is None else s1
        if out is None:
            out = np.zeros((max(0, tri_end - tri_begin), 6), dtype=np.float32)
        _check(lib().rtf_estimate_illumination_mesh(self.h, C.byref(env), C.byref(ray_props), C.c_uint32(seed),
                                                    C.c_float(push), int(front), int(back), C.c_uint64(tri_begin),
                                                    C.c_uint64(tri_end), int(s0), int(s1), int(normalise), _p(out)))
        return out

    def estimate_illumination_mesh_device(self, env, ray_props, seed, push, front, back, tri_begin, tri_end, s0, s1,
                                          dev_accum6_ptr):
        _check(lib().rtf_estimate_illumination_mesh_device(self.h, C.byref(env), C.byref(ray_props), C.c_uint32(seed),
                                                           C.c_float(push), int(front), int(back),
                                                           C.c_uint64(tri_begin), C.c_uint64(tri_end), int(s0), int(s1),
                                                           C.c_void_p(dev_accum6_ptr)))

    def illumination_finalize_device(self, dev_accum6_ptr, n_triangles, faces, samples):
        _check(lib().rtf_illumination_finalize_device(self.h, C.c_void_p(dev_accum6_ptr), C.c_uint64(n_triangles),
                                                      int(faces), int(samples)))

    def prepare_light_probes(self, tri_begin=0, tri_end=None, front=True, back=True):
        """TriangleIlluminationEstimator::PrepareLightProbeGroup on the device."""
        tri_end = self.triangle_count if tri_end is None else tri_end
        out = np.zeros(max(0, tri_end - tri_begin), dtype=S.ILLUM_SAMPLER_DTYPE)
        _check(lib().rtf_prepare_light_probes(self.h, C.c_uint64(tri_begin), C.c_uint64(tri_end), int(front), int(back),
                                              _p(out)))
        return out

    def illumination_flux(self, probes):
        """SampleLightProbeGroup's flux integration: {totalFlux[3], totalArea, averageFlux[3], 0}."""
        assert probes.dtype == S.ILLUM_SAMPLER_DTYPE and probes.flags.c_contiguous
        out = np.zeros(8, dtype=np.float32)
        _check(lib().rtf_illumination_flux(self.h, _p(probes), C.c_size_t(len(probes)), _p(out)))
        return out

    def illumination_flux_mesh_device(self, dev_energy6_ptr, tri_begin, tri_end, front=True, back=True):
        out = np.zeros(8, dtype=np.float32)
        _check(lib().rtf_illumination_flux_mesh_device(self.h, C.c_void_p(dev_energy6_ptr), C.c_uint64(tri_begin),
                                                       C.c_uint64(tri_end), int(front), int(back), _p(out)))
        return out

    def illumination_flux_mesh_device_static(self, dev_energy6_ptr, tri_begin, tri_end, dev_out8_ptr, front=True, back=True):
        """The flux integration with its scratch in the library's static device array: no allocation on any path, at most
        2^24 triangles per call, result left in the caller's 8 device floats (enqueued on the context's stream)."""
        _check(lib().rtf_illumination_flux_mesh_device_static(self.h, C.c_void_p(dev_energy6_ptr), C.c_uint64(tri_begin),
                                                              C.c_uint64(tri_end), int(front), int(back),
                                                              C.c_void_p(dev_out8_ptr)))

    def sss_probe(self, rays, handles):
        """Test hook: BSSRDF spatial probe (BSSDF.cuh:18-89); per ray {n, prim[4], t[4], u[4], v[4]} (uint32 view)."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        handles = np.ascontiguousarray(handles, dtype=np.uint64)
        out = np.zeros((len(rays), 17), dtype=np.uint32)
        _check(lib().rtf_sss_probe(self.h, _p(rays), _p(handles), C.c_size_t(len(rays)), _p(out)))
        return out

    def last_job_stats(self):
        r, p = C.c_uint64(), C.c_uint64()
        _check(lib().rtf_last_job_stats(self.h, C.byref(r), C.byref(p)))
        return r.value, p.value

    # ---- LiDAR (CudaModule::SamplePointCloud) ----
    def sample_point_cloud(self, env, samples, want_prim_ids=False):
        assert samples.dtype == S.PC_SAMPLE_DTYPE and samples.flags.c_contiguous
        prim = np.empty(len(samples), dtype=np.uint32) if want_prim_ids else None
        _check(lib().rtf_scan_point_cloud(self.h, C.byref(env), _p(samples), C.c_size_t(len(samples)), _p(prim)))
        return prim

    def scan_grid_ray_count(self, grid):
        n = C.c_uint64()
        _check(lib().rtf_scan_grid_ray_count(C.byref(grid), C.byref(n)))
        return n.value

    def scan_grid(self, grid, ray_begin=0, ray_end=None):
        ray_end = self.scan_grid_ray_count(grid) if ray_end is None else ray_end
        n = max(0, ray_end - ray_begin)
        t = np.empty(n, dtype=np.float32)
        prim = np.empty(n, dtype=np.uint32)
        inst = np.empty(n, dtype=np.uint32)
        _check(lib().rtf_scan_grid_trace(self.h, C.byref(grid), C.c_uint64(ray_begin), C.c_uint64(ray_end), _p(t),
                                         _p(prim), _p(inst)))
        return t, prim, inst

    def scan_grid_device(self, grid, ray_begin, ray_end, dev_t_ptr, dev_prim_ptr):
        _check(lib().rtf_scan_grid_trace_device(self.h, C.byref(grid), C.c_uint64(ray_begin), C.c_uint64(ray_end),
                                                C.c_void_p(dev_t_ptr), C.c_void_p(dev_prim_ptr)))

    # ---- raw closest hit ----
    def trace_closest(self, rays, counts=False, out=None):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        n = len(rays)
        hits = out if out is not None else np.empty(n, dtype=S.HIT_DTYPE)
        if counts:
            nv, tv = np.empty(n, dtype=np.uint32), np.empty(n, dtype=np.uint32)
            _check(lib().rtf_trace_closest_counts(self.h, _p(rays), C.c_size_t(n), _p(hits), _p(nv), _p(tv)))
            return hits, nv, tv
        _check(lib().rtf_trace_closest(self.h, _p(rays), C.c_size_t(n), _p(hits)))
        return hits

    def trace_closest_device(self, dev_rays_ptr, n, dev_hits_ptr):
        _check(lib().rtf_trace_closest_device(self.h, C.c_void_p(dev_rays_ptr), C.c_size_t(n),
                                              C.c_void_p(dev_hits_ptr)))

    # ---- camera (RayTracer::RenderToCamera) ----
    def create_camera(self):
        return Camera(self)

    def render_to_camera(self, env, camera, ray_props):
        """Returns True when a frame was rendered, False when there is no acceleration structure."""
        return _check(lib().rtf_render_camera(self.h, C.byref(env), camera.h, C.byref(ray_props))) == 1

    def render_to_camera_rows(self, env, camera, ray_props, row_begin, row_end):
        """RenderToCamera for the image rows [row_begin, row_end) only (multi-GPU bands); same pixel values as the
        full-frame call."""
        return _check(lib().rtf_render_camera_rows(self.h, C.byref(env), camera.h, C.byref(ray_props), int(row_begin),
                                                   int(row_end))) == 1


    # ---- multi-GPU, one process per GPU: the library's own NCCL communicator (csrc/multi.cu) ----
    @staticmethod
    def comm_unique_id():
        buf = (C.c_ubyte * 128)()
        _check(lib().rtf_comm_unique_id(buf, C.c_size_t(128)))
        return bytes(buf)

    def comm_init_rank(self, nranks, rank, unique_id):
        idb = (C.c_ubyte * 128).from_buffer_copy(unique_id) if unique_id is not None else None
        _check(lib().rtf_comm_init_rank(self.h, int(nranks), int(rank), idb))

    def comm_destroy(self):
        _check(lib().rtf_comm_destroy(self.h))

    def comm_rank(self):
        r, n = C.c_int(), C.c_int()
        _check(lib().rtf_comm_rank(self.h, C.byref(r), C.byref(n)))
        return r.value, n.value

    def comm_barrier(self):
        _check(lib().rtf_comm_barrier(self.h))

    def comm_reduce(self, values, op="max"):
        v = np.ascontiguousarray(values, dtype=np.float64).reshape(-1).copy()
        _check(lib().rtf_comm_reduce_f64(self.h, _p(v), int(len(v)), 1 if op == "max" else 0))
        return v

    def estimate_illumination_mesh_sharded_device(self, env, ray_props, seed, push, front, back, mode, dev_out6_ptr):
        _check(lib().rtf_estimate_illumination_mesh_sharded_device(self.h, C.byref(env), C.byref(ray_props), C.c_uint32(seed),
                                                                   C.c_float(push), int(front), int(back), int(mode),
                                                                   C.c_void_p(dev_out6_ptr)))

    def estimate_illumination_mesh_sharded(self, env, ray_props, seed, push=0.001, front=True, back=True, mode=0):
        out = np.zeros((self.triangle_count, 6), dtype=np.float32)
        _check(lib().rtf_estimate_illumination_mesh_sharded(self.h, C.byref(env), C.byref(ray_props), C.c_uint32(seed),
                                                            C.c_float(push), int(front), int(back), int(mode), _p(out)))
        return out

    def scan_grid_sharded_device(self, grid, gather, dev_t_ptr, dev_prim_ptr, dev_inst_ptr=None):
        """Traces this rank's interleaved blocks into full-size device arrays; returns (block_rays, n_blocks): block k of
        the grid (rays [k * block_rays, (k + 1) * block_rays)) belongs to rank k mod nranks."""
        blk, nb = C.c_uint64(), C.c_uint64()
        _check(lib().rtf_scan_grid_trace_sharded_device(self.h, C.byref(grid), int(gather), C.c_void_p(dev_t_ptr),
                                                        C.c_void_p(dev_prim_ptr), C.c_void_p(dev_inst_ptr), C.byref(blk),
                                                        C.byref(nb)))
        return blk.value, nb.value

    def render_to_camera_sharded(self, env, camera, ray_props, outputs_mask=1):
        return _check(lib().rtf_render_camera_sharded(self.h, C.byref(env), camera.h, C.byref(ray_props),
                                                      int(outputs_mask))) == 1

    def profile_read(self):
        """Device ms per kernel category since the last read (set_option("profile", 1) first)."""
        out = (C.c_float * 6)()
        _check(lib().rtf_profile_read(self.h, out))
        return dict(zip(("gen_ms", "trace_ms", "shade_ms", "reduce_ms", "sort_ms", "other_ms"), [float(x) for x in out]))

    def phase_timings(self):
        """{compute, collective, finalize} device ms of the last sharded call."""
        out = (C.c_float * 3)()
        _check(lib().rtf_phase_timings(self.h, out))
        return {"compute_ms": float(out[0]), "collective_ms": float(out[1]), "finalize_ms": float(out[2])}

    # ---- parity read-back ----
    def debug(self, what):
        nbytes = C.c_uint64()
        _check(lib().rtf_debug_size(self.h, int(what), C.byref(nbytes)))
        dt = S.DBG_DTYPES[what]
        out = np.empty(nbytes.value // dt.itemsize, dtype=dt)
        _check(lib().rtf_debug_read(self.h, int(what), _p(out), nbytes))
        return out

    def debug_bytes(self, what):
        """Size in bytes of a build-stage array without reading it back."""
        nbytes = C.c_uint64()
        _check(lib().rtf_debug_size(self.h, int(what), C.byref(nbytes)))
        return int(nbytes.value)

    def debug_upload_bvh(self, nodes, tris, depth):
        nodes, tris = np.ascontiguousarray(nodes), np.ascontiguousarray(tris)
        _check(lib().rtf_debug_upload_bvh(self.h, _p(nodes), C.c_uint64(len(nodes)), _p(tris), C.c_uint64(len(tris)),
                                          int(depth)))

    def microbench_fp32(self):
        v = C.c_double()
        _check(lib().rtf_microbench_fp32(self.h, C.byref(v)))
        return v.value

    def microbench_l2(self, footprint_bytes):
        v = C.c_double()
        _check(lib().rtf_microbench_l2(self.h, C.c_double(footprint_bytes), C.byref(v)))
        return v.value


def shard_range(n, nranks, rank):
    b, e = C.c_uint64(), C.c_uint64()
    _check(lib().rtf_shard_range(C.c_uint64(int(n)), int(nranks), int(rank), C.byref(b), C.byref(e)))
    return b.value, e.value


class _Borrowed(RayTracer):
    """A context owned by a MultiRayTracer (never destroyed from here)."""

    def __init__(self, handle, device):
        self.h, self.device = handle, device

    def destroy(self):
        self.h = C.c_void_p()


class MultiCamera:
    def __init__(self, mrt):
        self.h = C.c_void_p()
        _check(lib().rtf_multi_camera_create(mrt.h, C.byref(self.h)))
        pp = C.POINTER(S.CameraProps)()
        _check(lib().rtf_multi_camera_get_props(self.h, C.byref(pp)))
        self.props = pp.contents

    def resize(self, width, height):
        _check(lib().rtf_multi_camera_resize(self.h, int(width), int(height)))

    def set(self, position, rotation_wxyz=(1.0, 0.0, 0.0, 0.0)):
        _check(lib().rtf_multi_camera_set(self.h, (C.c_float * 3)(*position), (C.c_float * 4)(*rotation_wxyz)))

    def download(self, which=S.OUT_COLOR):
        out = np.empty((self.props.height, self.props.width, 4), dtype=np.float32)
        _check(lib().rtf_multi_camera_download(self.h, int(which), _p(out)))
        return out

    def destroy(self):
        if self.h:
            lib().rtf_multi_camera_destroy(self.h)
            self.h = C.c_void_p()


class MultiRayTracer:
    """One process driving several GPUs through rtf_create_multi (ncclCommInitAll inside the library)."""

    def __init__(self, devices):
        self.devices = [int(d) for d in devices]
        self.h = C.c_void_p()
        arr = (C.c_int * len(self.devices))(*self.devices)
        _check(lib().rtf_create_multi(arr, len(self.devices), C.byref(self.h)))

    def destroy(self):
        if self.h:
            lib().rtf_multi_destroy(self.h)
            self.h = C.c_void_p()

    def ctx(self, rank):
        c = C.c_void_p()
        _check(lib().rtf_multi_ctx(self.h, int(rank), C.byref(c)))
        return _Borrowed(c, self.devices[rank])

    def set_option(self, name, value):
        _check(lib().rtf_multi_set_option(self.h, name.encode(), C.c_int64(int(value))))

    def load_scene(self, desc, morton_bits=63):
        for gh, (v, t) in desc.geometries.items():
            v = np.ascontiguousarray(v)
            t = np.ascontiguousarray(t, dtype=np.uint32).reshape(-1, 3)
            _check(lib().rtf_multi_set_geometry(self.h, C.c_uint64(gh), _p(v), C.c_size_t(len(v)), C.c_size_t(80), _p(t),
                                                C.c_size_t(len(t)), C.c_uint64(0)))
        for mh, m in desc.materials.items():
            _check(lib().rtf_multi_set_material(self.h, C.c_uint64(mh), C.byref(m)))
        for ih, (g, m, x, comp) in desc.instances.items():
            x = np.ascontiguousarray(x, dtype=np.float32).reshape(16)
            _check(lib().rtf_multi_set_instance(self.h, C.c_uint64(ih), C.c_uint64(g), C.c_uint64(m), _p(x), C.c_uint64(comp)))
        for r in range(len(self.devices)):  # textures / BTFs / cube map: per-context calls, repeated for every device
            c = self.ctx(r)
            for th, img in getattr(desc, "textures", {}).items():
                c.set_texture(th, img)
            for mh, tex in getattr(desc, "material_textures", {}).items():
                c.set_material_textures(mh, *tex)
            for bh, tables in getattr(desc, "btfs", {}).items():
                c.set_btf(bh, tables)
            for mh, bh in getattr(desc, "material_btf", {}).items():
                c.set_material_btf(mh, bh)
            c.set_environment_map(getattr(desc, "env_map", None))
        _check(lib().rtf_multi_commit(self.h, int(morton_bits)))

    def estimate_illumination_mesh(self, env, ray_props, seed, push=0.001, front=True, back=True, mode=0):
        out = np.zeros((self.ctx(0).triangle_count, 6), dtype=np.float32)
        _check(lib().rtf_multi_estimate_illumination_mesh(self.h, C.byref(env), C.byref(ray_props), C.c_uint32(seed),
                                                          C.c_float(push), int(front), int(back), int(mode), _p(out)))
        return out

    def scan_grid(self, grid):
        n = self.ctx(0).scan_grid_ray_count(grid)
        t, prim, inst = np.empty(n, np.float32), np.empty(n, np.uint32), np.empty(n, np.uint32)
        _check(lib().rtf_multi_scan_grid_trace(self.h, C.byref(grid), _p(t), _p(prim), _p(inst)))
        return t, prim, inst

    def trace_closest(self, rays, out=None):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        hits = out if out is not None else np.empty(len(rays), dtype=S.HIT_DTYPE)
        _check(lib().rtf_multi_trace_closest(self.h, _p(rays), C.c_size_t(len(rays)), _p(hits)))
        return hits

    def create_camera(self):
        return MultiCamera(self)

    def render_to_camera(self, env, camera, ray_props, outputs_mask=1):
        return _check(lib().rtf_multi_render_camera(self.h, C.byref(env), camera.h, C.byref(ray_props), int(outputs_mask))) == 1


class CudaModule:
    """Static façade with the reference's names (CUDAModule.hpp:30-45)."""
    _rt = None

    @classmethod
    def Init(cls, device=0):
        if cls._rt is None:
            cls._rt = RayTracer(device)

    @classmethod
    def Terminate(cls):
        if cls._rt is not None:
            cls._rt.destroy()
            cls._rt = None

    @classmethod
    def GetRayTracer(cls):
        if cls._rt is None:
            raise RtfError(ERR_NOT_READY, "CudaModule::Init has not been called")
        return cls._rt

    @classmethod
    def EstimateIlluminationRayTracing(cls, environment_properties, ray_properties, light_probes, seed,
                                       push_normal_distance):
        cls.GetRayTracer().estimate_illumination(environment_properties, ray_properties, light_probes, seed,
                                                 push_normal_distance)

    @classmethod
    def SamplePointCloud(cls, environment_properties, samples):
        cls.GetRayTracer().sample_point_cloud(environment_properties, samples)


def probes_from_scene(desc, front=True, back=True):
    """TriangleIlluminationEstimator::PrepareLightProbeGroup (TriangleIlluminationEstimator.cpp:94-139):
    one IlluminationSampler per world-space triangle, position = M*(p,1), normal = M*(n,0) (not renormalised),
    instances in ascending handle order.  Float32 arithmetic in glm's order."""
    out = []
    for ih in sorted(desc.instances):
        g, _, x, _ = desc.instances[ih]
        v, t = desc.geometries[g]
        m = np.asarray(x, dtype=np.float32).reshape(4, 4)  # m[c] = column c
        pr = np.zeros(len(t), dtype=S.ILLUM_SAMPLER_DTYPE)
        for k, name in enumerate(("a", "b", "c")):
            vk = v[t[:, k]]
            pr[name] = vk
            p, n = vk["position"], vk["normal"]
            one = np.float32(1.0)
            pw = (m[0][None, :3] * p[:, 0:1] + m[1][None, :3] * p[:, 1:2]) + (m[2][None, :3] * p[:, 2:3] + m[3][None, :3] * one)
            nw = (m[0][None, :3] * n[:, 0:1] + m[1][None, :3] * n[:, 1:2]) + (m[2][None, :3] * n[:, 2:3] + m[3][None, :3] * np.float32(0.0))
            pr[name]["position"] = pw.astype(np.float32)
            pr[name]["normal"] = nw.astype(np.float32)
        pr["front_face"] = 1 if front else 0
        pr["back_face"] = 1 if back else 0
        out.append(pr)
    return np.ascontiguousarray(np.concatenate(out)) if out else np.zeros(0, dtype=S.ILLUM_SAMPLER_DTYPE)


def probe_area(probes):
    """IlluminationSampler::GetArea (RayTracer.hpp:204-210), Heron."""
    a = np.linalg.norm(probes["a"]["position"] - probes["b"]["position"], axis=1)
    b = np.linalg.norm(probes["b"]["position"] - probes["c"]["position"], axis=1)
    c = np.linalg.norm(probes["c"]["position"] - probes["a"]["position"], axis=1)
    p = (a + b + c) * np.float32(0.5)
    return np.sqrt(np.maximum(p * (p - a) * (p - b) * (p - c), 0)).astype(np.float32)
