"""ctypes mirrors of the reference-layout POD records declared in include/rtf/rtf.h.

Reference layouts (relative to /root/reference):
  IlluminationSampler / PointCloudSample / EnvironmentProperties / RayProperties / CameraProperties:
      include/RayTracerFacility/RayTracer/RayTracer.hpp:30-232
  HitInfo: include/RayTracerFacility/RayTracer/HitInfo.hpp:4-11
  UniEngine::Vertex (80 B): inferred from src/RayTracerFacility/RayTracer.cu:1163-1173
"""
import ctypes as C

import numpy as np

f32, u32, i32, u64, u8 = C.c_float, C.c_uint32, C.c_int32, C.c_uint64, C.c_uint8


class Vertex(C.Structure):
    _fields_ = [("position", f32 * 3), ("position_padding", f32), ("normal", f32 * 3), ("normal_padding", f32),
                ("tangent", f32 * 3), ("tangent_padding", f32), ("color", f32 * 4), ("tex_coord", f32 * 2),
                ("tex_coord_padding", f32 * 2)]


class IlluminationSampler(C.Structure):
    _fields_ = [("a", Vertex), ("b", Vertex), ("c", Vertex), ("direction", f32 * 3), ("energy", f32 * 3),
                ("front_face", u8), ("back_face", u8), ("pad_", u8 * 2)]


class HitInfo(C.Structure):
    _fields_ = [("position", f32 * 3), ("normal", f32 * 3), ("tangent", f32 * 3), ("color", f32 * 4),
                ("tex_coord", f32 * 2), ("data", f32 * 4)]


class PointCloudSample(C.Structure):
    _fields_ = [("direction", f32 * 3), ("start", f32 * 3), ("handle", u64), ("hit", u8), ("pad_", u8 * 3),
                ("hit_info", HitInfo)]


class Atmosphere(C.Structure):
    _fields_ = [("earth_radius", f32), ("atmosphere_radius", f32), ("hr", f32), ("hm", f32), ("g", f32),
                ("num_samples", i32), ("num_samples_light", i32)]


ENV_SCENE, ENV_SKYDOME, ENV_SINGLE_LIGHT_SOURCE = 0, 1, 2
MAT_DEFAULT, MAT_VERTEX_COLOR, MAT_COMPRESSED_BTF = 0, 1, 2
OUT_COLOR, OUT_NORMAL, OUT_ALBEDO, OUT_DEPTH = 0, 1, 2, 3
BRDF_REFERENCE, BRDF_LAMBERT_GGX = 0, 1
RNG_PHILOX, RNG_REFERENCE_LCG = 0, 1  # rtf_set_option(ctx, "rng", ...): rtf.h


class EnvironmentProperties(C.Structure):
    """Defaults = RayTracer.hpp:104-126."""
    _fields_ = [("lighting_type", i32), ("skylight_intensity", f32), ("ambient_light_intensity", f32),
                ("light_size", f32), ("gamma", f32), ("sun_direction", f32 * 3), ("color", f32 * 3),
                ("environmental_map_id", u32), ("environmental_maps", u64 * 6), ("atmosphere", Atmosphere),
                ("pad_", u32)]

    def __init__(self, **kw):
        super().__init__()
        self.lighting_type = ENV_SCENE
        self.skylight_intensity = 1.0
        self.ambient_light_intensity = 0.1
        self.light_size = 0.0
        self.gamma = 1.0
        self.sun_direction[:] = (0.0, 1.0, 0.0)
        self.color[:] = (1.0, 1.0, 1.0)
        self.atmosphere = Atmosphere(6360.0, 6420.0, 7994.0, 1200.0, 0.76, 16, 8)
        for k, v in kw.items():
            if isinstance(v, (tuple, list)):
                getattr(self, k)[:] = v
            else:
                setattr(self, k, v)


class RayProperties(C.Structure):
    """Defaults = RayTracer.hpp:154-155."""
    _fields_ = [("bounces", i32), ("samples", i32)]

    def __init__(self, bounces=4, samples=1):
        super().__init__(bounces, samples)


class Material(C.Structure):
    _fields_ = [("material_type", i32), ("albedo_color", f32 * 3), ("subsurface_color", f32 * 3),
                ("subsurface_factor", f32), ("subsurface_radius", f32 * 3), ("metallic", f32), ("roughness", f32),
                ("emission", f32), ("transmission", f32), ("pad_", f32)]

    def __init__(self, albedo=(1.0, 1.0, 1.0), metallic=0.0, roughness=1.0, emission=0.0, transmission=0.0,
                 material_type=MAT_DEFAULT):
        super().__init__()
        self.material_type = material_type
        self.albedo_color[:] = albedo
        self.subsurface_color[:] = (1.0, 1.0, 1.0)
        self.subsurface_factor = 0.0
        self.subsurface_radius[:] = (1.0, 0.2, 0.1)
        self.metallic, self.roughness, self.emission, self.transmission = metallic, roughness, emission, transmission


class CameraProps(C.Structure):
    _fields_ = [("accumulate", i32), ("fov", f32), ("from_", f32 * 3), ("horizontal", f32 * 3), ("vertical", f32 * 3),
                ("inverse_projection_view", f32 * 16), ("max_distance", f32), ("output_type", i32), ("gamma", f32),
                ("width", i32), ("height", i32), ("frame_id", u64), ("modified", i32), ("aperture", f32),
                ("focal_length", f32)]


class ScanGrid(C.Structure):
    _fields_ = [("center", f32 * 3), ("left", f32 * 3), ("up", f32 * 3), ("direction", f32 * 3), ("size", f32 * 2),
                ("distance", f32 * 2)]


class JobStats(C.Structure):
    _fields_ = [("rays", u64), ("nodes", u64), ("tris", u64), ("paths", u64)]


assert C.sizeof(Vertex) == 80
assert C.sizeof(IlluminationSampler) == 268
assert C.sizeof(HitInfo) == 76
assert C.sizeof(PointCloudSample) == 112
assert C.sizeof(EnvironmentProperties) == 128
assert C.sizeof(Material) == 64

# numpy views of the same layouts (zero-copy bulk access)
VERTEX_DTYPE = np.dtype([("position", "<f4", 3), ("position_padding", "<f4"), ("normal", "<f4", 3),
                         ("normal_padding", "<f4"), ("tangent", "<f4", 3), ("tangent_padding", "<f4"),
                         ("color", "<f4", 4), ("tex_coord", "<f4", 2), ("tex_coord_padding", "<f4", 2)])
ILLUM_SAMPLER_DTYPE = np.dtype([("a", VERTEX_DTYPE), ("b", VERTEX_DTYPE), ("c", VERTEX_DTYPE), ("direction", "<f4", 3),
                                ("energy", "<f4", 3), ("front_face", "u1"), ("back_face", "u1"), ("pad_", "u1", 2)])
HIT_INFO_DTYPE = np.dtype([("position", "<f4", 3), ("normal", "<f4", 3), ("tangent", "<f4", 3), ("color", "<f4", 4),
                           ("tex_coord", "<f4", 2), ("data", "<f4", 4)])
PC_SAMPLE_DTYPE = np.dtype([("direction", "<f4", 3), ("start", "<f4", 3), ("handle", "<u8"), ("hit", "u1"),
                            ("pad_", "u1", 3), ("hit_info", HIT_INFO_DTYPE)])
HIT_DTYPE = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("prim", "<u4")])
NODE_DTYPE = np.dtype([("p", "<f4", 3), ("e", "u1", 3), ("imask", "u1"), ("child_base", "<u4"), ("tri_base", "<u4"),
                       ("meta", "u1", 8), ("qlo", "u1", (3, 8)), ("qhi", "u1", (3, 8))])
TRI_DTYPE = np.dtype([("v0", "<f4", 3), ("prim", "<u4"), ("v1", "<f4", 3), ("inst", "<u4"), ("v2", "<f4", 3),
                      ("flags", "<u4")])
# UniEngine::SkinnedVertex (layout inferred from CopySkinnedVerticesKernel, RayTracer.cu:1194-1247): 144 B
SKINNED_VERTEX_DTYPE = np.dtype([("v", VERTEX_DTYPE), ("bone_id", "<i4", 4), ("weight", "<f4", 4), ("bone_id2", "<i4", 4),
                                 ("weight2", "<f4", 4)])
assert VERTEX_DTYPE.itemsize == 80 and ILLUM_SAMPLER_DTYPE.itemsize == 268 and SKINNED_VERTEX_DTYPE.itemsize == 144


def node_leaf_slots(node):
    """Leaf slots of one wide node as (slot, offset of its first triangle from tri_base, triangle count); node meta
    (csrc/rtf_bvh.cuh): word 0 = 24-bit valid mask (bit 3s+j: slot s holds more than j triangles), word 1 = imask |
    eight 2-bit triangle counts << 8; a node's leaf triangles lie consecutively in slot order."""
    m = [int(x) for x in node["meta"]]
    w0 = m[0] | m[1] << 8 | m[2] << 16 | m[3] << 24
    w1 = m[4] | m[5] << 8 | m[6] << 16 | m[7] << 24
    assert (w1 & 0xff) == int(node["imask"]) and (w0 >> 24) == 0 and (w1 >> 24) == 0
    out, off = [], 0
    for s in range(8):
        cnt = (w1 >> (8 + 2 * s)) & 3
        assert ((w0 >> (3 * s)) & 7) == (0, 1, 3, 7)[cnt]
        assert not (cnt and (int(node["imask"]) >> s) & 1)
        if cnt:
            out.append((s, off, cnt))
            off += cnt
    return out


assert PC_SAMPLE_DTYPE.itemsize == 112 and NODE_DTYPE.itemsize == 80 and TRI_DTYPE.itemsize == 48

DBG_MORTON, DBG_SORTED_IDS, DBG_BIN_CHILDREN, DBG_BIN_PARENT, DBG_BIN_BOXES = 0, 1, 2, 3, 4
DBG_WIDE_NODES, DBG_WIDE_TRIS, DBG_FLAT_POS, DBG_FLAT_NRM, DBG_FLAT_INST, DBG_SCENE_BOUNDS = 5, 6, 7, 8, 9, 10
DBG_REF_TRI, DBG_REF_BOX = 11, 12
DEFAULT_PRESPLIT = 8  # librtf_b200's default for rtf_set_option("presplit")
DBG_DTYPES = {DBG_REF_TRI: np.dtype("<u4"), DBG_REF_BOX: np.dtype("<f4"),
              DBG_MORTON: np.dtype("<u8"), DBG_SORTED_IDS: np.dtype("<u4"), DBG_BIN_CHILDREN: np.dtype("<i4"),
              DBG_BIN_PARENT: np.dtype("<i4"), DBG_BIN_BOXES: np.dtype("<f4"), DBG_WIDE_NODES: NODE_DTYPE,
              DBG_WIDE_TRIS: TRI_DTYPE, DBG_FLAT_POS: np.dtype("<f4"), DBG_FLAT_NRM: np.dtype("<f4"),
              DBG_FLAT_INST: np.dtype("<u4"), DBG_SCENE_BOUNDS: np.dtype("<f4")}


class BTFDesc(C.Structure):
    """rtf_btf_desc (include/rtf/rtf.h): scalars + host pointers to the 13 tables of CompressedBTF::ImportFromFolder
    (/root/reference/src/RayTracerFacility/MLVQ/CompressedBTF.cpp:339-377)."""
    _fields_ = [("use_cos_beta", i32), ("num_beta", i32), ("num_alpha", i32), ("num_theta", i32), ("num_phi", i32),
                ("code_btf_flag", i32), ("hdr", i32), ("hdr_value", f32), ("multiplier", f32),
                ("tex_coord_multiplier", f32), ("gamma", f32), ("rows_offset", i32), ("cols_offset", i32),
                ("pdf6_rows", i32), ("pdf6_cols", i32), ("n_pdf1d", i32), ("n_colors", i32), ("n_channels", i32),
                ("n_index_slices", i32), ("n_pdf2d_color", i32), ("n_pdf2d_lum", i32), ("n_pdf2d", i32),
                ("pdf2d_slice_len", i32), ("n_pdf3d", i32), ("n_pdf4d", i32),
                ("pdf6_slices", C.c_void_p), ("pdf6_scales", C.c_void_p), ("pdf1d_basis", C.c_void_p),
                ("colors", C.c_void_p), ("index_ab", C.c_void_p), ("pdf2d_colors", C.c_void_p),
                ("pdf2d_slices", C.c_void_p), ("pdf2d_scales", C.c_void_p), ("pdf2d_indices", C.c_void_p),
                ("pdf3d_scales", C.c_void_p), ("pdf3d_slices", C.c_void_p), ("pdf4d_scales", C.c_void_p),
                ("pdf4d_slices", C.c_void_p)]


BTF_INT_TABLES = ("pdf6_slices", "index_ab", "pdf2d_colors", "pdf2d_slices", "pdf2d_indices", "pdf3d_slices", "pdf4d_slices")
BTF_FLOAT_TABLES = ("pdf6_scales", "pdf1d_basis", "colors", "pdf2d_scales", "pdf3d_scales", "pdf4d_scales")


class BTFTables:
    """A compressed-BTF database as numpy tables (2-D, row-major) + scalar parameters; desc() builds the ABI record
    (the arrays stay referenced by this object)."""

    def __init__(self, tables, use_cos_beta=True, code_btf_flag=0, hdr=False, hdr_value=1.0, multiplier=1.0,
                 tex_coord_multiplier=1.0, gamma=2.2, rows_offset=0, cols_offset=0):
        self.t = {}
        for k in BTF_INT_TABLES:
            self.t[k] = np.ascontiguousarray(tables[k], dtype=np.int32)
        for k in BTF_FLOAT_TABLES:
            self.t[k] = np.ascontiguousarray(tables[k], dtype=np.float32)
        self.params = dict(use_cos_beta=int(use_cos_beta), code_btf_flag=int(code_btf_flag), hdr=int(hdr),
                           hdr_value=float(hdr_value), multiplier=float(multiplier),
                           tex_coord_multiplier=float(tex_coord_multiplier), gamma=float(gamma),
                           rows_offset=int(rows_offset), cols_offset=int(cols_offset))

    def desc(self):
        t, d = self.t, BTFDesc()
        for k, v in self.params.items():
            setattr(d, k, v)
        d.num_beta, d.num_alpha = t["pdf1d_basis"].shape[1], t["pdf2d_slices"].shape[1]
        d.num_theta, d.num_phi = t["pdf3d_slices"].shape[1], t["pdf4d_slices"].shape[1]
        d.pdf6_rows, d.pdf6_cols = t["pdf6_slices"].shape
        d.n_pdf1d = t["pdf1d_basis"].shape[0]
        d.n_colors, d.n_channels = t["colors"].shape
        d.n_index_slices = t["index_ab"].shape[0]
        d.n_pdf2d_color = t["pdf2d_colors"].shape[0]
        d.n_pdf2d_lum = t["pdf2d_slices"].shape[0]
        d.n_pdf2d, d.pdf2d_slice_len = t["pdf2d_indices"].shape
        d.n_pdf3d = t["pdf3d_slices"].shape[0]
        d.n_pdf4d = t["pdf4d_slices"].shape[0]
        for k in BTF_INT_TABLES + BTF_FLOAT_TABLES:
            setattr(d, k, t[k].ctypes.data)
        return d
