/*
 * rtf.h — C ABI of librtf_b200.so, the B200-native ray-tracing core that sits behind
 * RayTracerFacility's module façade (layer L3 in SURVEY.md §1).
 *
 * Every entry point names the reference interface it replaces (paths relative to the
 * reference checkout):
 *   RT.hpp   = include/RayTracerFacility/RayTracer/RayTracer.hpp
 *   CM.hpp   = include/RayTracerFacility/RayTracer/CUDAModule.hpp
 *   CM.cpp   = src/RayTracerFacility/CUDAModule.cpp
 *   RT.cu    = src/RayTracerFacility/RayTracer.cu
 *   PCS.cpp  = src/RayTracerFacility/PointCloudScanner.cpp
 *
 * Conventions: plain pointers and sizes only; every function returns an int status
 * (RTF_OK == 0, negative on error) and never throws or exits (the reference throws on CUDA
 * errors and exit(2)s on OptiX errors, Optix7.hpp:10-56); rtf_last_error() returns a
 * thread-local message.  One host thread drives one context.  All host pointers are plain
 * (pageable or pinned) memory; `_device` variants take device pointers of the context's GPU.
 */
#ifndef RTF_RTF_H
#define RTF_RTF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define RTF_API __declspec(dllexport)
#else
#define RTF_API __attribute__((visibility("default")))
#endif

/* ---- status codes -------------------------------------------------------------------- */
#define RTF_OK 0
#define RTF_ERR_INVALID (-1)   /* bad argument                                              */
#define RTF_ERR_CUDA (-2)      /* CUDA runtime error (reference: throw, Optix7.hpp:10-22)   */
#define RTF_ERR_NOT_READY (-3) /* no acceleration structure yet (reference: RT.cu:248,610,760) */
#define RTF_ERR_NOT_FOUND (-4) /* unknown handle                                            */
#define RTF_ERR_LIMIT (-5)     /* internal capacity exceeded (e.g. BVH deeper than the stack) */

/* ---- reference-layout POD records ---------------------------------------------------- */

/* UniEngine::Vertex, 80 B.  Layout inferred from its uses at RT.cu:1163-1173, RDD.hpp:132-161. */
typedef struct rtf_vertex {
  float position[3];
  float position_padding;
  float normal[3];
  float normal_padding;
  float tangent[3];
  float tangent_padding;
  float color[4];
  float tex_coord[2];
  float tex_coord_padding[2];
} rtf_vertex;

/* IlluminationSampler<glm::vec3>, RT.hpp:187-211, 268 B, in-out. */
typedef struct rtf_illum_sampler {
  rtf_vertex a, b, c;
  float direction[3]; /* out */
  float energy[3];    /* out */
  uint8_t front_face;
  uint8_t back_face;
  uint8_t pad_[2];
} rtf_illum_sampler;

/* HitInfo, HitInfo.hpp:4-11, 76 B. */
typedef struct rtf_hit_info {
  float position[3];
  float normal[3];
  float tangent[3];
  float color[4];
  float tex_coord[2];
  float data[4];
} rtf_hit_info;

/* PointCloudSample, RT.hpp:222-232, 112 B, in-out. */
typedef struct rtf_pc_sample {
  float direction[3]; /* in  */
  float start[3];     /* in  */
  uint64_t handle;    /* out: instance's private-component handle (RT.cu:1859) */
  uint8_t hit;        /* out */
  uint8_t pad_[3];
  rtf_hit_info hit_info; /* out */
} rtf_pc_sample;

/* EnvironmentalLightingType, RT.hpp:99-101 */
enum { RTF_ENV_SCENE = 0, RTF_ENV_SKYDOME = 1, RTF_ENV_SINGLE_LIGHT_SOURCE = 2 };

/* EnvironmentProperties, RT.hpp:103-151.  Numeric fields keep the reference order; the
 * GL cube-map id keeps its meaning "non-zero = sample the cube map" (Env.cuh:153-158); the six
 * faces are linear-memory images registered with rtf_set_environment_map (no GL on this path);
 * the cudaTextureObject_t[6] slot is kept for layout and ignored. */
typedef struct rtf_env {
  int32_t lighting_type;
  float skylight_intensity;
  float ambient_light_intensity;
  float light_size;
  float gamma;
  float sun_direction[3];
  float color[3];
  uint32_t environmental_map_id; /* != 0: Scene lighting samples the cube map (rtf_set_environment_map) */
  uint64_t environmental_maps[6];
  struct {
    float earth_radius;
    float atmosphere_radius;
    float hr;
    float hm;
    float g;
    int32_t num_samples;
    int32_t num_samples_light;
  } atmosphere;
  uint32_t pad_;
} rtf_env;

/* RayProperties, RT.hpp:153-163 */
typedef struct rtf_ray_props {
  int32_t bounces;
  int32_t samples;
} rtf_ray_props;

/* MaterialType, Enums.hpp:4-8 */
enum { RTF_MAT_DEFAULT = 0, RTF_MAT_VERTEX_COLOR = 1, RTF_MAT_COMPRESSED_BTF = 2 };

/* The MaterialProperties fields the path reads (RF.cuh:97-98,122-123,152,170; RDD.hpp:250-264)
 * + RayTracedMaterial::m_materialType (RT.hpp:248).
 * BSSRDF (RF.cuh:97-133, BSSDF.cuh:42-89): with subsurface_factor > 0 and subsurface_radius[0] > 0 the
 * FIRST hit of a path probes the same object (private component handle) from a disk of that radius
 * above the surface; if the drawn record lies within the radius, a second radiance path leaves the
 * surface there (weight factor*subsurface_color*k), traced BEFORE the ordinary bounce, and — as in the
 * reference, whose two paths share PerRayData::m_hitCount — the bounce path continues from the hit
 * count the subsurface path ended with.  Deviation (documented in csrc/rtf_sss.cuh): the reference's
 * record set depends on the order in which OptiX reports any-hit candidates; here it is the up-to-four
 * nearest same-handle intersections, which is a function of the scene alone. */
typedef struct rtf_material {
  int32_t material_type;
  float albedo_color[3];
  float subsurface_color[3];
  float subsurface_factor;
  float subsurface_radius[3];
  float metallic;
  float roughness;
  float emission;
  float transmission;
  float pad_;
} rtf_material;

/* OutputType, RT.hpp:26-28 */
enum { RTF_OUT_COLOR = 0, RTF_OUT_NORMAL = 1, RTF_OUT_ALBEDO = 2, RTF_OUT_DEPTH = 3 };

/* The numeric part of CameraProperties (RT.hpp:30-96).  Frame buffers live in headless
 * device float4 arrays owned by the camera object (GL texture target: unavailable). */
typedef struct rtf_camera_props {
  int32_t accumulate; /* m_accumulate */
  float fov;          /* m_fov, degrees; halved by Set() like RT.cu:32 */
  float from[3];
  float horizontal[3];
  float vertical[3];
  float inverse_projection_view[16]; /* column-major glm::mat4 */
  float max_distance;
  int32_t output_type;
  float gamma;
  int32_t width, height;
  uint64_t frame_id;
  int32_t modified;
  float aperture;
  float focal_length;
} rtf_camera_props;

/* Parameters of PointCloudScanner::Scan (PCS.cpp:70-92) so that 10^8-ray scans never
 * materialise 112 B/ray on the host. */
typedef struct rtf_scan_grid {
  float center[3];
  float left[3];
  float up[3];
  float direction[3]; /* normalised scan direction (actualVector) */
  float size[2];      /* m_size */
  float distance[2];  /* m_distance */
} rtf_scan_grid;

/* BRDF model of the surface materials, rtf_set_option(ctx, "brdf", ...).  RTF_BRDF_REFERENCE = mirror lobe of
 * BSDF.cuh:6-13 + RF.cuh:152-161 (the reference's only model).  RTF_BRDF_LAMBERT_GGX (BASELINE config 2) = importance-
 * sampled Lambert + GGX (Schlick Fresnel, Smith G); a new capability, checked against its own closed forms only. */
enum { RTF_BRDF_REFERENCE = 0, RTF_BRDF_LAMBERT_GGX = 1 };
/* Random numbers of rtf_estimate_illumination and rtf_render_camera[_rows], rtf_set_option(ctx, "rng", ...).  RTF_RNG_PHILOX (default, every entry
 * point): counter-based Philox-4x32-10 keyed (seed, stream, unit, bounce, primitive) — a draw does not depend on what other
 * rays drew, which is what lets samples run in parallel and results be independent of batching and GPU count.
 * RTF_RNG_REFERENCE_LCG (verification mode; the AoS illumination entry point and the single-context camera calls): the
 * reference's own generator (LinearCongruenceGenerator.hpp:6-35), one stream per probe seeded (probe index, seed) as
 * __raygen__IE does (IlluminationEstimation.cu:47), one per pixel seeded (pixel, frame id) as __raygen__CR does
 * (CameraRendering.cu:42-44), consumed in the reference's order — the samples of a probe / pixel run one after the other, so
 * it is slow; results are comparable with the reference's programs directly (tests/test_gpu_lcg.py).  Scenes with a
 * subsurface material and the sharded camera calls are refused. */
enum { RTF_RNG_PHILOX = 0, RTF_RNG_REFERENCE_LCG = 1 };

typedef struct rtf_ctx rtf_ctx;
typedef struct rtf_camera rtf_camera;

/* ---- lifecycle: CudaModule::Init / Terminate (CM.hpp:30-36, CM.cpp:23-36) -------------- */
RTF_API int rtf_create(int device, rtf_ctx** out);
RTF_API int rtf_destroy(rtf_ctx* ctx);
RTF_API const char* rtf_last_error(void);
RTF_API const char* rtf_version(void);
/* Use an externally owned CUDA stream for all work.  NULL = a private non-blocking stream of the context (NOT the legacy
 * default stream: torch's default stream handle is NULL too — pass a real stream to share one).  Stream contract: every
 * *_device entry point only ENQUEUES on the context's stream; a caller that reads or overwrites the buffers from another
 * stream (a collective, a zero-fill) must order the two itself (share the stream, or events).  Host-buffer entry points
 * block until their results are in host memory. */
RTF_API int rtf_set_stream(rtf_ctx* ctx, void* cuda_stream);
/* Number of kernel launches issued by the context so far (bench.py's gpu_launches claim). */
RTF_API uint64_t rtf_launch_count(rtf_ctx* ctx);

/* ---- scene sync: RayTracer::m_geometries / m_materials / m_instances (RT.hpp:361-363),
 *      filled by RayTracerLayer::UpdateMeshesStorage (RayTracerLayer.cpp:18-346) ----------- */
/* Copies immediately (the reference borrows host vectors until BuildGAS, RT.hpp:280-292). */
RTF_API int rtf_set_geometry(rtf_ctx* ctx, uint64_t handle, const void* vertices, size_t n_vertices,
                             size_t vertex_stride_bytes, const uint32_t* triangles, size_t n_triangles,
                             uint64_t version);
/* RendererType::Instanced geometry (RT.hpp:277-292): one mesh drawn under n_instances matrices.  The
 * device pre-kernel restates CopyVerticesInstancedKernel (RT.cu:1148-1175): vertex k of instance i =
 * {position = M_i*(p,1), N = normalize(M_i*(n,0)), T = normalize(M_i*(t,0)) then
 * T = normalize(T - dot(T,N)*N), texCoord and colour copied, the instance colour in the four padding
 * floats (position/normal/tangent padding, tex_coord_padding[0] -> HitInfo::m_data) and the instance index
 * in tex_coord_padding[1]}; triangle j of instance i = mesh triangle j + i*n_vertices (RT.cu:1477-1487).
 * The resulting geometry has n_instances*n_triangles triangles and is referenced by rtf_set_instance
 * like any other.  matrices: 16 floats each, column-major; colors: 4 floats each, NULL = (1,1,1,1).
 * Waived quirk: the reference's index vector is resize(N) THEN push_back (RT.cu:1477-1483), so its
 * first N triples are the degenerate triangle (0,0,0) — they can never be hit (zero determinant)
 * and are not emitted here, i.e. primitive numbering starts at the first real triangle. */
RTF_API int rtf_set_geometry_instanced(rtf_ctx* ctx, uint64_t handle, const void* vertices, size_t n_vertices,
                                       size_t vertex_stride_bytes, const uint32_t* triangles, size_t n_triangles,
                                       const float* matrices, const float* colors, size_t n_instances,
                                       uint64_t version);
/* RendererType::Skinned geometry: device pre-kernel restating CopySkinnedVerticesKernel
 * (RT.cu:1194-1247): boneTransform = sum_k bone[id_k]*w_k over the up-to-8 influences (the first always,
 * the others when their id != -1), then the same position / N / T arithmetic as above; paddings,
 * texCoord and colour copied.  rtf_skinned_vertex is UniEngine::SkinnedVertex (layout inferred, 144 B). */
typedef struct rtf_skinned_vertex {
  rtf_vertex v;
  int32_t bone_id[4];
  float weight[4];
  int32_t bone_id2[4];
  float weight2[4];
} rtf_skinned_vertex;
RTF_API int rtf_set_geometry_skinned(rtf_ctx* ctx, uint64_t handle, const void* skinned_vertices, size_t n_vertices,
                                     size_t vertex_stride_bytes, const uint32_t* triangles, size_t n_triangles,
                                     const float* bone_matrices, size_t n_bones, uint64_t version);
/* Test hook: the baked (device-resident) vertices / indices of an instanced or skinned geometry. */
RTF_API int rtf_geometry_read_baked(rtf_ctx* ctx, uint64_t handle, rtf_vertex* vertices_out, size_t n_vertices,
                                    uint32_t* triangles_out, size_t n_triangles);
RTF_API int rtf_set_material(rtf_ctx* ctx, uint64_t handle, const rtf_material* material);
RTF_API int rtf_set_instance(rtf_ctx* ctx, uint64_t handle, uint64_t geometry_handle, uint64_t material_handle,
                             const float transform[16] /* column-major glm::mat4 m_globalTransform */,
                             uint64_t private_component_handle);
RTF_API int rtf_remove_geometry(rtf_ctx* ctx, uint64_t handle);
RTF_API int rtf_remove_material(rtf_ctx* ctx, uint64_t handle);
RTF_API int rtf_remove_instance(rtf_ctx* ctx, uint64_t handle);
/* RayTracer::BuildIAS (RT.hpp:386, RT.cu:1618-1715) + BuildGAS (RT.cu:1251-1598):
 * flatten instances (ascending instance handle) to a world-space triangle soup and build the
 * LBVH -> 8-wide compressed BVH on the device.  morton_bits: 30 or 63 (0 = 63). */
RTF_API int rtf_commit(rtf_ctx* ctx, int morton_bits);
RTF_API int rtf_triangle_count(rtf_ctx* ctx, uint64_t* n_triangles);
/* Scene-sync diffing (the reference's m_updateFlag / m_version bookkeeping, RayTracerLayer.cpp:47-92,
 * RT.hpp:305-327, done on this side of the ABI): rtf_set_geometry with a non-zero version equal to the
 * stored one, and rtf_set_material / rtf_set_instance with identical contents, are no-ops; rtf_commit
 * rebuilds the BVH only if a geometry, an instance or "presplit" changed, re-resolves the material-side
 * tables only (materials, textures, BTFs, cube map — microseconds) if nothing else did, and returns at once
 * if nothing changed.  With option "refit" = 1, a commit in which only vertex positions or instance transforms
 * moved (same geometries, instances and triangle counts) takes the REFIT path: the sorted reference order
 * and the radix tree of the last full build are reused, reference boxes, node boxes, the 8-wide collapse
 * and the triangle records are recomputed (no bounds / split planning / Morton / sort / Karras).  Hits do
 * not depend on the tree; its quality decays as the scene deforms, so rebuild fully from time to time
 * (any topology change does).  out5 = {full builds, material-only commits, skipped commits, skipped
 * geometry uploads, refits}. */
RTF_API int rtf_commit_stats(rtf_ctx* ctx, uint64_t* out5);
/* Device time (ms, CUDA events) of the last rtf_commit per stage:
 * {flatten, bounds+morton, sort, karras+refit, collapse+encode, total}; n >= 6.  (The reference
 * has no counterpart: its builds are opaque optixAccelBuild calls, RT.cu:1562-1584.) */
RTF_API int rtf_build_timings(rtf_ctx* ctx, float* ms_out, int n);
/* Tuning knobs (not in the reference): "batch_paths" = wavefront batch capacity in paths,
 * "sort_rays" = 0/1 sort bounce rays by (direction octant, origin cell) before tracing (default 0:
 * measured slower than the unsorted persistent-lane traversal; results are bit-identical),
 * "presplit" = most references a long thin triangle may be cut into before the Morton sort
 * (default 8, 1 = off; applies at the next rtf_commit; hits do not depend on it),
 * "brdf" = RTF_BRDF_REFERENCE (default) or RTF_BRDF_LAMBERT_GGX,
 * "rng" = RTF_RNG_PHILOX (default) or RTF_RNG_REFERENCE_LCG (rtf_estimate_illumination, rtf_render_camera[_rows]),
 * "refit" = 0/1 allow the refit path of rtf_commit (see rtf_commit_stats; default 0). */
RTF_API int rtf_set_option(rtf_ctx* ctx, const char* name, int64_t value);

/* ---- images in linear memory ---------------------------------------------------------------
 * Replace the GL texture ids of RayTracedMaterial (RT.hpp:254-263; bound by BindTexture,
 * RT.cu:2233-2256) and the cube-map faces of EnvironmentProperties (RT.hpp:112-113; RT.cu:270-347).
 * Sampling restates tex2D<float4> under the reference's descriptor (cudaAddressModeWrap,
 * cudaFilterModeLinear, normalizedCoords, RT.cu:2246-2252) with full float weights (the texture
 * unit keeps 8 fractional bits).  Texels: width*height float4, row-major, row 0 at v = 0.
 * Changes take effect at the next rtf_commit. */
RTF_API int rtf_set_texture(rtf_ctx* ctx, uint64_t handle, int width, int height, const float* rgba);
RTF_API int rtf_remove_texture(rtf_ctx* ctx, uint64_t handle);
/* RayTracedMaterial::m_albedoTexture / m_normalTexture / m_metallicTexture / m_roughnessTexture
 * (RT.hpp:254-263; SurfaceMaterial::GetAlbedo/GetRoughness/GetMetallic/ApplyNormalTexture,
 * RDD.hpp:248-281).  0 = none. */
RTF_API int rtf_set_material_textures(rtf_ctx* ctx, uint64_t material_handle, uint64_t albedo, uint64_t normal,
                                      uint64_t metallic, uint64_t roughness);
/* Cube map of the Scene lighting type: six texture handles in the order +X -X +Y -Y +Z -Z
 * (SampleCubeMap's faceIndex, RTU.cuh:54-80); NULL clears it. */
RTF_API int rtf_set_environment_map(rtf_ctx* ctx, const uint64_t face_textures[6]);
/* Test hooks: n bilinear samples of a texture (uv: 2 floats each -> rgba: 4 floats each); n
 * evaluations of CalculateEnvironmentalLight (Env.cuh:147-175) for {position.xyz, direction.xyz}
 * -> rgb.  Host buffers. */
RTF_API int rtf_sample_texture(rtf_ctx* ctx, uint64_t handle, size_t n, const float* uv, float* rgba_out);
RTF_API int rtf_sample_environment(rtf_ctx* ctx, const rtf_env* env, size_t n, const float* pos_dir6, float* rgb_out);

/* ---- compressed BTF material (MLVQ), CompressedBTF.hpp / BTFBase.cuh:13-81 ------------------
 * The 13 tables of CompressedBTF::ImportFromFolder (CompressedBTF.cpp:339-377) as host arrays,
 * row-major, with the row/column counts the loader reads from each file's header. */
typedef struct rtf_btf_desc {
  int32_t use_cos_beta;   /* SharedCoordinates::m_useCosBeta */
  int32_t num_beta, num_alpha, num_theta, num_phi;
  int32_t code_btf_flag;  /* SharedCoordinates::m_codeBtfFlag (12th parameterisation int, CompressedBTF.cpp:197) */
  int32_t hdr;            /* BTFBase::m_hdr / SharedCoordinates::m_hdrFlag */
  float hdr_value, multiplier, tex_coord_multiplier, gamma; /* BTFBase.cuh:19-24 */
  int32_t rows_offset, cols_offset;       /* PDF6D::m_rowsOffset / m_colsOffset */
  int32_t pdf6_rows, pdf6_cols;           /* <mat>_PDF6Dslices.txt / _PDF6Dscale.txt */
  int32_t n_pdf1d;                        /* all_PDF1Dslice.txt: n_pdf1d x num_beta floats */
  int32_t n_colors, n_channels;           /* all_colors.txt */
  int32_t n_index_slices;                 /* all_indexAB.txt: x num_beta ints */
  int32_t n_pdf2d_color;                  /* all_PDF2Dcolours.txt: x num_alpha ints */
  int32_t n_pdf2d_lum;                    /* all_PDF2Dslices.txt / all_PDF2Dscale.txt: x num_alpha */
  int32_t n_pdf2d, pdf2d_slice_len;       /* all_PDF2Dindices.txt: n_pdf2d x 2 {luminance, colour} */
  int32_t n_pdf3d;                        /* all_PDF3Dslices.txt / _PDF3Dscale.txt: x num_theta */
  int32_t n_pdf4d;                        /* all_PDF4Dslices.txt / _PDF4Dscale.txt: x num_phi */
  const int32_t* pdf6_slices;
  const float* pdf6_scales;
  const float* pdf1d_basis;
  const float* colors;
  const int32_t* index_ab;
  const int32_t* pdf2d_colors;
  const int32_t* pdf2d_slices;
  const float* pdf2d_scales;
  const int32_t* pdf2d_indices;
  const float* pdf3d_scales;
  const int32_t* pdf3d_slices;
  const float* pdf4d_scales;
  const int32_t* pdf4d_slices;
} rtf_btf_desc;
/* Validates every index table against the sizes (the reference only asserts) and uploads. */
RTF_API int rtf_set_btf(rtf_ctx* ctx, uint64_t handle, const rtf_btf_desc* desc);
/* CompressedBTF::ImportFromFolder (CompressedBTF.cpp:74-390): all_materialInfo.txt,
 * <material>_materialInfo.txt and the 13 "rows cols min max" ASCII tables. */
RTF_API int rtf_btf_import_folder(rtf_ctx* ctx, uint64_t handle, const char* folder);
RTF_API int rtf_remove_btf(rtf_ctx* ctx, uint64_t handle);
/* A material of type RTF_MAT_COMPRESSED_BTF shades with this table set (SBT::m_material ->
 * SurfaceCompressedBTF, RT.cu:1888-1894). */
RTF_API int rtf_set_material_btf(rtf_ctx* ctx, uint64_t material_handle, uint64_t btf_handle);
/* Test hook = SurfaceCompressedBTF::GetValue (RDD.hpp:329-354) for n host-side queries:
 * tex_coord 2, view_dir 3 (direction of the incoming ray), illum_dir 3, normal 3, tangent 3 -> rgb 3. */
RTF_API int rtf_btf_eval(rtf_ctx* ctx, uint64_t btf_handle, size_t n, const float* tex_coord, const float* view_dir,
                         const float* illum_dir, const float* normal, const float* tangent, float* rgb_out);

/* ---- illumination: CudaModule::EstimateIlluminationRayTracing (CM.hpp:38-41, CM.cpp:39-50)
 *      -> RayTracer::EstimateIllumination (RT.cu:605-756) -> __raygen__IE (IE.cu:30-129) ---- */
RTF_API int rtf_estimate_illumination(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* ray_props,
                                      rtf_illum_sampler* probes /* host, in-out */, size_t n, uint32_t seed,
                                      float push_normal_distance);
/* Scale variant: the probes are the committed world-space triangles themselves
 * (TriangleIlluminationEstimator::PrepareLightProbeGroup, TriangleIlluminationEstimator.cpp:94-139),
 * global triangle order = ascending instance handle, then mesh triangle order.
 * Traces samples [sample_begin, sample_end) of ray_props->samples for triangles [tri_begin, tri_end).
 * out6: host, 6 floats per triangle in the range {E.x,E.y,E.z,D.x,D.y,D.z}.
 * normalise != 0: energy = sum/(faces*samples), direction = normalize(D) (IE.cu:125-128);
 * normalise == 0: raw partial sums (for cross-GPU sample sharding, summed then finalised). */
RTF_API int rtf_estimate_illumination_mesh(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* ray_props,
                                           uint32_t seed, float push_normal_distance, int front_face, int back_face,
                                           uint64_t tri_begin, uint64_t tri_end, int sample_begin, int sample_end,
                                           int normalise, float* out6);
/* Same, result accumulated (+=) into a device array of 6 floats per committed triangle (all triangles). */
RTF_API int rtf_estimate_illumination_mesh_device(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* ray_props,
                                                  uint32_t seed, float push_normal_distance, int front_face,
                                                  int back_face, uint64_t tri_begin, uint64_t tri_end,
                                                  int sample_begin, int sample_end, float* dev_accum6);
/* energy = sum/(faces*samples); direction = normalize(D) over a device accumulator (IE.cu:125-128). */
RTF_API int rtf_illumination_finalize_device(rtf_ctx* ctx, float* dev_accum6, uint64_t n_triangles, int faces,
                                             int samples);
/* TriangleIlluminationEstimator::PrepareLightProbeGroup (TriangleIlluminationEstimator.cpp:94-139) on the
 * device: the 268 B probes of committed triangles [tri_begin, tri_end) — the three mesh vertices with
 * position = M*(p,1) and normal = M*(n,0), direction = energy = 0, the two face flags as given (the
 * caller maps Material::m_drawSettings cull modes to them, :120-135). */
RTF_API int rtf_prepare_light_probes(rtf_ctx* ctx, uint64_t tri_begin, uint64_t tri_end, int front_face,
                                     int back_face, rtf_illum_sampler* probes_out /* host */);
/* TriangleIlluminationEstimator::SampleLightProbeGroup's flux integration (:84-92):
 * out8 = {totalFlux.xyz = sum energy*GetArea() (Heron, RT.hpp:204-210), totalArea (a two-sided probe
 * counts twice, :124), averageFlux.xyz = totalFlux/totalArea, 0}.  The reference adds 10^7 floats
 * serially on the host; here the terms are added in double by a fixed-shape tree (1024-term tiles:
 * 4 strided terms per thread, xor butterfly, 8 warp sums in order; tiles reduced the same way), so the
 * result does not depend on the GPU count or batch size and the oracle reproduces it bit for bit.
 * rtf_illumination_flux: host probes (as returned by rtf_estimate_illumination).
 * rtf_illumination_flux_mesh_device: finalised device array of 6 floats per committed triangle (the
 * mesh-range entry point's output), areas from the committed world-space triangles. */
RTF_API int rtf_illumination_flux(rtf_ctx* ctx, const rtf_illum_sampler* probes /* host */, size_t n, float* out8);
RTF_API int rtf_illumination_flux_mesh_device(rtf_ctx* ctx, const float* dev_energy6, uint64_t tri_begin,
                                              uint64_t tri_end, int front_face, int back_face, float* out8);
/* No counterpart in the reference.  rtf_illumination_flux_mesh_device for hosts that cannot let this library allocate
 * device memory: an engine that claimed the whole GPU for a pool of its own at start-up (cudaMalloc then fails, and with
 * it the workspace the ordinary call grows on first use), or a frame loop that cannot take the device-wide
 * synchronisation cudaMalloc / cudaFree imply.  This call never allocates or frees, on the device or through any
 * library: the partial sums of the reduction tree live in a 525 312-byte array that is part of the library image, so
 * every GPU that loads librtf_b200.so holds it until the process exits.
 *   - Limit: tri_end - tri_begin <= 16 777 216 (2^24) triangles per call; more returns RTF_ERR_LIMIT with a message
 *     (split the range, or use rtf_illumination_flux_mesh_device).
 *   - dev_out8: DEVICE pointer to 8 floats, same contents and the same bits as the ordinary call's out8.  The call only
 *     enqueues on the context's stream, like every *_device entry point; nothing is copied to the host.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress, with that device current)
 *     and keeps it.  Make that first call during set-up, before any stream capture begins.
 *   - One array per GPU is shared by every call and every context on that GPU.  It keeps what the last call wrote; no
 *     part of it needs clearing, because each level of the tree reads only what the level below wrote in the same
 *     call.  Two calls whose kernels can overlap on the GPU (different streams, no ordering between them) race on it:
 *     order them. */
RTF_API int rtf_illumination_flux_mesh_device_static(rtf_ctx* ctx, const float* dev_energy6, uint64_t tri_begin,
                                                     uint64_t tri_end, int front_face, int back_face, float* dev_out8);
/* Statistics of the last illumination / render call: rays traced, paths started. */
RTF_API int rtf_last_job_stats(rtf_ctx* ctx, uint64_t* rays_traced, uint64_t* paths);

/* ---- LiDAR: CudaModule::SamplePointCloud (CM.hpp:43-45, CM.cpp:52-64)
 *      -> RayTracer::ScanPointCloud (RT.cu:758-795) -> __raygen__PCS (PCS.cu:44-75) --------- */
RTF_API int rtf_scan_point_cloud(rtf_ctx* ctx, const rtf_env* env, rtf_pc_sample* samples /* host, in-out */,
                                 size_t n, uint32_t* prim_ids_out /* host, nullable */);
/* PointCloudScanner::Scan's ray grid (PCS.cpp:70-92) generated on the device; rays
 * [ray_begin, ray_end) of the column*row grid; SoA outputs (host, each nullable):
 * t (float, <0 = miss... see hit), prim_id (0xFFFFFFFF = miss), inst_id. */
/* column = uint(size.x / distance.x), row = uint(size.y / distance.y) (PCS.cpp:71-75).  RTF_ERR_INVALID unless both
 * distances are > 0, both sizes >= 0, all finite and the quotients below 2^32 (the reference converts unchecked). */
RTF_API int rtf_scan_grid_dims(const rtf_scan_grid* grid, uint32_t* column, uint32_t* row);
RTF_API int rtf_scan_grid_ray_count(const rtf_scan_grid* grid, uint64_t* n_rays);
RTF_API int rtf_scan_grid_trace(rtf_ctx* ctx, const rtf_scan_grid* grid, uint64_t ray_begin, uint64_t ray_end,
                          float* t_out, uint32_t* prim_id_out, uint32_t* inst_id_out);
RTF_API int rtf_scan_grid_trace_device(rtf_ctx* ctx, const rtf_scan_grid* grid, uint64_t ray_begin, uint64_t ray_end,
                                 float* dev_t_out, uint32_t* dev_prim_id_out);

/* ---- raw closest-hit (the M1 metric; K8 optixTrace) ------------------------ */
/* rays: n x 8 floats {o.xyz, tmin, d.xyz, tmax}; hits: n x 4 x 32-bit {t, u, v, prim_id(u32)};
 * prim_id 0xFFFFFFFF = miss.  u,v are OptiX-style barycentrics (weights of vertex 1 and 2). */
RTF_API int rtf_trace_closest(rtf_ctx* ctx, const float* rays, size_t n, void* hits);
RTF_API int rtf_trace_closest_device(rtf_ctx* ctx, const float* dev_rays, size_t n, void* dev_hits);
/* Instrumented variant: also returns per-ray node / triangle visit counts (u32 each, device or NULL). */
RTF_API int rtf_trace_closest_counts(rtf_ctx* ctx, const float* rays, size_t n, void* hits, uint32_t* node_visits,
                                     uint32_t* tri_visits);

/* ---- camera: CameraProperties (RT.hpp:30-96, RT.cu:28-92), RayTracer::RenderToCamera
 *      (RT.hpp:368-371, RT.cu:244-603) -> __raygen__CR (CR.cu:32-147) ---------------------- */
RTF_API int rtf_camera_create(rtf_ctx* ctx, rtf_camera** out);
RTF_API int rtf_camera_destroy(rtf_camera* cam);
RTF_API int rtf_camera_get_props(rtf_camera* cam, rtf_camera_props** props); /* borrowed, mutable */
RTF_API int rtf_camera_resize(rtf_camera* cam, int width, int height);   /* CameraProperties::Resize */
/* CameraProperties::Set(position, rotation quaternion {w,x,y,z}) RT.cu:28-46 */
RTF_API int rtf_camera_set(rtf_camera* cam, const float position[3], const float rotation_wxyz[4]);
/* returns 1 rendered / 0 no acceleration structure (bool of the reference) or <0 error */
RTF_API int rtf_render_camera(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* ray_props);
/* Same call restricted to image rows [row_begin, row_end): pixels keep their global index, so their
 * RNG streams and values are those of the full-frame render; rows outside the band keep their
 * previous contents.  This is how image bands are split across GPUs (the reference is single-GPU,
 * CUDAModule.cpp:25). */
RTF_API int rtf_render_camera_rows(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* ray_props,
                                   int row_begin, int row_end);
/* which: RTF_OUT_COLOR / NORMAL / ALBEDO(also depth when output_type == DEPTH); host float4 per pixel */
RTF_API int rtf_camera_download(rtf_camera* cam, int which, float* host_rgba);
/* host -> frame buffer (width * height float4): restores a buffer, e.g. the colour buffer progressive accumulation blends into */
RTF_API int rtf_camera_upload(rtf_camera* cam, int which, const float* host);
RTF_API int rtf_camera_device_buffer(rtf_camera* cam, int which, float** dev_rgba);

/* ---- multi-GPU (no reference counterpart: CudaModule::Init is cudaSetDevice(0), CM.cpp:25; RT.cu:827) -------------
 * The BVH is replicated (every device commits the same scene); an entry point's work is split and at most one NCCL
 * collective per job runs on the context's stream.  Sharding modes of the illumination estimate: */
enum {
  RTF_SHARD_SAMPLES = 0,   /* rank r traces samples [r*S/G, (r+1)*S/G) of every triangle; ncclAllReduce(sum) of the raw
                              6*nTris sums, normalisation after it (BASELINE config 3).  <= 1e-5 of the 1-GPU result */
  RTF_SHARD_TRIANGLES = 1  /* contiguous triangle ranges; broadcasts of the finished ranges.  Bit-identical to 1 GPU */
};
#define RTF_COMM_ID_BYTES 128
/* (a) one process per GPU: rank 0 calls rtf_comm_unique_id, ships the 128 bytes to the other ranks by any means (the
 * launcher's store, MPI, a file), every rank calls rtf_comm_init_rank (collective: ncclCommInitRank). */
RTF_API int rtf_comm_unique_id(void* out, size_t bytes);
RTF_API int rtf_comm_init_rank(rtf_ctx* ctx, int nranks, int rank, const void* unique_id);
RTF_API int rtf_comm_destroy(rtf_ctx* ctx);
RTF_API int rtf_comm_rank(rtf_ctx* ctx, int* rank, int* nranks);
RTF_API int rtf_comm_barrier(rtf_ctx* ctx);
/* all-reduce (max if op_max else sum) of n <= 64 host doubles over the communicator: timings, ray counts */
RTF_API int rtf_comm_reduce_f64(rtf_ctx* ctx, double* values, int n, int op_max);
/* contiguous balanced split of n units: the first n % nranks ranks get one more */
RTF_API int rtf_shard_range(uint64_t n, int nranks, int rank, uint64_t* begin, uint64_t* end);
/* Sharded entry points of a context that belongs to a communicator (either kind); with nranks == 1 they are the plain
 * calls.  *_device: results stay in device memory (dev_out6: 6 floats per committed triangle, identical on all ranks). */
RTF_API int rtf_estimate_illumination_mesh_sharded_device(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* ray_props,
                                                          uint32_t seed, float push_normal_distance, int front_face,
                                                          int back_face, int mode, float* dev_out6);
RTF_API int rtf_estimate_illumination_mesh_sharded(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* ray_props,
                                                   uint32_t seed, float push_normal_distance, int front_face, int back_face,
                                                   int mode, float* out6 /* host */);
/* LiDAR shards are interleaved blocks of *block_rays consecutive rays (a power of two, about 64 blocks per rank): block k
 * of the *n_blocks blocks of the grid belongs to rank k mod nranks, the last block may be partial.  The rank traces its
 * blocks into the rays' own elements of the FULL-SIZE device arrays (each nullable); gather != 0 completes them on every
 * rank (grouped broadcasts).  Contiguous ranges per rank left the ranks over a field's bare rim idle. */
RTF_API int rtf_scan_grid_trace_sharded_device(rtf_ctx* ctx, const rtf_scan_grid* grid, int gather, float* dev_t,
                                               uint32_t* dev_prim, uint32_t* dev_inst, uint64_t* block_rays,
                                               uint64_t* n_blocks);
/* The layout above for any (nranks, rank), without a device: block size, number of blocks of the grid, rays of `rank`. */
RTF_API int rtf_scan_grid_shard_layout(const rtf_scan_grid* grid, int nranks, int rank, uint64_t* block_rays,
                                       uint64_t* n_blocks, uint64_t* n_local);
/* RenderToCamera over interleaved row bands; outputs_mask: bit RTF_OUT_COLOR / NORMAL / ALBEDO = complete that buffer on
 * every rank (one group of broadcasts each).  Pixels keep their global RNG streams: bit-identical to the 1-GPU frame. */
RTF_API int rtf_render_camera_sharded(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* ray_props,
                                      int outputs_mask);
/* rtf_set_option(ctx, "profile", 1): an event pair around every wavefront kernel; rtf_profile_read returns the device
 * milliseconds per category since the last read {ray generation, traversal, shading, reduction / resolve, ray sort,
 * other} and consumes the records (SURVEY §5 tracing row: the reference has no timing at all). */
RTF_API int rtf_profile_read(rtf_ctx* ctx, float* ms6);
/* {compute, collective, finalize} device milliseconds of the context's last sharded call (CUDA events) */
RTF_API int rtf_phase_timings(rtf_ctx* ctx, float* ms3);

/* (b) one process, n devices (rtf_create(devices, n)): ncclCommInitAll, one context + stream + host thread per
 * device.  Scene calls are replicated; rtf_multi_commit builds the n BVHs concurrently.  rtf_multi_ctx lends the context of
 * one device for anything not wrapped here (textures, BTFs, options): repeat such calls for every rank. */
typedef struct rtf_multi rtf_multi;
typedef struct rtf_multi_camera rtf_multi_camera;
RTF_API int rtf_create_multi(const int* devices, int n, rtf_multi** out);
RTF_API int rtf_multi_destroy(rtf_multi* m);
RTF_API int rtf_multi_size(rtf_multi* m, int* n);
RTF_API int rtf_multi_ctx(rtf_multi* m, int rank, rtf_ctx** ctx);
RTF_API int rtf_multi_set_geometry(rtf_multi* m, uint64_t handle, const void* vertices, size_t n_vertices,
                                   size_t vertex_stride_bytes, const uint32_t* triangles, size_t n_triangles, uint64_t version);
RTF_API int rtf_multi_set_material(rtf_multi* m, uint64_t handle, const rtf_material* material);
RTF_API int rtf_multi_set_instance(rtf_multi* m, uint64_t handle, uint64_t geometry, uint64_t material, const float xform[16],
                                   uint64_t component_handle);
RTF_API int rtf_multi_set_option(rtf_multi* m, const char* name, int64_t value);
RTF_API int rtf_multi_commit(rtf_multi* m, int morton_bits);
RTF_API int rtf_multi_estimate_illumination_mesh(rtf_multi* m, const rtf_env* env, const rtf_ray_props* ray_props, uint32_t seed,
                                                 float push_normal_distance, int front_face, int back_face, int mode,
                                                 float* out6 /* host, 6 floats per triangle */);
/* every device traces its interleaved blocks and copies them straight to their places in the caller's arrays (n PCIe
 * links, no collective) */
RTF_API int rtf_multi_scan_grid_trace(rtf_multi* m, const rtf_scan_grid* grid, float* t_out, uint32_t* prim_id_out,
                                      uint32_t* inst_id_out);
RTF_API int rtf_multi_trace_closest(rtf_multi* m, const float* rays, size_t n, void* hits);
RTF_API int rtf_multi_camera_create(rtf_multi* m, rtf_multi_camera** out);
RTF_API int rtf_multi_camera_destroy(rtf_multi_camera* cam);
RTF_API int rtf_multi_camera_resize(rtf_multi_camera* cam, int width, int height);
RTF_API int rtf_multi_camera_set(rtf_multi_camera* cam, const float position[3], const float rotation_wxyz[4]);
RTF_API int rtf_multi_camera_get_props(rtf_multi_camera* cam, rtf_camera_props** props);
RTF_API int rtf_multi_render_camera(rtf_multi* m, const rtf_env* env, rtf_multi_camera* cam, const rtf_ray_props* ray_props,
                                    int outputs_mask);
RTF_API int rtf_multi_camera_download(rtf_multi_camera* cam, int which, float* host);

/* ---- build-stage readback for parity tests (tests/ compares each with the oracle) --------- */
enum {
  RTF_DBG_MORTON = 0,       /* u64 per triangle reference (unsorted order)       */
  RTF_DBG_SORTED_IDS = 1,   /* u32 per triangle reference: sorted position -> reference */
  RTF_DBG_BIN_CHILDREN = 2, /* 2 x i32 per internal binary node (>=0 internal, <0 = ~leaf) */
  RTF_DBG_BIN_PARENT = 3,   /* oracle only (i32 per binary node); the CUDA builder has no parent array: 0 bytes */
  RTF_DBG_BIN_BOXES = 4,    /* 6 floats per binary node, same numbering            */
  RTF_DBG_WIDE_NODES = 5,   /* 80 B per wide node                                */
  RTF_DBG_WIDE_TRIS = 6,    /* 48 B per triangle reference, BVH (leaf-slot) order */
  RTF_DBG_FLAT_POS = 7,     /* 9 floats per triangle, commit order               */
  RTF_DBG_FLAT_NRM = 8,     /* 9 floats per triangle, commit order               */
  RTF_DBG_FLAT_INST = 9,    /* u32 per triangle                                  */
  RTF_DBG_SCENE_BOUNDS = 10, /* 6 floats                                         */
  RTF_DBG_REF_TRI = 11,     /* u32 per triangle reference: its triangle          */
  RTF_DBG_REF_BOX = 12      /* 6 floats per triangle reference: clipped bounds   */
};
RTF_API int rtf_debug_size(rtf_ctx* ctx, int what, uint64_t* bytes);
RTF_API int rtf_debug_read(rtf_ctx* ctx, int what, void* host_out, uint64_t bytes);
/* Upload a BVH built elsewhere (the oracle, in tests) so traversal can be checked in isolation. */
RTF_API int rtf_debug_upload_bvh(rtf_ctx* ctx, const void* wide_nodes, uint64_t n_nodes, const void* wide_tris,
                                 uint64_t n_tris, int depth);
/* Test hook: the BSSRDF spatial probe (SSAnyHit / BSSRDF, BSSDF.cuh:18-89) on its own.  rays: n x
 * {o.xyz, tmin, d.xyz, tmax}; handles: the private component handle each probe is restricted to;
 * out17: per probe {n_records, prim[4], t[4], u[4], v[4]} (float bits) = the up-to-four nearest
 * same-handle intersections in [tmin, tmax], ascending (t, prim).  See rtf_material.subsurface_*. */
RTF_API int rtf_sss_probe(rtf_ctx* ctx, const float* rays, const uint64_t* handles, size_t n, uint32_t* out17);
/* microbenchmarks for the roofline denominators BASELINE.md asks the builder to measure */
RTF_API int rtf_microbench_fp32(rtf_ctx* ctx, double* tflops);
RTF_API int rtf_microbench_l2(rtf_ctx* ctx, double bytes_footprint, double* gbs);

#ifdef __cplusplus
}
#endif
#endif /* RTF_RTF_H */
