// ref_optix_driver.cpp — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Runs the reference's OWN device programs of the three entry points on the host, compiled unmodified from where they
// lie.  One library per entry point, as the reference builds one PTX module per .cu file (RayFunctions.cuh has no include
// guard): -DREF_ENTRY_IE -> _ref/libie_ref.so, -DREF_ENTRY_PCS -> _ref/libpcs_ref.so, -DREF_ENTRY_CR -> _ref/libcr_ref.so.
//   /root/reference/src/RayTracerFacility/ptx/PointCloudScanning.cu          __raygen__PCS, __closesthit__PCS_R, __miss__PCS_R
//   /root/reference/src/RayTracerFacility/ptx/CameraRendering.cu             __raygen__CR, __closesthit__CR_R, __anyhit__CR_R,
//                                                                           __miss__CR_R
//   /root/reference/src/RayTracerFacility/ptx/IlluminationEstimation.cu      __raygen__IE, __closesthit__IE_R, __anyhit__IE_R,
//                                                                           __miss__IE_R (+ the _SS programs)
//   /root/reference/include/RayTracerFacility/RayTracer/RayFunctions.cuh     ClosestHitFunc, AnyHitFunc, MissFunc
//   /root/reference/include/RayTracerFacility/RayTracer/{BSDF,BSSDF,Environment,RayTracerUtilities}.cuh
//   /root/reference/include/RayTracerFacility/RayTracer/RayDataDefinations.hpp  SBT, TriangularMesh::GetHitInfo, SurfaceMaterial
//   /root/reference/include/RayTracerFacility/RayTracer/{RayTracer,LinearCongruenceGenerator}.hpp
// What the OptiX SDK and driver would provide is emulated here, and only that:
//   * the device intrinsics (optixGet*, optixIgnoreIntersection, optixTrace) as a per-thread "current ray / current hit"
//     record that optixTrace fills before it calls the reference's any-hit / closest-hit / miss program, saved and
//     restored around the recursion the reference's closest-hit program does;
//   * BVH traversal and ray/triangle intersection (closed source inside OptiX) = the oracle's orc::trace on the oracle's
//     BVH, any-hit program called per candidate in that traversal's order;
//   * tex2D<float4> = the oracle's tex2d on the scene's linear-memory textures (handle = texture index + 1);
//   * the shader binding table = one SBT record per instance, filled the way RayTracer::BuildSBT does
//     (src/RayTracerFacility/RayTracer.cu): handle, global transform, TriangularMesh{vertices, triangles},
//     SurfaceMaterial{MaterialProperties, texture objects}.
// Random numbers are the reference's own LinearCongruenceGenerator<16>, seeded as __raygen__IE does.  The oracle is run
// with the same generator in the same draw order (orc_illum_probes_lcg) for the comparison in tests/test_ref_programs.py.
// Compressed-BTF materials get a SurfaceCompressedBTF whose BTFBase is filled as in ref_btf_driver.cpp (ref_btf_fill.h).
// Not exercised: curves, and the BSSRDF spatial probe (the SS programs compile, but the repo's BSSRDF deliberately
// differs: DESIGN.md §4.4).
#include <cstddef>
#include <cstring>
#include <vector>

#include <optix_device.h>  // stand-in: declarations of the intrinsics defined below

#include "oracle_shade.h"  // the scene and the tracer

namespace emu {
struct RayCtx {
  float3 o, d;
  float tmax;
  unsigned prim;  // primitive index inside the instance's geometry
  float2 bary;
  const void* sbt;
  unsigned p0, p1;
  bool ignore;
};
static thread_local RayCtx g;
static thread_local uint3 g_launch;
static const orc::Scene* g_scene = nullptr;
static unsigned long long g_rays = 0;
// the "pipeline": the hit-group and miss programs of the entry point being launched, per ray type (Radiance,
// SpacialSampling), as RayTracer::CreateHitGroupPrograms / CreateMissPrograms bind them by name
struct Programs {
  void (*anyhit[2])();
  void (*closesthit[2])();
  void (*miss[2])();
};
static Programs g_prog;
}  // namespace emu

unsigned int optixGetPayload_0() { return emu::g.p0; }
unsigned int optixGetPayload_1() { return emu::g.p1; }
uint3 optixGetLaunchIndex() { return emu::g_launch; }
float3 optixGetWorldRayOrigin() { return emu::g.o; }
float3 optixGetWorldRayDirection() { return emu::g.d; }
float optixGetRayTmax() { return emu::g.tmax; }
unsigned int optixGetPrimitiveIndex() { return emu::g.prim; }
OptixPrimitiveType optixGetPrimitiveType() { return OPTIX_PRIMITIVE_TYPE_TRIANGLE; }
float2 optixGetTriangleBarycentrics() { return emu::g.bary; }
float optixGetCurveParameter() { return 0.0f; }
float3 optixTransformPointFromWorldToObjectSpace(float3 p) { return p; }
CUdeviceptr optixGetSbtDataPointer() { return (CUdeviceptr)emu::g.sbt; }
// On the device neither of these returns to the any-hit program: emulated by unwinding out of it.
struct IgnoreIntersection {};
struct TerminateRay {};
void optixIgnoreIntersection() {
  emu::g.ignore = true;
  throw IgnoreIntersection();
}
void optixTerminateRay() { throw TerminateRay(); }
template <>
float4 tex2D<float4>(cudaTextureObject_t tex, float u, float v) {
  float t[4];
  orc::tex2d(emu::g_scene->textures[(size_t)tex - 1], u, v, t);
  return make_float4(t[0], t[1], t[2], t[3]);
}

#include <RayTracerUtilities.cuh>
#ifndef REF_ENTRY_PCS  // PointCloudScanning.cu includes RayTracerUtilities.cuh only
#include <BSDF.cuh>
#include <BSSDF.cuh>
// Environment.cuh:77 says `const const EnvironmentProperties &` (nvcc accepts it, g++ does not): that one file is
// included with `const` defined away (see ref_env_driver.cpp), and the call in MissFunc, which passes a const
// environment, goes through the forwarding overload below.
#define const
#include <Environment.cuh>
#undef const
namespace RayTracerFacility {
static inline glm::vec3 CalculateEnvironmentalLight(const glm::vec3& position, const glm::vec3& rayDir,
                                                    const EnvironmentProperties& environment) {
  glm::vec3 p = position, d = rayDir;
  EnvironmentProperties e = environment;
  return CalculateEnvironmentalLight(p, d, e);
}
}  // namespace RayTracerFacility
#endif
namespace RayTracerFacility {
// `extern "C" __constant__ X x;` is a definition for nvcc and a declaration for g++: define the launch parameters here.
extern "C" {
IlluminationEstimationLaunchParams illuminationEstimationLaunchParams;
PointCloudScanningLaunchParams pointCloudScanningLaunchParams;
CameraRenderingLaunchParams cameraRenderingLaunchParams;
}
}  // namespace RayTracerFacility
#if defined(REF_ENTRY_IE)
#include <IlluminationEstimation.cu>
#elif defined(REF_ENTRY_PCS)
#include <PointCloudScanning.cu>
#elif defined(REF_ENTRY_CR)
#include <CameraRendering.cu>
#else
#error "define REF_ENTRY_IE, REF_ENTRY_PCS or REF_ENTRY_CR"
#endif

#include "ref_btf_fill.h"

using namespace RayTracerFacility;
#define REF_API extern "C" __attribute__((visibility("default")))

// The C ABI's records are documented as byte-compatible with the reference's: check them against the real ones.
static_assert(sizeof(UniEngine::Vertex) == sizeof(rtf_vertex) && sizeof(rtf_vertex) == 80, "vertex");
static_assert(sizeof(IlluminationSampler<glm::vec3>) == sizeof(rtf_illum_sampler), "light probe size");
static_assert(offsetof(IlluminationSampler<glm::vec3>, m_direction) == offsetof(rtf_illum_sampler, direction), "direction");
static_assert(offsetof(IlluminationSampler<glm::vec3>, m_energy) == offsetof(rtf_illum_sampler, energy), "energy");
static_assert(offsetof(IlluminationSampler<glm::vec3>, m_frontFace) == offsetof(rtf_illum_sampler, front_face), "front");
static_assert(offsetof(IlluminationSampler<glm::vec3>, m_backFace) == offsetof(rtf_illum_sampler, back_face), "back");
static_assert(sizeof(HitInfo) == sizeof(rtf_hit_info), "hit info");
static_assert(sizeof(PointCloudSample) == sizeof(rtf_pc_sample), "point cloud sample");
static_assert(offsetof(PointCloudSample, m_handle) == offsetof(rtf_pc_sample, handle), "sample handle");
static_assert(offsetof(PointCloudSample, m_hitInfo) == offsetof(rtf_pc_sample, hit_info), "sample hit info");
static_assert(sizeof(RayProperties) == sizeof(rtf_ray_props), "ray properties");

namespace emu {
static std::vector<SBT> g_sbt;  // one record per instance
static std::vector<TriangularMesh> g_meshes;
static std::vector<SurfaceMaterial> g_materials;
static std::vector<SurfaceCompressedBTF> g_btf_materials;  // parallel to g_materials; used where the type is CompressedBTF

static void set_hit(const float3& o, const float3& d, unsigned p0, unsigned p1, uint32_t inst, uint32_t prim, float t,
                    float u, float v) {
  g.o = o, g.d = d, g.tmax = t, g.p0 = p0, g.p1 = p1, g.ignore = false;
  g.prim = prim - g_scene->insts[inst].first_prim;
  g.bary = make_float2(u, v);
  g.sbt = &g_sbt[inst];
}
}  // namespace emu

void optixTrace(OptixTraversableHandle, float3 origin, float3 direction, float tmin, float tmax, float, OptixVisibilityMask,
                unsigned int flags, unsigned int sbt_offset, unsigned int, unsigned int miss_index, unsigned int& p0,
                unsigned int& p1) {
  using namespace emu;
  const orc::Scene& sc = *g_scene;
  const RayCtx saved = g;
  const bool radiance = sbt_offset == (unsigned)RayType::Radiance;
  const orc::Ray ray{orc::v3(origin.x, origin.y, origin.z), tmin, orc::v3(direction.x, direction.y, direction.z), tmax};
  if (radiance) g_rays++;  // the spatial-sampling probes of the BSSRDF are not counted as rays (neither does the oracle)
  orc::Hit h;
  if (flags & OPTIX_RAY_FLAG_DISABLE_ANYHIT) {
    h = orc::trace(sc.bvh, ray, nullptr, (flags & OPTIX_RAY_FLAG_TERMINATE_ON_FIRST_HIT) != 0);
  } else {
    const unsigned q0 = p0, q1 = p1;
    bool terminated = false;  // optixTerminateRay: the current hit is kept, nothing further is reported
    h = orc::trace(sc.bvh, ray, nullptr, false, [&](const orc::Tri48& tri, const orc::Hit& cand) {
      if (terminated) return false;
      set_hit(origin, direction, q0, q1, tri.inst, tri.prim, cand.t, cand.u, cand.v);
      try {
        g_prog.anyhit[radiance ? 0 : 1]();
      } catch (const IgnoreIntersection&) {
      } catch (const TerminateRay&) {
        terminated = true;
      }
      return !g.ignore;
    });
  }
  if (h.prim != 0xFFFFFFFFu) {
    if (!(flags & OPTIX_RAY_FLAG_DISABLE_CLOSESTHIT)) {
      set_hit(origin, direction, p0, p1, sc.tri_inst[h.prim], h.prim, h.t, h.u, h.v);
      g_prog.closesthit[radiance ? 0 : 1]();
    }
  } else {
    g.o = origin, g.d = direction, g.tmax = tmax, g.p0 = p0, g.p1 = p1;
    g_prog.miss[miss_index == (unsigned)RayType::Radiance ? 0 : 1]();
  }
  g = saved;
}

// The scene side of the launch: texture objects, materials, geometry and one SBT record per instance.
static int setup_scene(const orc::Scene* sc) {
  using namespace emu;
  g_scene = sc;
  g_rays = 0;
  g_meshes.assign(sc->geoms.size(), TriangularMesh());
  for (size_t k = 0; k < sc->geoms.size(); k++) {
    g_meshes[k].m_vertices = (UniEngine::Vertex*)sc->geoms[k].verts.data();
    g_meshes[k].m_triangles = (glm::uvec3*)sc->geoms[k].tris.data();
  }
  g_materials.assign(sc->mats.size(), SurfaceMaterial());
  g_btf_materials.assign(sc->mats.size(), SurfaceCompressedBTF());
  for (size_t k = 0; k < sc->mats.size(); k++) {
    const rtf_material& m = sc->mats[k];
    if (m.material_type == RTF_MAT_COMPRESSED_BTF) {
      const int bi = orc::mat_extra(*sc, (uint32_t)k).btf;
      if (bi < 0) return -1;
      const orc::BTF& btf = sc->btfs[(size_t)bi];
      rtf_btf_desc d = btf.d;  // scalars; the tables are the scene's own copies (btf.d still points at the caller's)
      d.pdf6_slices = btf.pdf6_slices.data(), d.pdf6_scales = btf.pdf6_scales.data(), d.pdf1d_basis = btf.pdf1d_basis.data();
      d.colors = btf.colors.data(), d.index_ab = btf.index_ab.data(), d.pdf2d_colors = btf.pdf2d_colors.data();
      d.pdf2d_slices = btf.pdf2d_slices.data(), d.pdf2d_scales = btf.pdf2d_scales.data();
      d.pdf2d_indices = btf.pdf2d_indices.data(), d.pdf3d_scales = btf.pdf3d_scales.data();
      d.pdf3d_slices = btf.pdf3d_slices.data(), d.pdf4d_scales = btf.pdf4d_scales.data();
      d.pdf4d_slices = btf.pdf4d_slices.data();
      ref_fill_btf(g_btf_materials[k].m_btf, &d, btf.beta_angles.data());
      continue;
    }
    UniEngine::MaterialProperties& mp = g_materials[k].m_materialProperties;
    mp.m_albedoColor = glm::vec3(m.albedo_color[0], m.albedo_color[1], m.albedo_color[2]);
    mp.m_subsurfaceColor = glm::vec3(m.subsurface_color[0], m.subsurface_color[1], m.subsurface_color[2]);
    mp.m_subsurfaceFactor = m.subsurface_factor;
    mp.m_subsurfaceRadius = glm::vec3(m.subsurface_radius[0], m.subsurface_radius[1], m.subsurface_radius[2]);
    mp.m_metallic = m.metallic, mp.m_roughness = m.roughness, mp.m_emission = m.emission;
    mp.m_transmission = m.transmission;
    const orc::MatExtra& mx = orc::mat_extra(*sc, (uint32_t)k);
    g_materials[k].m_albedoTexture = (cudaTextureObject_t)(mx.tex[0] + 1);
    g_materials[k].m_normalTexture = (cudaTextureObject_t)(mx.tex[1] + 1);
    g_materials[k].m_metallicTexture = (cudaTextureObject_t)(mx.tex[2] + 1);
    g_materials[k].m_roughnessTexture = (cudaTextureObject_t)(mx.tex[3] + 1);
  }
  g_sbt.assign(sc->insts.size(), SBT());
  for (size_t k = 0; k < sc->insts.size(); k++) {
    const orc::Instance& in = sc->insts[k];
    SBT& s = g_sbt[k];
    s.m_handle = in.handle;
    std::memcpy((void*)&s.m_globalTransform, in.xform, 64);  // column-major, like glm::mat4
    s.m_geometryType = RendererType::Default;
    s.m_geometry = &g_meshes[in.geom];
    s.m_materialType = (MaterialType)sc->mats[in.mat].material_type;
    s.m_material = s.m_materialType == MaterialType::CompressedBTF ? (void*)&g_btf_materials[in.mat] : (void*)&g_materials[in.mat];
  }
  return 0;
}
static void bind_environment(EnvironmentProperties& e, const rtf_env* env, const orc::Scene* sc) {
  std::memcpy((void*)&e, env, sizeof(EnvironmentProperties));
  for (int k = 0; k < 6; k++)  // the cube map bound to the scene (rtf_set_environment_map), as texture objects
    e.m_environmentalMaps[k] = (cudaTextureObject_t)(sc->env_faces[k] + 1);
  if (sc->env_faces[0] < 0) e.m_environmentalMapId = 0;
}

#ifdef REF_ENTRY_IE
// __raygen__IE over `n` light probes (in-out, the reference's 268 B records) on a committed oracle scene.  Returns 0, or
// -1 when a BTF material has no table set.
REF_API int ref_ie_run(const orc::Scene* sc, const rtf_env* env, const rtf_ray_props* rp, rtf_illum_sampler* probes,
                       unsigned long long n, unsigned seed, float push, unsigned long long* rays) {
  using namespace emu;
  if (setup_scene(sc) != 0) return -1;
  g_prog = Programs{{__anyhit__IE_R, __anyhit__IE_SS}, {__closesthit__IE_R, __closesthit__IE_SS}, {__miss__IE_R, __miss__IE_SS}};
  IlluminationEstimationLaunchParams& lp = illuminationEstimationLaunchParams;
  lp.m_seed = seed, lp.m_pushNormalDistance = push, lp.m_size = n;
  bind_environment(lp.m_rayTracerProperties.m_environment, env, sc);
  lp.m_rayTracerProperties.m_rayProperties.m_bounces = rp->bounces;
  lp.m_rayTracerProperties.m_rayProperties.m_samples = rp->samples;
  lp.m_lightProbes = (IlluminationSampler<glm::vec3>*)probes;
  lp.m_traversable = 1;
  for (unsigned long long ix = 0; ix < n; ix++) {
    g_launch = make_uint3((unsigned)ix, 0, 0);
    __raygen__IE();
  }
  if (rays) *rays = g_rays;
  return 0;
}
#endif
#ifdef REF_ENTRY_PCS
// __raygen__PCS over `n` point-cloud samples (in-out, the reference's 112 B records).
REF_API int ref_pcs_run(const orc::Scene* sc, rtf_pc_sample* samples, unsigned long long n) {
  using namespace emu;
  if (setup_scene(sc) != 0) return -1;
  g_prog = Programs{{__anyhit__PCS_R, __anyhit__PCS_SS}, {__closesthit__PCS_R, __closesthit__PCS_SS}, {__miss__PCS_R, __miss__PCS_SS}};
  PointCloudScanningLaunchParams& lp = pointCloudScanningLaunchParams;
  lp.m_size = n, lp.m_samples = (PointCloudSample*)samples, lp.m_traversable = 1;
  for (unsigned long long ix = 0; ix < n; ix++) {
    g_launch = make_uint3((unsigned)ix, 0, 0);
    __raygen__PCS();
  }
  return 0;
}
#endif
#ifdef REF_ENTRY_CR
// __raygen__CR over the whole frame.  color / normal / albedo: width * height float4, color in-out (the accumulation
// reads the previous frame from it, CR.cu:112-127).
REF_API int ref_cr_run(const orc::Scene* sc, const rtf_env* env, const rtf_ray_props* rp, const rtf_camera_props* cam,
                       float* color, float* normal, float* albedo, unsigned long long* rays) {
  using namespace emu;
  if (setup_scene(sc) != 0) return -1;
  g_rays = 0;
  g_prog = Programs{{__anyhit__CR_R, __anyhit__CR_SS}, {__closesthit__CR_R, __closesthit__CR_SS}, {__miss__CR_R, __miss__CR_SS}};
  CameraRenderingLaunchParams& lp = cameraRenderingLaunchParams;
  CameraProperties& c = lp.m_cameraProperties;
  c.m_accumulate = cam->accumulate != 0, c.m_fov = cam->fov;
  c.m_from = glm::vec3(cam->from[0], cam->from[1], cam->from[2]);
  c.m_horizontal = glm::vec3(cam->horizontal[0], cam->horizontal[1], cam->horizontal[2]);
  c.m_vertical = glm::vec3(cam->vertical[0], cam->vertical[1], cam->vertical[2]);
  std::memcpy((void*)&c.m_inverseProjectionView, cam->inverse_projection_view, 64);
  c.m_maxDistance = cam->max_distance, c.m_outputType = (OutputType)cam->output_type, c.m_gamma = cam->gamma;
  c.m_frame.m_colorBuffer = (glm::vec4*)color, c.m_frame.m_normalBuffer = (glm::vec4*)normal;
  c.m_frame.m_albedoBuffer = (glm::vec4*)albedo;
  c.m_frame.m_size = glm::ivec2(cam->width, cam->height), c.m_frame.m_frameId = cam->frame_id;
  c.m_aperture = cam->aperture, c.m_focalLength = cam->focal_length;
  bind_environment(lp.m_rayTracerProperties.m_environment, env, sc);
  lp.m_rayTracerProperties.m_rayProperties.m_bounces = rp->bounces;
  lp.m_rayTracerProperties.m_rayProperties.m_samples = rp->samples;
  lp.m_traversable = 1;
  for (int iy = 0; iy < cam->height; iy++)
    for (int ix = 0; ix < cam->width; ix++) {
      g_launch = make_uint3((unsigned)ix, (unsigned)iy, 0);
      __raygen__CR();
    }
  if (rays) *rays = g_rays;
  return 0;
}
#endif
