// ref_env_driver.cpp — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Exposes the reference's OWN sampling helpers, sky model and environment lookup, compiled unmodified on the host from
// where they lie:
//   /root/reference/include/RayTracerFacility/RayTracer/RayTracerUtilities.cuh:54-157  SampleCubeMap, Reflect, Refract,
//        GetTangentSpace, RandomSampleHemisphere (both), RandomSampleSphere, RandomSampleDisk
//   /root/reference/include/RayTracerFacility/RayTracer/Environment.cuh:8-175          CIESkyIntensity, RaySphereIntersect,
//        NishitaSkyIncidentLight, CalculateEnvironmentalLight
//   /root/reference/include/RayTracerFacility/RayTracer/BSDF.cuh:6-13                  BRDF
//   /root/reference/include/RayTracerFacility/RayTracer/RayTracer.hpp:98-147           EnvironmentProperties
// against the stand-in headers of oracle/ref_stubs/ (glm subset, opaque OptiX handles, UniEngine vertex / material
// structs, an empty RayDataDefinations.hpp); cuda_runtime.h comes from the toolkit.  The output goes to oracle/_ref/ and
// is used to (1) generate tests/golden/env_ref.npz and (2) validate the restatements in oracle_shade.h / oracle_btf.h.
// The random draws come from the reference's own LinearCongruenceGenerator<16>; each sampler export also returns the
// uniforms it consumed so that the restatement can be fed the same numbers.
#include <cstddef>
#include <cstring>

#include <optix_device.h>  // stand-in: declares tex2D<T>

// SampleCubeMap's texture fetch: record which face handle was addressed and where; return a colour that encodes them.
static thread_local unsigned long long g_face;
static thread_local float g_u, g_v;
template <>
float4 tex2D<float4>(cudaTextureObject_t tex, float u, float v) {
  g_face = tex, g_u = u, g_v = v;
  return make_float4((float)tex, u, v, 1.0f);
}

#include <RayTracerUtilities.cuh>
#include <BSDF.cuh>
// Environment.cuh:77 spells a parameter `const const EnvironmentProperties &`, which nvcc's front end accepts with a
// warning and g++ rejects.  The file is compiled as it lies, with `const` defined away for its extent only (everything it
// includes has been included above, so no other header sees the macro); dropping const qualifiers from valid code does
// not change any arithmetic.  The exports below therefore hand it lvalues.
#define const
#include <Environment.cuh>
#undef const

#include "../include/rtf/rtf.h"

using namespace RayTracerFacility;
#define REF_API extern "C" __attribute__((visibility("default")))

// rtf_env is documented as byte-compatible with EnvironmentProperties (include/rtf/rtf.h): check it against the real one.
static_assert(sizeof(EnvironmentProperties) == sizeof(rtf_env), "rtf_env size");
static_assert(offsetof(EnvironmentProperties, m_sunDirection) == offsetof(rtf_env, sun_direction), "sun_direction");
static_assert(offsetof(EnvironmentProperties, m_color) == offsetof(rtf_env, color), "color");
static_assert(offsetof(EnvironmentProperties, m_environmentalMapId) == offsetof(rtf_env, environmental_map_id), "map id");
static_assert(offsetof(EnvironmentProperties, m_environmentalMaps) == offsetof(rtf_env, environmental_maps), "maps");
static_assert(offsetof(EnvironmentProperties, m_atmosphere) == offsetof(rtf_env, atmosphere), "atmosphere");

static inline glm::vec3 V(const float* p) { return glm::vec3(p[0], p[1], p[2]); }
static inline void S(float* o, const glm::vec3& v) { o[0] = v.x, o[1] = v.y, o[2] = v.z; }
static inline Random rng(unsigned v0, unsigned v1, float* r2) {
  Random r;
  r.Init(v0, v1);
  if (r2) {
    Random c = r;
    r2[0] = c(), r2[1] = c();
  }
  return r;
}

REF_API void ref_reflect(const float* i, const float* n, float* out) { S(out, Reflect(V(i), V(n))); }
REF_API void ref_refract(const float* i, const float* n, float ior, float* out) { S(out, Refract(V(i), V(n), ior)); }
REF_API void ref_tangent_space(const float* n, float* out9) {
  const glm::mat3x3 m = GetTangentSpace(V(n));
  S(out9, m[0]), S(out9 + 3, m[1]), S(out9 + 6, m[2]);
}
REF_API void ref_sample_hemisphere(unsigned v0, unsigned v1, const float* n, float* r2, float* out) {
  Random r = rng(v0, v1, r2);
  S(out, RandomSampleHemisphere(r, V(n)));
}
REF_API void ref_sample_lobe(unsigned v0, unsigned v1, const float* n, float alpha, float* r2, float* out) {
  Random r = rng(v0, v1, r2);
  S(out, RandomSampleHemisphere(r, V(n), alpha));
}
REF_API void ref_sample_sphere(unsigned v0, unsigned v1, float* r2, float* out) {
  Random r = rng(v0, v1, r2);
  S(out, RandomSampleSphere(r));
}
REF_API void ref_sample_disk(unsigned v0, unsigned v1, float* r2, float* out2) {
  Random r = rng(v0, v1, r2);
  const glm::vec2 d = RandomSampleDisk(r);
  out2[0] = d.x, out2[1] = d.y;
}
REF_API void ref_brdf(float metallic, unsigned v0, unsigned v1, const float* in_dir, const float* normal, float* r2,
                      float* out) {
  Random r = rng(v0, v1, r2);
  float3 o;
  BRDF(metallic, r, V(in_dir), V(normal), o);
  out[0] = o.x, out[1] = o.y, out[2] = o.z;
}
REF_API void ref_cube_map_lookup(const float* dir, int* face, float* uv) {
  cudaTextureObject_t faces[6] = {0, 1, 2, 3, 4, 5};
  SampleCubeMap<float4>(faces, V(dir));
  *face = (int)g_face, uv[0] = g_u, uv[1] = g_v;
}
REF_API int ref_ray_sphere(const float* o, const float* d, float radius, float* t01) {
  glm::vec3 vo = V(o), vd = V(d);
  return RaySphereIntersect(vo, vd, radius, t01[0], t01[1]) ? 1 : 0;
}
REF_API float ref_cie_sky(const float* dir, const float* sun, const float* zenith) {
  glm::vec3 vs = V(sun), vz = V(zenith);
  return CIESkyIntensity(V(dir), vs, vz);
}
static inline EnvironmentProperties E(const void* env_bytes) {
  EnvironmentProperties e;
  std::memcpy((void*)&e, env_bytes, sizeof(e));
  return e;
}
REF_API void ref_nishita_sky(const void* env_bytes, const float* pos, const float* dir, float* out) {
  glm::vec3 p = V(pos), d = V(dir);
  EnvironmentProperties e = E(env_bytes);
  S(out, NishitaSkyIncidentLight(p, d, e));
}
// With a cube map bound (environmental_map_id != 0) the colour returned is (face handle, u, v) through the lighting
// type's intensity / gamma arithmetic: the handles in env_bytes' environmental_maps[] are returned as colours.
REF_API void ref_env_light(const void* env_bytes, const float* pos, const float* dir, float* out) {
  glm::vec3 p = V(pos), d = V(dir);
  EnvironmentProperties e = E(env_bytes);
  S(out, CalculateEnvironmentalLight(p, d, e));
}
