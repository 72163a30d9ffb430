// rtf_shade.cuh — device shading math: environment light, hit reconstruction.
// Restates (as wavefront-friendly device functions) /root/reference/include/RayTracerFacility/RayTracer/
// Environment.cuh:28-175 (miss program / sky), RayDataDefinations.hpp:127-164 + :364-382 (hit info).
#pragma once
#include "rtf_internal.h"

namespace rtf {

// ---- Environment.cuh:28-65 -------------------------------------------------------------------------
RTF_HD bool solve_quadratic(float a, float b, float c, float& x1, float& x2) {
  if (b == 0) {
    if (a == 0) return false;
    x1 = 0;
    x2 = sqrtf(-c / a);
    return true;
  }
  float discr = b * b - 4 * a * c;
  if (discr < 0) return false;
  float q = (b < 0.f) ? -0.5f * (b - sqrtf(discr)) : -0.5f * (b + sqrtf(discr));
  x1 = q / a;
  x2 = c / q;
  return true;
}
RTF_HD bool ray_sphere(float3 o, float3 d, float radius, float& t0, float& t1) {
  float A = (d.x * d.x + d.y * d.y) + d.z * d.z;
  float B = 2.0f * ((d.x * o.x + d.y * o.y) + d.z * o.z);
  float C = ((o.x * o.x + o.y * o.y) + o.z * o.z) - radius * radius;
  if (!solve_quadratic(A, B, C, t0, t1)) return false;
  if (t0 > t1) {
    float t = t0;
    t0 = t1, t1 = t;
  }
  return true;
}
// NishitaSkyIncidentLight, Environment.cuh:75-145 (exp/pow: tolerance class, not bit-exact vs libm)
RTF_HD float3 nishita_sky(float3 position, float3 dir, const rtf_env& env) {
  float earth = env.atmosphere.earth_radius * 1000.0f;
  float atmos = env.atmosphere.atmosphere_radius * 1000.0f;
  float Hr = env.atmosphere.hr, Hm = env.atmosphere.hm;
  float3 betaR = f3(3.8e-6f, 13.5e-6f, 33.1e-6f);
  float3 betaM = f3(21e-6f, 21e-6f, 21e-6f);
  float3 sun = f3(env.sun_direction[0], env.sun_direction[1], env.sun_direction[2]);
  float tmin = 0, tmax = 999999999999.0f;
  float3 orig = position + f3(0.0f, earth, 0.0f);
  float t0, t1;
  if (!ray_sphere(orig, dir, atmos, t0, t1) || t1 < 0.0f) return f3(0, 0, 0);
  if (t0 > tmin && t0 > 0.0f) tmin = t0;
  if (t1 < tmax) tmax = t1;
  unsigned ns = (unsigned)env.atmosphere.num_samples, nl = (unsigned)env.atmosphere.num_samples_light;
  float seg = (tmax - tmin) / ns;
  float tcur = tmin;
  float3 sumR = f3(0, 0, 0), sumM = f3(0, 0, 0);
  float odR = 0, odM = 0;
  float mu = dot3(dir, sun);
  float phaseR = 3.f / (16.f * 3.1415926f) * (1.0f + mu * mu);
  float g = env.atmosphere.g;
  float phaseM = 3.f / (8.f * 3.1415926f) * ((1.f - g * g) * (1.f + mu * mu)) /
                 ((2.f + g * g) * powf(1.f + g * g - 2.f * g * mu, 1.5f));
  for (unsigned i = 0; i < ns; ++i) {
    float3 sp = orig + (tcur + seg * 0.5f) * dir;
    float height = length3(sp) - earth;
    float hr = expf(-height / Hr) * seg;
    float hm = expf(-height / Hm) * seg;
    odR += hr;
    odM += hm;
    float t0l = 0, t1l = 0;
    ray_sphere(sp, sun, atmos, t0l, t1l);
    float segl = t1l / nl, tcl = 0;
    float olR = 0, olM = 0;
    unsigned j;
    for (j = 0; j < nl; ++j) {
      float3 spl = sp + (tcl + segl * 0.5f) * sun;
      float hl = length3(spl) - earth;
      if (hl < 0) break;
      olR += expf(-hl / Hr) * segl;
      olM += expf(-hl / Hm) * segl;
      tcl += segl;
    }
    if (j == nl) {
      float3 tau = betaR * (odR + olR) + (betaM * 1.1f) * (odM + olM);
      float3 att = f3(expf(-tau.x), expf(-tau.y), expf(-tau.z));
      sumR = sumR + att * hr;
      sumM = sumM + att * hm;
    }
    tcur += seg;
  }
  return ((sumR * betaR) * phaseR + (sumM * betaM) * phaseM) * 20.0f;
}
// CalculateEnvironmentalLight, Environment.cuh:147-175.  env_faces: the six cube-map faces (rtf_set_environment_map) or
// nullptr; sampled when m_environmentalMapId != 0 (Env.cuh:153-158).
RTF_D float3 env_light(float3 position, float3 dir, const rtf_env& env, const DevTexture* env_faces) {
  float3 c = f3(1, 1, 1);
  float3 col = f3(env.color[0], env.color[1], env.color[2]);
  switch (env.lighting_type) {
    case RTF_ENV_SCENE:
      if (env.environmental_map_id != 0 && env_faces != nullptr) {
        const float4 t = sample_cube_map(env_faces, dir);
        col = f3(t.x, t.y, t.z);
      }
      c = col * env.skylight_intensity;
      break;
    case RTF_ENV_SKYDOME: c = nishita_sky(position, dir, env) * env.skylight_intensity; break;
    case RTF_ENV_SINGLE_LIGHT_SOURCE: c = col * env.skylight_intensity; break;
  }
  if (env.gamma != 1.0f) {
    float ig = 1.0f / env.gamma;
    c = f3(powf(c.x, ig), powf(c.y, ig), powf(c.z, ig));
  }
  return max3(f3(0, 0, 0), c);
}

// ---- hit reconstruction: TriangularMesh::GetHitInfo (RayDataDefinations.hpp:127-149) interpolates the OBJECT-space
//      vertices; SBT::GetHitInfo (RayDataDefinations.hpp:375-380) then transforms: normal = normalize(M * (n, 0)), faced
//      against the ray, position = M * (p, 1).  Same association as the reference, so the result is bit-equal to it under
//      any instance transform (interpolating pre-transformed vertices would differ by an ulp). -------
struct SurfaceHit {
  float3 position, normal;
  uint32_t inst;
};
RTF_D SurfaceHit surface_hit(const DevScene& sc, uint32_t prim, float u, float v, float3 ray_dir) {
  SurfaceHit s;
  s.inst = __ldg(sc.tri_inst + prim);
  const InstDev& in = sc.insts[s.inst];
  const uint32_t lt = prim - in.first_prim;
  const uint32_t i0 = __ldg(in.tris + 3 * lt), i1 = __ldg(in.tris + 3 * lt + 1), i2 = __ldg(in.tris + 3 * lt + 2);
  // position|padding and normal|padding are the first two 16 B quads of the 80 B vertex
  const float4* q0 = reinterpret_cast<const float4*>(in.verts + i0);
  const float4* q1 = reinterpret_cast<const float4*>(in.verts + i1);
  const float4* q2 = reinterpret_cast<const float4*>(in.verts + i2);
  const float4 a0 = __ldg(q0), b0 = __ldg(q0 + 1), a1 = __ldg(q1), b1 = __ldg(q1 + 1), a2 = __ldg(q2), b2 = __ldg(q2 + 1);
  const float w0 = (1.0f - u) - v;
  const float3 pl = (w0 * f3(a0.x, a0.y, a0.z) + u * f3(a1.x, a1.y, a1.z)) + v * f3(a2.x, a2.y, a2.z);
  const float3 nl = (w0 * f3(b0.x, b0.y, b0.z) + u * f3(b1.x, b1.y, b1.z)) + v * f3(b2.x, b2.y, b2.z);
  float m[16];
#pragma unroll
  for (int k = 0; k < 16; k++) m[k] = in.xform[k];
  float3 nn = normalize3(mat4_mul(m, nl, 0.0f));
  if (dot3(ray_dir, nn) > 0.0f) nn = neg3(nn);
  s.normal = nn;
  s.position = mat4_mul(m, pl, 1.0f);
  return s;
}
// Interpolated vertex attributes for the materials that read them (TriangularMesh::GetHitInfo, RDD.hpp:127-164;
// tangent through SBT::GetHitInfo, RDD.hpp:379).
struct HitAttrs {
  float u, v;      // texture coordinate
  float3 tangent;  // normalize(M * t)
  float3 color;    // colour of the dominant-barycentric vertex (RDD.hpp:150-160, as written)
};
RTF_D HitAttrs hit_attrs(const DevScene& sc, uint32_t prim, float bu, float bv, uint32_t inst) {
  const InstDev& in = sc.insts[inst];
  const uint32_t lt = prim - in.first_prim;
  const rtf_vertex& vx = in.verts[in.tris[3 * lt]];
  const rtf_vertex& vy = in.verts[in.tris[3 * lt + 1]];
  const rtf_vertex& vz = in.verts[in.tris[3 * lt + 2]];
  const float z = (1.0f - bu) - bv;
  HitAttrs a;
  float3 tg = (z * f3(vx.tangent[0], vx.tangent[1], vx.tangent[2]) + bu * f3(vy.tangent[0], vy.tangent[1], vy.tangent[2])) +
              bv * f3(vz.tangent[0], vz.tangent[1], vz.tangent[2]);
  float m[16];
#pragma unroll
  for (int k = 0; k < 16; k++) m[k] = in.xform[k];
  a.tangent = normalize3(mat4_mul(m, tg, 0.0f));
  a.u = (z * vx.tex_coord[0] + bu * vy.tex_coord[0]) + bv * vz.tex_coord[0];
  a.v = (z * vx.tex_coord[1] + bu * vy.tex_coord[1]) + bv * vz.tex_coord[1];
  const rtf_vertex* pick = (bu > z && bu > bv) ? &vx : (bv > z ? &vy : &vz);
  a.color = f3(pick->color[0], pick->color[1], pick->color[2]);
  return a;
}
// Full HitInfo as __closesthit__PCS_R writes it (PointCloudScanning.cu:15-28).
RTF_D void full_hit_info(const DevScene& sc, uint32_t prim, float u, float v, float3 ray_dir, rtf_hit_info& out,
                         uint64_t& handle) {
  const SurfaceHit s = surface_hit(sc, prim, u, v, ray_dir);
  const InstDev& in = sc.insts[s.inst];
  handle = in.handle;
  const uint32_t lt = prim - in.first_prim;
  const rtf_vertex& vx = in.verts[in.tris[3 * lt]];
  const rtf_vertex& vy = in.verts[in.tris[3 * lt + 1]];
  const rtf_vertex& vz = in.verts[in.tris[3 * lt + 2]];
  const float z = (1.0f - u) - v;
  float3 tg = (z * f3(vx.tangent[0], vx.tangent[1], vx.tangent[2]) + u * f3(vy.tangent[0], vy.tangent[1], vy.tangent[2])) +
              v * f3(vz.tangent[0], vz.tangent[1], vz.tangent[2]);
  float m[16];
#pragma unroll
  for (int k = 0; k < 16; k++) m[k] = in.xform[k];
  tg = normalize3(mat4_mul(m, tg, 0.0f));
  out.position[0] = s.position.x, out.position[1] = s.position.y, out.position[2] = s.position.z;
  out.normal[0] = s.normal.x, out.normal[1] = s.normal.y, out.normal[2] = s.normal.z;
  out.tangent[0] = tg.x, out.tangent[1] = tg.y, out.tangent[2] = tg.z;
  out.tex_coord[0] = (z * vx.tex_coord[0] + u * vy.tex_coord[0]) + v * vz.tex_coord[0];
  out.tex_coord[1] = (z * vx.tex_coord[1] + u * vy.tex_coord[1]) + v * vz.tex_coord[1];
  const rtf_vertex* pick = (u > z && u > v) ? &vx : (v > z ? &vy : &vz);  // RayDataDefinations.hpp:150-160 as written
#pragma unroll
  for (int k = 0; k < 4; k++) out.color[k] = pick->color[k];
  out.data[0] = pick->position_padding, out.data[1] = pick->normal_padding;
  out.data[2] = pick->tangent_padding, out.data[3] = pick->tex_coord_padding[0];
}

}  // namespace rtf
