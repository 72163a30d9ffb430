// oracle_shade.h — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Scalar CPU restatement of the reference's device programs for the hot path.  Paths are relative
// to /root/reference; RTU.cuh/RF.cuh/Env.cuh/BSDF.cuh/RDD.hpp live in
// include/RayTracerFacility/RayTracer/, IE.cu/PCS.cu/CR.cu in src/RayTracerFacility/ptx/.
//
// Deliberate, documented departures (SURVEY §7 "Hard parts", §8 "Quirks"):
//  * RNG: the per-thread LCG stream (LCG.hpp) is replaced by Philox keyed (seed, unit, bounce), so a
//    path's draws do not depend on other paths or on traversal order (the reference's own sequence is
//    not reproducible: AnyHitFunc draws per candidate hit in OptiX traversal order, RF.cuh:19).
//  * Recursion (RF.cuh:136-161 calls optixTrace inside closest-hit) is unrolled into the equivalent
//    throughput/accumulator loop; products are associated outside-in instead of inside-out.
//  * sin/cos(2*pi*u) use the portable polynomial in oracle_math.h.
//  * pow(c, 1/gamma) is skipped when gamma == 1 (identity).
// Everything else — barycentric sample not folded (IE.cu:56-59), un-normalised probe normal
// (IE.cu:60-62,81), lobe around the mirror direction (BSDF.cuh:6-13), bounce origin not offset
// (RF.cuh:138), normal transformed by M and flipped against the ray (RDD.hpp:375-378), dominant-vertex
// colour pick (RDD.hpp:150-160) — follows the reference.
#pragma once
#include <cstdint>
#include <vector>

#include "../include/rtf/rtf.h"
#include "oracle_btf.h"
#include "oracle_bvh.h"
#include "oracle_sss.h"

namespace orc {

struct Geometry {
  std::vector<rtf_vertex> verts;
  std::vector<uint32_t> tris;  // 3 per triangle
};
struct Instance {
  uint32_t geom, mat;
  float xform[16];
  uint64_t handle;
  uint32_t first_prim;
};

// RayTracedMaterial's texture ids (RT.hpp:254-263) and its BTF (RT.cu:1888-1894), as indices into the scene tables.
struct MatExtra {
  int tex[4] = {-1, -1, -1, -1};  // albedo, normal, metallic, roughness
  int btf = -1;
};
struct Scene {
  std::vector<Geometry> geoms;
  std::vector<rtf_material> mats;
  std::vector<MatExtra> mat_extra;  // parallel to mats (resized on demand)
  std::vector<Texture> textures;
  std::vector<BTF> btfs;
  int env_faces[6] = {-1, -1, -1, -1, -1, -1};  // cube map of the Scene lighting type (+X -X +Y -Y +Z -Z)
  int brdf = RTF_BRDF_REFERENCE;                // option "brdf"
  std::vector<Instance> insts;
  // flattened, commit order (global primitive id)
  std::vector<float> pos9, nrm9;
  std::vector<uint32_t> tri_inst;
  BVH bvh;
  bool committed = false;
};

// TriangleIlluminationEstimator.cpp:110-118: position = M*(p,1), normal = M*(n,0), NOT renormalised.
static inline void flatten(Scene& sc) {
  sc.pos9.clear(), sc.nrm9.clear(), sc.tri_inst.clear();
  for (size_t ii = 0; ii < sc.insts.size(); ii++) {
    Instance& in = sc.insts[ii];
    const Geometry& g = sc.geoms[in.geom];
    in.first_prim = (uint32_t)sc.tri_inst.size();
    for (size_t t = 0; t < g.tris.size() / 3; t++) {
      for (int k = 0; k < 3; k++) {
        const rtf_vertex& v = g.verts[g.tris[3 * t + k]];
        V3 p = mat4_mul_point(in.xform, v3(v.position[0], v.position[1], v.position[2]), 1.0f);
        V3 n = mat4_mul_point(in.xform, v3(v.normal[0], v.normal[1], v.normal[2]), 0.0f);
        sc.pos9.push_back(p.x), sc.pos9.push_back(p.y), sc.pos9.push_back(p.z);
        sc.nrm9.push_back(n.x), sc.nrm9.push_back(n.y), sc.nrm9.push_back(n.z);
      }
      sc.tri_inst.push_back((uint32_t)ii);
    }
  }
}
static inline void commit(Scene& sc, int bits, int max_split = 1) {
  flatten(sc);
  build(sc.bvh, (uint32_t)sc.tri_inst.size(), sc.pos9.data(), sc.tri_inst.data(), bits, max_split);
  sc.committed = true;
}

// ---- RTU.cuh:89-145 ------------------------------------------------------------------------------
static inline V3 reflect(V3 i, V3 n) { return i - (2.0f * dot(i, n)) * n; }
static inline void tangent_space(V3 n, V3& t, V3& b) {  // RTU.cuh:110-120
  V3 helper = v3(1, 0, 0);
  if (fabsf(n.x) > 0.99f) helper = v3(0, 0, 1);
  t = normalize(cross(n, helper));
  b = normalize(cross(n, t));
}
static inline V3 to_world(V3 n, float cos_theta, float r_phi) {
  float sin_theta = sqrtf(fmaxf(0.0f, 1.0f - cos_theta * cos_theta));
  float s, c;
  if (t_seq_rng) {  // test mode (oracle_math.h): the reference's own libm calls, glm::cos / glm::sin(2 pi u), RTU.cuh:127-131
    const float phi = 2.0f * 3.14159265358979323846264338327950288f * r_phi;
    c = cosf(phi), s = sinf(phi);
  } else {
    sincos2pi(r_phi, &s, &c);
  }
  V3 l = v3(c * sin_theta, s * sin_theta, cos_theta);
  V3 t, b;
  tangent_space(n, t, b);
  return (t * l.x + b * l.y) + n * l.z;
}
// RandomSampleHemisphere(rng, n): cosTheta = rnd, phi = 2*pi*rnd (RTU.cuh:135-145)
static inline V3 sample_hemisphere(float r0, float r1, V3 n) { return to_world(n, r0, r1); }
// RandomSampleHemisphere(rng, n, alpha): cosTheta = 1 - rnd*(1-alpha)^2 (RTU.cuh:122-133)
static inline V3 sample_lobe(float r0, float r1, V3 n, float alpha) {
  float cos_theta = 1.0f - (r0 * (1.0f - alpha)) * (1.0f - alpha);
  return to_world(n, cos_theta, r1);
}

// ---- Env.cuh:28-175 ------------------------------------------------------------------------------
static inline bool solve_quadratic(float a, float b, float c, float& x1, float& x2) {
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
static inline bool ray_sphere(V3 o, V3 d, float radius, float& t0, float& t1) {
  float A = (d.x * d.x + d.y * d.y) + d.z * d.z;
  float B = 2.0f * ((d.x * o.x + d.y * o.y) + d.z * o.z);
  float C = ((o.x * o.x + o.y * o.y) + o.z * o.z) - radius * radius;
  if (!solve_quadratic(A, B, C, t0, t1)) return false;
  if (t0 > t1) std::swap(t0, t1);
  return true;
}
static inline V3 nishita_sky(V3 position, V3 dir, const rtf_env& env) {  // Env.cuh:75-145
  float earth = env.atmosphere.earth_radius * 1000.0f;
  float atmos = env.atmosphere.atmosphere_radius * 1000.0f;
  float Hr = env.atmosphere.hr, Hm = env.atmosphere.hm;
  V3 betaR = v3(3.8e-6f, 13.5e-6f, 33.1e-6f);
  V3 betaM = v3(21e-6f, 21e-6f, 21e-6f);
  V3 sun = v3(env.sun_direction[0], env.sun_direction[1], env.sun_direction[2]);
  float tmin = 0, tmax = 999999999999.0f;
  V3 orig = position + v3(0.0f, earth, 0.0f);
  float t0, t1;
  if (!ray_sphere(orig, dir, atmos, t0, t1) || t1 < 0.0f) return v3(0, 0, 0);
  if (t0 > tmin && t0 > 0.0f) tmin = t0;
  if (t1 < tmax) tmax = t1;
  unsigned ns = (unsigned)env.atmosphere.num_samples, nl = (unsigned)env.atmosphere.num_samples_light;
  float seg = (tmax - tmin) / ns;
  float tcur = tmin;
  V3 sumR = v3(0, 0, 0), sumM = v3(0, 0, 0);
  float odR = 0, odM = 0;
  float mu = dot(dir, sun);
  float phaseR = 3.f / (16.f * 3.1415926f) * (1.0f + mu * mu);
  float g = env.atmosphere.g;
  float phaseM = 3.f / (8.f * 3.1415926f) * ((1.f - g * g) * (1.f + mu * mu)) /
                 ((2.f + g * g) * powf(1.f + g * g - 2.f * g * mu, 1.5f));
  for (unsigned i = 0; i < ns; ++i) {
    V3 sp = orig + (tcur + seg * 0.5f) * dir;
    float height = length(sp) - earth;
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
      V3 spl = sp + (tcl + segl * 0.5f) * sun;
      float hl = length(spl) - earth;
      if (hl < 0) break;
      olR += expf(-hl / Hr) * segl;
      olM += expf(-hl / Hm) * segl;
      tcl += segl;
    }
    if (j == nl) {
      V3 tau = betaR * (odR + olR) + (betaM * 1.1f) * (odM + olM);
      V3 att = v3(expf(-tau.x), expf(-tau.y), expf(-tau.z));
      sumR = sumR + att * hr;
      sumM = sumM + att * hm;
    }
    tcur += seg;
  }
  return ((sumR * betaR) * phaseR + (sumM * betaM) * phaseM) * 20.0f;
}
static inline V3 env_light(V3 position, V3 dir, const rtf_env& env, const Scene* sc = nullptr) {  // Env.cuh:147-175
  V3 c = v3(1, 1, 1);
  V3 col = v3(env.color[0], env.color[1], env.color[2]);
  switch (env.lighting_type) {
    case RTF_ENV_SCENE:
      if (env.environmental_map_id != 0 && sc && sc->env_faces[0] >= 0) {
        const Texture* faces[6];
        for (int k = 0; k < 6; k++) faces[k] = &sc->textures[sc->env_faces[k]];
        float t[4];
        sample_cube_map(faces, dir, t);
        col = v3(t[0], t[1], t[2]);
      }
      c = col * env.skylight_intensity;
      break;
    case RTF_ENV_SKYDOME: c = nishita_sky(position, dir, env) * env.skylight_intensity; break;
    case RTF_ENV_SINGLE_LIGHT_SOURCE: c = col * env.skylight_intensity; break;
  }
  if (env.gamma != 1.0f) {
    float ig = 1.0f / env.gamma;
    c = v3(powf(c.x, ig), powf(c.y, ig), powf(c.z, ig));
  }
  return vmax(v3(0, 0, 0), c);
}

// Lambert + GGX importance sampling (option "brdf" = RTF_BRDF_LAMBERT_GGX, BASELINE config 2).  Not in the reference (its
// only BRDF is the mirror lobe, BSDF.cuh:6-13): cosine-weighted Lambert with albedo*(1-metallic)*0.96, GGX visible-normal
// sampling, Schlick Fresnel with F0 = mix(0.04, albedo, metallic), separable Smith G, p_spec = (1 + metallic)/2.
// Validated against its own closed forms in tests/ (nothing in the reference to compare with).
static inline V3 sample_lambert_ggx(V3 albedo, float metallic, float roughness, V3 n, V3 v, float r0, float r1, float r2,
                                    V3& l) {
  const float p_spec = 0.5f + 0.5f * clampf(metallic, 0.0f, 1.0f);
  const float nv = dot(n, v);
  if (r2 >= p_spec) {
    l = to_world(n, sqrtf(1.0f - r0), r1);
    const float k = ((1.0f - metallic) * 0.96f) / (1.0f - p_spec);
    return albedo * k;
  }
  // specular lobe: GGX distribution of visible normals (Heitz 2018) — the weight F * G1(l) never exceeds 1
  if (!(nv > 0.0f)) return v3(0.0f, 0.0f, 0.0f);
  const float a = fmaxf(roughness * roughness, 1e-3f), a2 = a * a;
  V3 helper = v3(1.0f, 0.0f, 0.0f);
  if (fabsf(n.x) > 0.99f) helper = v3(0.0f, 0.0f, 1.0f);
  const V3 tx = normalize(cross(n, helper)), ty = normalize(cross(n, tx));
  const float vx = dot(v, tx), vy = dot(v, ty);
  // stretch the view vector, build an orthonormal basis around it
  const float sx = a * vx, sy = a * vy;
  const float inv_len = 1.0f / sqrtf((sx * sx + sy * sy) + nv * nv);
  const float hx = sx * inv_len, hy = sy * inv_len, hz = nv * inv_len;
  const float lensq = hx * hx + hy * hy;
  float t1x = 1.0f, t1y = 0.0f;
  if (lensq > 0.0f) {
    const float il = 1.0f / sqrtf(lensq);
    t1x = -hy * il, t1y = hx * il;
  }
  const float t2x = -hz * t1y, t2y = hz * t1x, t2z = hx * t1y - hy * t1x;  // cross(Vh, T1)
  // sample the projected disk, warp the half that faces away
  const float rr = sqrtf(r0);
  float sn, cs;
  sincos2pi(r1, &sn, &cs);
  const float p1 = rr * cs;
  float p2 = rr * sn;
  const float sw = 0.5f * (1.0f + hz);
  p2 = (1.0f - sw) * sqrtf(fmaxf(0.0f, 1.0f - p1 * p1)) + sw * p2;
  const float p3 = sqrtf(fmaxf(0.0f, (1.0f - p1 * p1) - p2 * p2));
  const float nhx = (p1 * t1x + p2 * t2x) + p3 * hx, nhy = (p1 * t1y + p2 * t2y) + p3 * hy, nhz = p2 * t2z + p3 * hz;
  // unstretch
  const float ex = a * nhx, ey = a * nhy, ez = fmaxf(0.0f, nhz);
  const float ie = 1.0f / sqrtf((ex * ex + ey * ey) + ez * ez);
  const V3 h = (tx * (ex * ie) + ty * (ey * ie)) + n * (ez * ie);
  const float vh = dot(v, h);
  l = (2.0f * vh) * h - v;
  const float nl = dot(n, l);
  if (!(nl > 0.0f) || !(vh > 0.0f)) return v3(0.0f, 0.0f, 0.0f);
  const float g1l = (2.0f * nl) / (nl + sqrtf(a2 + (1.0f - a2) * (nl * nl)));
  const float m = 1.0f - vh, m2 = m * m, m5 = (m2 * m2) * m;
  const V3 f0 = v3(0.04f, 0.04f, 0.04f) * (1.0f - metallic) + albedo * metallic;
  const V3 fr = f0 + (v3(1.0f, 1.0f, 1.0f) - f0) * m5;
  return fr * (g1l / p_spec);
}

// ---- hit reconstruction: RDD.hpp:127-164 (TriangularMesh::GetHitInfo) + :364-382 (SBT::GetHitInfo) ----
struct SurfaceHit {
  V3 position, normal;
  uint32_t inst;
};
static inline SurfaceHit surface_hit(const Scene& sc, const Hit& h, V3 ray_dir) {
  // TriangularMesh::GetHitInfo interpolates the OBJECT-space vertices (RDD.hpp:127-149); SBT::GetHitInfo then transforms:
  // normal = normalize(M * (n, 0)), faced against the ray, position = M * (p, 1) (RDD.hpp:375-380).
  SurfaceHit s;
  s.inst = sc.tri_inst[h.prim];
  const Instance& in = sc.insts[s.inst];
  const Geometry& g = sc.geoms[in.geom];
  const uint32_t lt = h.prim - in.first_prim;
  const rtf_vertex& vx = g.verts[g.tris[3 * lt]];
  const rtf_vertex& vy = g.verts[g.tris[3 * lt + 1]];
  const rtf_vertex& vz = g.verts[g.tris[3 * lt + 2]];
  float w0 = (1.0f - h.u) - h.v;
  V3 pl = (w0 * v3(vx.position[0], vx.position[1], vx.position[2]) + h.u * v3(vy.position[0], vy.position[1], vy.position[2])) +
          h.v * v3(vz.position[0], vz.position[1], vz.position[2]);
  V3 nl = (w0 * v3(vx.normal[0], vx.normal[1], vx.normal[2]) + h.u * v3(vy.normal[0], vy.normal[1], vy.normal[2])) +
          h.v * v3(vz.normal[0], vz.normal[1], vz.normal[2]);
  V3 nn = normalize(mat4_mul_point(in.xform, nl, 0.0f));
  if (dot(ray_dir, nn) > 0.0f) nn = neg(nn);
  s.normal = nn;
  s.position = mat4_mul_point(in.xform, pl, 1.0f);
  return s;
}
// Full HitInfo as written by __closesthit__PCS_R (PCS.cu:15-28).
static inline void full_hit_info(const Scene& sc, const Hit& h, V3 ray_dir, rtf_hit_info& out) {
  SurfaceHit s = surface_hit(sc, h, ray_dir);
  const Instance& in = sc.insts[s.inst];
  const Geometry& g = sc.geoms[in.geom];
  uint32_t lt = h.prim - in.first_prim;
  const rtf_vertex& vx = g.verts[g.tris[3 * lt]];
  const rtf_vertex& vy = g.verts[g.tris[3 * lt + 1]];
  const rtf_vertex& vz = g.verts[g.tris[3 * lt + 2]];
  float z = (1.0f - h.u) - h.v;
  V3 tg = (z * v3(vx.tangent[0], vx.tangent[1], vx.tangent[2]) + h.u * v3(vy.tangent[0], vy.tangent[1], vy.tangent[2])) +
          h.v * v3(vz.tangent[0], vz.tangent[1], vz.tangent[2]);
  tg = normalize(mat4_mul_point(in.xform, tg, 0.0f));
  out.position[0] = s.position.x, out.position[1] = s.position.y, out.position[2] = s.position.z;
  out.normal[0] = s.normal.x, out.normal[1] = s.normal.y, out.normal[2] = s.normal.z;
  out.tangent[0] = tg.x, out.tangent[1] = tg.y, out.tangent[2] = tg.z;
  out.tex_coord[0] = (z * vx.tex_coord[0] + h.u * vy.tex_coord[0]) + h.v * vz.tex_coord[0];
  out.tex_coord[1] = (z * vx.tex_coord[1] + h.u * vy.tex_coord[1]) + h.v * vz.tex_coord[1];
  const rtf_vertex* pick = (h.u > z && h.u > h.v) ? &vx : (h.v > z ? &vy : &vz);  // RDD.hpp:150-160 (as written)
  for (int k = 0; k < 4; k++) out.color[k] = pick->color[k];
  out.data[0] = pick->position_padding, out.data[1] = pick->normal_padding;
  out.data[2] = pick->tangent_padding, out.data[3] = pick->tex_coord_padding[0];
}

// Interpolated vertex attributes of a hit (TriangularMesh::GetHitInfo, RDD.hpp:127-164; tangent through
// SBT::GetHitInfo, RDD.hpp:379) for the materials that read them.
struct HitAttrs {
  float u, v;   // texture coordinate
  V3 tangent;   // normalize(M * t)
  V3 color;     // colour of the dominant-barycentric vertex
};
static inline HitAttrs hit_attrs(const Scene& sc, uint32_t prim, float bu, float bv, uint32_t inst) {
  const Instance& in = sc.insts[inst];
  const Geometry& g = sc.geoms[in.geom];
  const uint32_t lt = prim - in.first_prim;
  const rtf_vertex& vx = g.verts[g.tris[3 * lt]];
  const rtf_vertex& vy = g.verts[g.tris[3 * lt + 1]];
  const rtf_vertex& vz = g.verts[g.tris[3 * lt + 2]];
  const float z = (1.0f - bu) - bv;
  HitAttrs a;
  V3 tg = (z * v3(vx.tangent[0], vx.tangent[1], vx.tangent[2]) + bu * v3(vy.tangent[0], vy.tangent[1], vy.tangent[2])) +
          bv * v3(vz.tangent[0], vz.tangent[1], vz.tangent[2]);
  a.tangent = normalize(mat4_mul_point(in.xform, tg, 0.0f));
  a.u = (z * vx.tex_coord[0] + bu * vy.tex_coord[0]) + bv * vz.tex_coord[0];
  a.v = (z * vx.tex_coord[1] + bu * vy.tex_coord[1]) + bv * vz.tex_coord[1];
  const rtf_vertex* pick = (bu > z && bu > bv) ? &vx : (bv > z ? &vy : &vz);  // RDD.hpp:150-160 (as written)
  a.color = v3(pick->color[0], pick->color[1], pick->color[2]);
  return a;
}
static inline const MatExtra& mat_extra(const Scene& sc, uint32_t mat) {
  static const MatExtra none;
  return mat < sc.mat_extra.size() ? sc.mat_extra[mat] : none;
}

// any-hit alpha test (RF.cuh:5-23): Default material only; alpha = albedo texture's w, else 1 - transmission
// (SurfaceMaterial::GetAlbedo, RDD.hpp:248-254).
struct AlphaFilter {
  const Scene* sc;
  uint32_t seed;
  uint64_t unit;
  uint32_t bounce;
  bool operator()(const Tri48& tri, const Hit& cand) const {
    const uint32_t mi = sc->insts[tri.inst].mat;
    const rtf_material& m = sc->mats[mi];
    if (m.material_type != RTF_MAT_DEFAULT) return true;
    float alpha = 1.0f - m.transmission;
    const int at = mat_extra(*sc, mi).tex[0];
    if (at >= 0) {
      const HitAttrs a = hit_attrs(*sc, tri.prim, cand.u, cand.v, tri.inst);
      float t[4];
      tex2d(sc->textures[at], a.u, a.v, t);
      alpha = t[3];
    } else if (alpha >= 1.0f && !t_seq_rng) {
      return true;  // 1 <= rnd never holds; the reference still draws (RF.cuh:19), which only a sequential source sees
    }
    float r[4];
    path_draws(seed, STREAM_ALPHA, unit, bounce, tri.prim, r, 1);
    return !(alpha <= r[0]);
  }
};

struct PathResult {
  V3 energy, normal, albedo, position;  // PerRayData, RTU.cuh:26-34
  bool first_hit;
  uint32_t rays;
};

// __closesthit__*_R -> ClosestHitFunc (RF.cuh:25-258: surface materials :44-172, compressed BTF :172-258),
// __miss__*_R -> MissFunc (RF.cuh:260-272), unrolled.  `o,d,tmin` is the primary ray; returns radiance + first-hit
// records.
static inline PathResult trace_path(const Scene& sc, const rtf_env& env, int bounces, uint32_t seed, uint64_t unit,
                                    V3 o, V3 d, float tmin, TraceStats* stats) {
  PathResult pr;
  pr.energy = v3(0, 0, 0), pr.normal = v3(0, 0, 0), pr.albedo = v3(0, 0, 0), pr.position = v3(0, 0, 0);
  pr.first_hit = false;
  pr.rays = 0;
  V3 T = v3(1, 1, 1), acc = v3(0, 0, 0);
  uint32_t hit_count = 0;
  uint32_t wave = 0;  // radiance rays traced so far on this path = the wavefront index the GPU traces this ray in; keys the
                      // any-hit draws (equal to hit_count except after a BSSRDF resume, where two rays share a hit count)
  bool any_hit_enabled = true;  // BTF bounce rays run with OPTIX_RAY_FLAG_DISABLE_ANYHIT (RF.cuh:239)
  // BSSRDF (RF.cuh:97-133): the first hit may start TWO radiance paths, the subsurface one first.  Both share
  // PerRayData::m_hitCount, which is never reset: the ordinary bounce path is traced after the subsurface path has
  // returned and starts from the hit count the subsurface path ended with.
  struct {
    bool on = false;
    V3 o = v3(0, 0, 0), d = v3(0, 0, 0), T = v3(0, 0, 0);
    // test mode (t_seq_rng) only: the reference draws the ordinary bounce AFTER the subsurface path has returned
    // (RF.cuh:134-136), so with a sequential stream the draw is made at resume time from what is kept here
    bool deferred = false;
    V3 normal = v3(0, 0, 0), d_in = v3(0, 0, 0), base = v3(0, 0, 0);  // base = (1 - subsurface_factor) * albedo
    float metallic = 0.0f, roughness = 0.0f, f = 0.0f;
  } pend;
  auto resume = [&]() -> bool {
    if (!pend.on) return false;
    pend.on = false;
    if (pend.deferred) {
      pend.deferred = false;
      float rr[4];
      path_draws(seed, STREAM_PATH, unit, 0, 0, rr, 2);
      const V3 nd = sample_lobe(rr[0], rr[1], reflect(pend.d_in, pend.normal), pend.metallic);
      const float k = clampf(fabsf(dot(pend.normal, nd)) * pend.roughness + (1.0f - pend.roughness) * pend.f, 0.0f, 1.0f);
      pend.d = nd, pend.T = pend.T * (pend.base * k);
    }
    o = pend.o, d = pend.d, T = pend.T, tmin = 1e-3f;
    any_hit_enabled = true;
    return true;
  };
  for (;;) {
    Ray ray{o, tmin, d, 1e20f};
    Hit h;
    if (any_hit_enabled) {
      AlphaFilter flt{&sc, seed, unit, wave};
      h = trace(sc.bvh, ray, stats, false, flt);
    } else {
      h = trace(sc.bvh, ray, stats, false);
    }
    wave++;
    any_hit_enabled = true;
    pr.rays++;
    if (h.prim == 0xFFFFFFFFu) {
      V3 L = env_light(o, d, env, &sc);
      acc = acc + T * L;
      if (hit_count == 0) pr.albedo = L;  // MissFunc: m_albedo = m_energy = env colour
      if (resume()) continue;
      break;
    }
    hit_count++;
    SurfaceHit sh = surface_hit(sc, h, d);
    const uint32_t mi = sc.insts[sh.inst].mat;
    const rtf_material& m = sc.mats[mi];
    const MatExtra& mx = mat_extra(sc, mi);
    const bool rich = m.material_type != RTF_MAT_DEFAULT || mx.tex[0] >= 0 || mx.tex[1] >= 0 || mx.tex[2] >= 0 || mx.tex[3] >= 0;
    HitAttrs at{0.0f, 0.0f, v3(0, 0, 0), v3(0, 0, 0)};
    if (rich) at = hit_attrs(sc, h.prim, h.u, h.v, sh.inst);
    float r[4];
    if (m.material_type == RTF_MAT_COMPRESSED_BTF) {  // RF.cuh:172-258
      const BTF& btf = sc.btfs[mx.btf];
      V3 btf_color = v3(0, 0, 0);  // the reference leaves it uninitialised past the bounce limit: 0 here
      bool go_on = false;
      if ((int)hit_count <= bounces) {
        path_draws(seed, STREAM_PATH, unit, hit_count, 0, r, 2);
        if (env.lighting_type == RTF_ENV_SINGLE_LIGHT_SOURCE) {  // RF.cuh:180-218
          V3 sun = v3(env.sun_direction[0], env.sun_direction[1], env.sun_direction[2]);
          V3 nd = sample_lobe(r[0], r[1], sun, 1.0f - env.light_size);
          btf_color = btf_get_value(btf, at.u, at.v, d, nd, sh.normal, at.tangent);
          V3 col = v3(env.color[0], env.color[1], env.color[2]);
          acc = acc + T * ((col * env.ambient_light_intensity) * btf_color);
          float ndl = dot(sh.normal, nd);
          if (ndl > 0.0f) {
            Ray sr{sh.position + sh.normal * 1e-3f, 1e-3f, nd, 1e20f};
            Hit sh2 = trace(sc.bvh, sr, stats, true);
            pr.rays++;
            if (sh2.prim == 0xFFFFFFFFu) {
              V3 L = env_light(sr.o, nd, env, &sc);
              acc = acc + T * ((L * ndl) * btf_color);
            }
          }
        } else {  // RF.cuh:219-246
          V3 refl = reflect(d, sh.normal);
          V3 nd = sample_lobe(r[0], r[1], refl, 0.0f);
          btf_color = btf_get_value(btf, at.u, at.v, d, nd, sh.normal, at.tangent);
          T = T * btf_color;
          o = sh.position + sh.normal * 1e-3f, d = nd, tmin = 1e-3f;
          any_hit_enabled = false;
          go_on = true;
        }
      }
      if (hit_count == 1) pr.normal = sh.normal, pr.albedo = btf_color, pr.position = sh.position, pr.first_hit = true;
      if (go_on) continue;
      if (resume()) continue;
      break;
    }
    // ---- surface materials, RF.cuh:44-172 ----
    V3 normal = sh.normal;
    if (mx.tex[1] >= 0) {  // ApplyNormalTexture, RDD.hpp:268-281
      float t[4];
      tex2d(sc.textures[mx.tex[1]], at.u, at.v, t);
      V3 B = cross(normal, at.tangent);
      V3 tn = v3(t[0] * 2.0f - 1.0f, t[1] * 2.0f - 1.0f, t[2] * 2.0f - 1.0f);
      normal = normalize((at.tangent * tn.x + B * tn.y) + normal * tn.z);
    }
    float metallic = m.metallic, roughness = m.roughness;
    if (mx.tex[2] >= 0) {
      float t[4];
      tex2d(sc.textures[mx.tex[2]], at.u, at.v, t);
      metallic = t[0];
    }
    if (mx.tex[3] >= 0) {
      float t[4];
      tex2d(sc.textures[mx.tex[3]], at.u, at.v, t);
      roughness = t[0];
    }
    V3 albedo = v3(m.albedo_color[0], m.albedo_color[1], m.albedo_color[2]);
    if (m.material_type != RTF_MAT_DEFAULT) {
      albedo = at.color;  // VertexColor: hitInfo.m_color (RF.cuh:55-57)
    } else if (mx.tex[0] >= 0) {
      float t[4];
      tex2d(sc.textures[mx.tex[0]], at.u, at.v, t);
      albedo = v3(t[0], t[1], t[2]);
    }
    if (hit_count == 1) pr.normal = normal, pr.albedo = albedo, pr.position = sh.position, pr.first_hit = true;
    float f = 1.0f;
    if (metallic >= 0.0f) f = (metallic + 2) / (metallic + 1);
    V3 emis = m.emission * albedo;
    if (env.lighting_type == RTF_ENV_SINGLE_LIGHT_SOURCE) {  // RF.cuh:61-92
      path_draws(seed, STREAM_PATH, unit, hit_count, 0, r, 2);
      V3 sun = v3(env.sun_direction[0], env.sun_direction[1], env.sun_direction[2]);
      V3 nd = sample_lobe(r[0], r[1], sun, 1.0f - env.light_size);
      V3 col = v3(env.color[0], env.color[1], env.color[2]);
      acc = acc + T * ((col * env.ambient_light_intensity) * albedo);
      float ndl = dot(normal, nd);
      if (ndl > 0.0f) {
        Ray sr{sh.position, 1e-3f, nd, 1e20f};
        Hit sh2 = trace(sc.bvh, sr, stats, true);
        pr.rays++;
        if (sh2.prim == 0xFFFFFFFFu) {
          V3 L = env_light(sh.position, nd, env, &sc);
          acc = acc + T * ((L * ndl) * albedo);
        }
      }
      acc = acc + T * emis;
      break;
    } else if ((int)hit_count <= bounces) {  // RF.cuh:93-161
      // ---- BSSRDF, RF.cuh:97-133 + BSSDF.cuh:42-89 ----
      bool need_sample = false;
      V3 ss_o = v3(0, 0, 0), ss_d = v3(0, 0, 0), ss_w = v3(0, 0, 0);
      if (hit_count <= 1 && m.subsurface_factor > 0.0f && m.subsurface_radius[0] > 0.0f) {
        const float radius = m.subsurface_radius[0];
        float rs[4], rb[4];
        path_draws(seed, STREAM_SSS, unit, hit_count, 0, rs, 2);  // disk radius, rotation angle (BSSDF.cuh:51-54)
        const V3 disk_center = sh.position + (radius * normal) * 0.5f;
        const float disk_radius = radius * sqrtf(rs[0]);
        const float distance = sqrtf(radius * radius - disk_radius * disk_radius);
        const V3 sample_pos = disk_center + disk_radius * rotate_about(v3(normal.y, normal.z, normal.x), rs[1], normal);
        const uint64_t handle = sc.insts[sh.inst].handle;
        const SSRecords rec = ss_probe(sc.bvh, sample_pos, neg(normal), distance, radius + distance, handle,
                                       [&](uint32_t inst) { return sc.insts[inst].handle; });
        if (rec.n > 0) {
          // The reference picks the record with a COPY of the generator (BSSDF.cuh:58,77), which leaves the main stream
          // where it was: the pick is the very number the exit lobe's first draw returns afterwards (BSSDF.cuh:83).
          float pick;
          if (t_seq_rng) {
            LCG copy = *t_seq_rng;
            pick = copy.next();
          } else {
            path_rng(seed, STREAM_SSS, unit, hit_count, 1, rb);
            pick = rb[0];
          }
          int index = (int)(pick * (float)rec.n);
          index = index < 0 ? 0 : (index > rec.n - 1 ? rec.n - 1 : index);
          Hit rh{rec.t[index], rec.u[index], rec.v[index], rec.prim[index]};
          const SurfaceHit so = surface_hit(sc, rh, v3(0, 0, 0));  // GetHitInfo(rayDirection, false): no flip
          V3 out_n = so.normal;
          const uint32_t omi = sc.insts[so.inst].mat;
          const MatExtra& omx = mat_extra(sc, omi);
          if (omx.tex[1] >= 0) {  // ApplyNormalTexture of the record's own material (BSSDF.cuh:31-32)
            const HitAttrs oa = hit_attrs(sc, rh.prim, rh.u, rh.v, so.inst);
            float t[4];
            tex2d(sc.textures[omx.tex[1]], oa.u, oa.v, t);
            V3 B = cross(out_n, oa.tangent);
            V3 tn = v3(t[0] * 2.0f - 1.0f, t[1] * 2.0f - 1.0f, t[2] * 2.0f - 1.0f);
            out_n = normalize((oa.tangent * tn.x + B * tn.y) + out_n * tn.z);
          }
          if (length(sh.position - so.position) <= radius) {
            ss_o = so.position + out_n * 0.01f;
            path_draws(seed, STREAM_SSS, unit, hit_count, 1, rb, 2);
            ss_d = sample_lobe(rb[0], rb[1], reflect(neg(out_n), out_n), metallic);  // BRDF(metallic, rnd, -n, n)
            const float k2 = clampf(fabsf(dot(out_n, ss_d)) * roughness + (1.0f - roughness) * f, 0.0f, 1.0f);
            ss_w = (m.subsurface_factor * v3(m.subsurface_color[0], m.subsurface_color[1], m.subsurface_color[2])) * k2;
            need_sample = true;
          }
        }
      }
      if (t_seq_rng && need_sample) {  // test mode: subsurface path first, the bounce is drawn when it resumes
        acc = acc + T * emis;
        pend.on = true, pend.deferred = true, pend.o = sh.position, pend.T = T;
        pend.normal = normal, pend.d_in = d, pend.base = (1.0f - m.subsurface_factor) * albedo;
        pend.metallic = metallic, pend.roughness = roughness, pend.f = f;
        T = T * ss_w;
        o = ss_o, d = ss_d, tmin = 1e-3f;
        continue;
      }
      path_draws(seed, STREAM_PATH, unit, hit_count, 0, r, sc.brdf == RTF_BRDF_LAMBERT_GGX ? 3 : 2);
      V3 nd, wgt;
      bool absorbed = false;
      if (sc.brdf == RTF_BRDF_LAMBERT_GGX) {
        wgt = sample_lambert_ggx(albedo, metallic, roughness, normal, neg(d), r[0], r[1], r[2], nd);
      } else {
        V3 refl = reflect(d, normal);
        nd = sample_lobe(r[0], r[1], refl, metallic);
        float k = clampf(fabsf(dot(normal, nd)) * roughness + (1.0f - roughness) * f, 0.0f, 1.0f);
        wgt = ((1.0f - m.subsurface_factor) * albedo) * k;
      }
      absorbed = wgt.x == 0.0f && wgt.y == 0.0f && wgt.z == 0.0f;  // a zero-weight ray is not traced (result unchanged)
      acc = acc + T * emis;
      if (need_sample) {  // subsurface path first; the bounce waits for it
        if (!absorbed) pend.on = true, pend.o = sh.position, pend.d = nd, pend.T = T * wgt;
        T = T * ss_w;
        o = ss_o, d = ss_d, tmin = 1e-3f;
        continue;
      }
      if (absorbed) {
        if (resume()) continue;
        break;
      }
      T = T * wgt;
      o = sh.position, d = nd, tmin = 1e-3f;
      continue;
    } else {
      acc = acc + T * emis;
      if (resume()) continue;
      break;
    }
  }
  pr.energy = acc;
  return pr;
}

// 32-lane strided accumulation + xor butterfly: the fixed reduction tree shared with the GPU.
struct Lanes6 {
  float v[32][6];
  Lanes6() { memset(v, 0, sizeof(v)); }
  void add(int idx, const float x[6]) {
    for (int k = 0; k < 6; k++) v[idx & 31][k] = v[idx & 31][k] + x[k];
  }
  void reduce(float out[6]) {
    for (int off = 16; off >= 1; off >>= 1) {
      float nv[32][6];
      for (int l = 0; l < 32; l++)
        for (int k = 0; k < 6; k++) nv[l][k] = v[l][k] + v[l ^ off][k];
      memcpy(v, nv, sizeof(v));
    }
    for (int k = 0; k < 6; k++) out[k] = v[0][k];
  }
};

// __raygen__IE (IE.cu:30-129) for one probe.  Returns raw sums {E.xyz, D.xyz} over the enabled faces
// and samples [s0,s1) of S_total.
static inline void illum_probe(const Scene& sc, const rtf_env& env, const rtf_ray_props& rp, uint32_t seed, float push,
                               uint64_t probe_id, const V3 P[3], const V3 N[3], bool front, bool back, int s0, int s1,
                               float out6[6], uint64_t* rays, TraceStats* stats) {
  for (int k = 0; k < 6; k++) out6[k] = 0.0f;
  for (int face = 0; face < 2; face++) {
    if (face == 0 ? !front : !back) continue;
    Lanes6 lanes;
    for (int s = s0; s < s1; s++) {
      uint64_t unit = (probe_id * 2 + (uint64_t)face) * (uint64_t)rp.samples + (uint64_t)s;
      float r[4];
      path_draws(seed, STREAM_PATH, unit, 0, 0, r, 4);
      float a = r[0], b = r[1];
      float w = (1.0f - a) - b;
      V3 position = (w * P[0] + a * P[1]) + b * P[2];
      V3 normal = (w * N[0] + a * N[1]) + b * N[2];
      if (face == 1) normal = neg(normal);
      V3 dir = sample_hemisphere(r[2], r[3], normal);
      V3 origin = position + normal * push;
      PathResult pr = trace_path(sc, env, rp.bounces, seed, unit, origin, dir, 1e-3f, stats);
      if (rays) *rays += pr.rays;
      V3 e = pr.energy * fabsf(dot(normal, dir));
      V3 dd = dir * length(e);
      float x[6] = {e.x, e.y, e.z, dd.x, dd.y, dd.z};
      lanes.add(s - s0, x);
    }
    float red[6];
    lanes.reduce(red);
    for (int k = 0; k < 6; k++) out6[k] = out6[k] + red[k];
  }
}
static inline void illum_finalize(float v[6], int faces, int samples) {  // IE.cu:125-128
  int n = faces * samples;
  if (n == 0) return;
  float fn = (float)n;
  v[0] = v[0] / fn, v[1] = v[1] / fn, v[2] = v[2] / fn;
  V3 d = normalize(v3(v[3], v[4], v[5]));
  v[3] = d.x, v[4] = d.y, v[5] = d.z;
}

// __raygen__CR (CR.cu:32-147) for one pixel: per-sample path results summed in sample order.
struct PixelOut {
  V3 color, normal, albedo, position;
};
static inline PixelOut camera_pixel(const Scene& sc, const rtf_env& env, const rtf_ray_props& rp,
                                    const rtf_camera_props& cam, int ix, int iy, uint64_t* rays, TraceStats* stats) {
  PixelOut po{v3(0, 0, 0), v3(0, 0, 0), v3(0, 0, 0), v3(0, 0, 0)};
  const int W = cam.width, H = cam.height;
  const uint64_t pixel = (uint64_t)ix + (uint64_t)W * (uint64_t)iy;
  const float halfX = W / 2.0f, halfY = H / 2.0f;
  const float* m = cam.inverse_projection_view;
  const uint32_t seed = (uint32_t)cam.frame_id;
  const float nsf = (float)rp.samples;
  V3 hor = v3(cam.horizontal[0], cam.horizontal[1], cam.horizontal[2]);
  V3 ver = v3(cam.vertical[0], cam.vertical[1], cam.vertical[2]);
  for (int s = 0; s < rp.samples; s++) {
    uint64_t unit = pixel * (uint64_t)rp.samples + (uint64_t)s;
    float r[4];
    if (t_seq_rng) {
      // test mode (oracle_math.h): CR.cu:60-61 draws both jitters inside one constructor call, glm::vec2(f(rnd()), f(rnd())),
      // whose argument order the language leaves open; g++, which compiles the reference's program for the host run
      // (ref_optix_driver.cpp), evaluates the second argument first.  Then the aperture angle (CR.cu:75).
      r[1] = t_seq_rng->next(), r[0] = t_seq_rng->next(), r[2] = t_seq_rng->next(), r[3] = 0.0f;
    } else {
      path_rng(seed, STREAM_CAMERA, unit, 0, 0, r);
    }
    float sx = (((float)ix + r[0]) - halfX) / halfX;
    float sy = (((float)iy + r[1]) - halfY) / halfY;
    float st[4], en[4];
    for (int k = 0; k < 4; k++) {
      st[k] = (m[k] * sx + m[4 + k] * sy) + (m[8 + k] * -1.0f + m[12 + k] * 1.0f);
      en[k] = (m[k] * sx + m[4 + k] * sy) + (m[8 + k] * 1.0f + m[12 + k] * 1.0f);
    }
    V3 start = v3(st[0] / st[3], st[1] / st[3], st[2] / st[3]);
    V3 end = v3(en[0] / en[3], en[1] / en[3], en[2] / en[3]);
    V3 pdir = normalize(end - start);
    V3 conv = start + pdir * cam.focal_length;
    float sn, cs;
    if (t_seq_rng) {  // test mode (oracle_math.h): the reference's own libm calls, CR.cu:75-79
      const float angle = r[2] * 3.1415927f * 2.0f;
      sn = sinf(angle), cs = cosf(angle);
    } else {
      sincos2pi(r[2], &sn, &cs);
    }
    V3 ap = start + cam.aperture * (hor * sn + ver * cs);
    V3 dir = normalize(conv - ap);
    PathResult pr = trace_path(sc, env, rp.bounces, seed, unit, ap, dir, 0.0f, stats);
    if (rays) *rays += pr.rays;
    V3 pos = pr.first_hit ? pr.position : (s == 0 ? v3(999999.0f, 999999.0f, 999999.0f) : v3(0, 0, 0));  // CR.cu:48,108
    po.color = po.color + v3(pr.energy.x / nsf, pr.energy.y / nsf, pr.energy.z / nsf);
    po.normal = po.normal + v3(pr.normal.x / nsf, pr.normal.y / nsf, pr.normal.z / nsf);
    po.albedo = po.albedo + v3(pr.albedo.x / nsf, pr.albedo.y / nsf, pr.albedo.z / nsf);
    po.position = po.position + v3(pos.x / nsf, pos.y / nsf, pos.z / nsf);
  }
  return po;
}
// CR.cu:112-146: progressive accumulation in gamma space and output writes.  prev = current colour buffer texel.
static inline void camera_resolve(const rtf_camera_props& cam, const PixelOut& po, const float prev[4], float color[4],
                                  float normal[4], float albedo[4]) {
  V3 pc = po.color;
  if (cam.accumulate && cam.frame_id > 1) {
    V3 acc = v3(powf(prev[0], cam.gamma), powf(prev[1], cam.gamma), powf(prev[2], cam.gamma));
    float fid = (float)cam.frame_id;
    pc = pc + fid * acc;
    float den = (float)(cam.frame_id + 1);
    pc = v3(pc.x / den, pc.y / den, pc.z / den);
  }
  float ig = (float)(1.0 / (double)cam.gamma);
  color[0] = powf(pc.x, ig), color[1] = powf(pc.y, ig), color[2] = powf(pc.z, ig), color[3] = 1.0f;
  if (cam.output_type == RTF_OUT_DEPTH) {
    V3 from = v3(cam.from[0], cam.from[1], cam.from[2]);
    float dist = length(from - po.position);
    float dd = clampf(dist / cam.max_distance, 0.0f, 1.0f);
    albedo[0] = albedo[1] = albedo[2] = dd, albedo[3] = 1.0f;
  } else {
    albedo[0] = po.albedo.x, albedo[1] = po.albedo.y, albedo[2] = po.albedo.z, albedo[3] = 1.0f;
  }
  normal[0] = po.normal.x, normal[1] = po.normal.y, normal[2] = po.normal.z, normal[3] = 1.0f;
}

}  // namespace orc
