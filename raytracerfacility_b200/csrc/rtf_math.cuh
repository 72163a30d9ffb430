// rtf_math.cuh — host/device arithmetic of the B200 ray-tracing core.
//
// Bit-exactness contract (DESIGN.md "Numerics"): this library is compiled with -fmad=false, so the
// compiler never fuses a*b+c on its own; FMAs appear only where fmaf() is written.  Division and sqrt
// are IEEE (nvcc defaults -prec-div=true -prec-sqrt=true, no -use_fast_math).  With the same
// association on the CPU (-ffp-contract=off) every function here reproduces bit for bit.
//
// Reference anchors (relative to /root/reference/include/RayTracerFacility/RayTracer/):
//   glm conventions used by RayTracerUtilities.cuh / RayFunctions.cuh: dot = (x+y)+z,
//   normalize = v * (1/sqrt(dot(v,v))), mat4*vec4 = (c0*x + c1*y) + (c2*z + c3*w).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define RTF_HD __host__ __device__ __forceinline__
#define RTF_D __device__ __forceinline__
#else
#define RTF_HD inline
#define RTF_D inline
#endif

namespace rtf {

RTF_HD float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
RTF_HD float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
RTF_HD float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
RTF_HD float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
RTF_HD float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
RTF_HD float3 operator*(float s, float3 a) { return f3(a.x * s, a.y * s, a.z * s); }
RTF_HD float3 neg3(float3 a) { return f3(-a.x, -a.y, -a.z); }
RTF_HD float dot3(float3 a, float3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
RTF_HD float3 cross3(float3 a, float3 b) {
  return f3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
RTF_HD float length3(float3 a) { return sqrtf(dot3(a, a)); }
RTF_HD float3 normalize3(float3 a) {
  float inv = 1.0f / sqrtf(dot3(a, a));
  return a * inv;
}
RTF_HD float3 min3(float3 a, float3 b) { return f3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
RTF_HD float3 max3(float3 a, float3 b) { return f3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
RTF_HD float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

RTF_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  union {
    float f;
    uint32_t u;
  } c;
  c.f = f;
  return c.u;
#endif
}
RTF_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  union {
    float f;
    uint32_t u;
  } c;
  c.u = u;
  return c.f;
#endif
}

// glm::mat4 (column-major) * vec4(p, w)
RTF_HD float3 mat4_mul(const float* m, float3 p, float w) {
  float3 r;
  r.x = (m[0] * p.x + m[4] * p.y) + (m[8] * p.z + m[12] * w);
  r.y = (m[1] * p.x + m[5] * p.y) + (m[9] * p.z + m[13] * w);
  r.z = (m[2] * p.x + m[6] * p.y) + (m[10] * p.z + m[14] * w);
  return r;
}

// ---- Philox-4x32-10 counter RNG (replaces the per-thread LCG stream of LinearCongruenceGenerator.hpp:15-33,
//      seeded at IlluminationEstimation.cu:47 / CameraRendering.cu:42-44) ---------------------------
RTF_HD void mulhilo32(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
#if defined(__CUDA_ARCH__)
  lo = a * b;
  hi = __umulhi(a, b);
#else
  uint64_t p = (uint64_t)a * (uint64_t)b;
  lo = (uint32_t)p;
  hi = (uint32_t)(p >> 32);
#endif
}
RTF_HD uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; r++) {
    uint32_t hi0, lo0, hi1, lo1;
    mulhilo32(0xD2511F53u, c.x, hi0, lo0);
    mulhilo32(0xCD9E8D57u, c.z, hi1, lo1);
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
// 24-bit uniform in [0,1): same granularity as LinearCongruenceGenerator.hpp:32
RTF_HD float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

enum { STREAM_PATH = 0, STREAM_ALPHA = 1, STREAM_CAMERA = 2, STREAM_SSS = 3 };

RTF_HD float4 path_rng(uint32_t seed, uint32_t stream, uint64_t unit, uint32_t bounce, uint32_t extra) {
  uint4 o = philox4x32_10(make_uint4((uint32_t)unit, (uint32_t)(unit >> 32), bounce, extra), make_uint2(seed, stream));
  return make_float4(u01(o.x), u01(o.y), u01(o.z), u01(o.w));
}

// ---- the reference's per-thread generator (LinearCongruenceGenerator.hpp:6-35: 16 TEA rounds seed an LCG; 24 bits per
//      draw).  Used only by the option "rng" = RTF_RNG_REFERENCE_LCG of the AoS illumination entry point, which serves the
//      reference's own stream in the reference's own order so that the GPU can be compared with its programs directly.
RTF_HD uint32_t lcg_seed(uint32_t val0, uint32_t val1) {
  uint32_t v0 = val0, v1 = val1, s0 = 0u;
  for (int n = 0; n < 16; n++) {
    s0 += 0x9e3779b9u;
    v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
    v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
  }
  return v0;
}
RTF_HD float lcg_next(uint32_t& state) {
  state = 1664525u * state + 1013904223u;
  return (float)(state & 0x00FFFFFFu) / 16777216.0f;
}

// ---- portable sin/cos(2*pi*u): only + * fma, so host and device agree bit for bit.
//      Stands in for glm::sin/cos(2*pi*rnd) at RayTracerUtilities.cuh:127-131,139-143, CameraRendering.cu:75-79.
RTF_HD void sincos2pi(float u, float& s, float& c) {
  float x = u * 4.0f;
  float qf = floorf(x + 0.5f);
  float r = x - qf;
  int q = (int)qf & 3;
  float r2 = r * r;
  float ps = 0.00016044118092395365f;
  ps = fmaf(ps, r2, -0.004681753925979137f);
  ps = fmaf(ps, r2, 0.07969262450933456f);
  ps = fmaf(ps, r2, -0.6459640860557556f);
  ps = fmaf(ps, r2, 1.5707963705062866f);
  ps = ps * r;
  float pc = -2.520204179745633e-05f;
  pc = fmaf(pc, r2, 0.0009192602592520416f);
  pc = fmaf(pc, r2, -0.020863480865955353f);
  pc = fmaf(pc, r2, 0.25366950035095215f);
  pc = fmaf(pc, r2, -1.2337005138397217f);
  pc = fmaf(pc, r2, 1.0f);
  float ss = (q & 1) ? pc : ps;
  float cc = (q & 1) ? ps : pc;
  if (q == 2 || q == 3) ss = -ss;
  if (q == 1 || q == 2) cc = -cc;
  s = ss, c = cc;
}

// ---- RayTracerUtilities.cuh:89-145 -----------------------------------------------------------------
RTF_HD float3 reflect3(float3 i, float3 n) { return i - (2.0f * dot3(i, n)) * n; }
RTF_HD float3 to_world(float3 n, float cos_theta, float r_phi) {
  float sin_theta = sqrtf(fmaxf(0.0f, 1.0f - cos_theta * cos_theta));
  float s, c;
  sincos2pi(r_phi, s, c);
  float3 l = f3(c * sin_theta, s * sin_theta, cos_theta);
  float3 helper = f3(1.0f, 0.0f, 0.0f);
  if (fabsf(n.x) > 0.99f) helper = f3(0.0f, 0.0f, 1.0f);
  float3 t = normalize3(cross3(n, helper));
  float3 b = normalize3(cross3(n, t));
  return (t * l.x + b * l.y) + n * l.z;
}
RTF_HD float3 sample_hemisphere(float r0, float r1, float3 n) { return to_world(n, r0, r1); }
RTF_HD float3 sample_lobe(float r0, float r1, float3 n, float alpha) {
  float cos_theta = 1.0f - (r0 * (1.0f - alpha)) * (1.0f - alpha);
  return to_world(n, cos_theta, r1);
}

// ---- Lambert + GGX importance sampling (option "brdf" = RTF_BRDF_LAMBERT_GGX; BASELINE config 2) ----------------------
// Not in the reference (its only BRDF is the mirror lobe above, BSDF.cuh:6-13): diffuse = cosine-weighted Lambert with
// albedo*(1-metallic)*0.96, specular = GGX microfacet (Trowbridge-Reitz, visible-normal sampling, Schlick Fresnel
// with F0 = mix(0.04, albedo, metallic), separable Smith G), lobe chosen with probability p_spec = (1 + metallic)/2.
// Only + - * / sqrt and the portable sincos: bit-identical on host and device.  r = three uniforms (direction 2, lobe 1);
// v = direction towards the viewer (= -ray direction), n = shading normal facing v.  Returns the path weight
// f*cos/pdf (0 = absorbed) and the new direction in `l`.
RTF_HD float3 sample_lambert_ggx(float3 albedo, float metallic, float roughness, float3 n, float3 v, float r0, float r1,
                                 float r2, float3& l) {
  const float p_spec = 0.5f + 0.5f * clampf(metallic, 0.0f, 1.0f);
  const float nv = dot3(n, v);
  if (r2 >= p_spec) {  // diffuse lobe: cos(theta) = sqrt(1 - r0)
    l = to_world(n, sqrtf(1.0f - r0), r1);
    const float k = ((1.0f - metallic) * 0.96f) / (1.0f - p_spec);
    return albedo * k;
  }
  // specular lobe: GGX distribution of visible normals (Heitz 2018) — the weight F * G1(l) never exceeds 1
  if (!(nv > 0.0f)) return f3(0.0f, 0.0f, 0.0f);
  const float a = fmaxf(roughness * roughness, 1e-3f), a2 = a * a;
  float3 helper = f3(1.0f, 0.0f, 0.0f);
  if (fabsf(n.x) > 0.99f) helper = f3(0.0f, 0.0f, 1.0f);
  const float3 tx = normalize3(cross3(n, helper)), ty = normalize3(cross3(n, tx));
  const float vx = dot3(v, tx), vy = dot3(v, ty);
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
  sincos2pi(r1, sn, cs);
  const float p1 = rr * cs;
  float p2 = rr * sn;
  const float sw = 0.5f * (1.0f + hz);
  p2 = (1.0f - sw) * sqrtf(fmaxf(0.0f, 1.0f - p1 * p1)) + sw * p2;
  const float p3 = sqrtf(fmaxf(0.0f, (1.0f - p1 * p1) - p2 * p2));
  const float nhx = (p1 * t1x + p2 * t2x) + p3 * hx, nhy = (p1 * t1y + p2 * t2y) + p3 * hy, nhz = p2 * t2z + p3 * hz;
  // unstretch
  const float ex = a * nhx, ey = a * nhy, ez = fmaxf(0.0f, nhz);
  const float ie = 1.0f / sqrtf((ex * ex + ey * ey) + ez * ez);
  const float3 h = (tx * (ex * ie) + ty * (ey * ie)) + n * (ez * ie);
  const float vh = dot3(v, h);
  l = (2.0f * vh) * h - v;
  const float nl = dot3(n, l);
  if (!(nl > 0.0f) || !(vh > 0.0f)) return f3(0.0f, 0.0f, 0.0f);
  const float g1l = (2.0f * nl) / (nl + sqrtf(a2 + (1.0f - a2) * (nl * nl)));
  const float m = 1.0f - vh, m2 = m * m, m5 = (m2 * m2) * m;
  const float3 f0 = f3(0.04f, 0.04f, 0.04f) * (1.0f - metallic) + albedo * metallic;
  const float3 fr = f0 + (f3(1.0f, 1.0f, 1.0f) - f0) * m5;
  return fr * (g1l / p_spec);
}

}  // namespace rtf
