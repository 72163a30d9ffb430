// oracle_math.h — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Scalar CPU arithmetic primitives of the oracle.  Nothing under raytracerfacility_b200/ may
// include this file.  Build with -ffp-contract=off so that every expression below is evaluated
// exactly as written (FMAs only where fmaf() is spelled out).
//
// Reference anchors (relative to /root/reference):
//   LCG            include/RayTracerFacility/RayTracer/LinearCongruenceGenerator.hpp:6-35
//   glm semantics  dot = (x+y)+z, normalize = v * (1/sqrt(dot)), cross as glm/detail/func_geometric.inl
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

struct V3 {
  float x, y, z;
};
static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
static inline V3 operator*(float s, V3 a) { return V3{a.x * s, a.y * s, a.z * s}; }
static inline V3 neg(V3 a) { return V3{-a.x, -a.y, -a.z}; }
static inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) {
  return V3{a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
static inline float length(V3 a) { return sqrtf(dot(a, a)); }
static inline V3 normalize(V3 a) {
  float inv = 1.0f / sqrtf(dot(a, a));
  return a * inv;
}
static inline V3 vmin(V3 a, V3 b) { return V3{fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)}; }
static inline V3 vmax(V3 a, V3 b) { return V3{fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)}; }
static inline float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

static inline uint32_t f2u(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
}
static inline float u2f(uint32_t u) {
  float f;
  memcpy(&f, &u, 4);
  return f;
}

// glm::mat4 * vec4 (column-major), evaluation order of glm's scalar path: (c0*x + c1*y) + (c2*z + c3*w).
static inline V3 mat4_mul_point(const float* m, V3 p, float w) {
  V3 r;
  r.x = (m[0] * p.x + m[4] * p.y) + (m[8] * p.z + m[12] * w);
  r.y = (m[1] * p.x + m[5] * p.y) + (m[9] * p.z + m[13] * w);
  r.z = (m[2] * p.x + m[6] * p.y) + (m[10] * p.z + m[14] * w);
  return r;
}

// ---- RNG of the reference (kept for known-answer tests; LCG.hpp:15-33) -------------------
struct LCG {
  uint32_t state;
  void init(uint32_t val0, uint32_t val1) {
    uint32_t v0 = val0, v1 = val1, s0 = 0;
    for (unsigned n = 0; n < 16; n++) {
      s0 += 0x9e3779b9u;
      v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
      v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    state = v0;
  }
  float next() {
    state = 1664525u * state + 1013904223u;
    return (float)(state & 0x00FFFFFFu) / (float)0x01000000;
  }
};

// ---- counter-based RNG: Philox-4x32-10 (Salmon et al., SC'11; Random123 v1.14 constants) ---
// The wavefront design needs a draw that depends only on (seed, unit, bounce, dimension), not on
// how many draws other rays made (SURVEY §7 "RNG-order dependence"); it replaces the per-thread
// LCG stream of IE.cu:47 / CR.cu:42-44.
static inline void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; r++) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0, c1 = n1, c2 = n2, c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0, out[1] = c1, out[2] = c2, out[3] = c3;
}
// 24-bit uniform in [0,1), same granularity as the reference's generator (LCG.hpp:32).
static inline float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

// RNG streams (key[1])
enum { STREAM_PATH = 0, STREAM_ALPHA = 1, STREAM_CAMERA = 2, STREAM_SSS = 3 };

static inline void path_rng(uint32_t seed, uint32_t stream, uint64_t unit, uint32_t bounce, uint32_t extra,
                            float r[4]) {
  uint32_t ctr[4] = {(uint32_t)unit, (uint32_t)(unit >> 32), bounce, extra};
  uint32_t key[2] = {seed, stream};
  uint32_t o[4];
  philox4x32_10(ctr, key, o);
  r[0] = u01(o[0]), r[1] = u01(o[1]), r[2] = u01(o[2]), r[3] = u01(o[3]);
}

// Sequential source, used by ONE test mode only (orc_illum_probes_lcg, pinning against the reference's own device
// programs run on the host, ref_ie_driver.cpp): while set for the calling thread, path_draws() serves the next `n`
// numbers of the reference's LCG in call order — the reference's per-thread stream (IE.cu:47) — instead of the keyed
// Philox draws.  The product never draws this way.
static thread_local LCG* t_seq_rng = nullptr;
static inline void path_draws(uint32_t seed, uint32_t stream, uint64_t unit, uint32_t bounce, uint32_t extra, float r[4],
                              int n) {
  if (t_seq_rng) {
    for (int k = 0; k < 4; k++) r[k] = k < n ? t_seq_rng->next() : 0.0f;
    return;
  }
  path_rng(seed, stream, unit, bounce, extra, r);
}

// ---- portable sin/cos of 2*pi*u, u in [0,1]: only + * fma, identical on host and device -----
// Stands in for glm::sin/glm::cos(2*pi*rnd) of RTU.cuh:127-131,139-143 and CR.cu:75-79; max abs
// error ~1.2e-7, so results agree with libm to the last place or two while being reproducible.
static inline void sincos2pi(float u, float* s, float* c) {
  float x = u * 4.0f;                 // exact
  float qf = floorf(x + 0.5f);        // nearest quadrant boundary
  float r = x - qf;                   // exact, |r| <= 0.5
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
  float ss, cc;
  switch (q) {
    case 0: ss = ps, cc = pc; break;
    case 1: ss = pc, cc = -ps; break;
    case 2: ss = -ps, cc = -pc; break;
    default: ss = -pc, cc = ps; break;
  }
  *s = ss, *c = cc;
}

// ---- portable inverse trigonometry: only + - * / sqrt, identical on host and device ----------------
// Stand in for acosf/asin/atan2 of the BTF angle code (RDD.hpp:307-313, SharedCoordinates.cuh:131-165).  atan on
// [0,1] after Cephes' atanf: reduce above tan(pi/8) with (x-1)/(x+1), then an odd degree-9 polynomial (abs error
// ~2e-7, i.e. the last place or two of libm, and reproducible).
static inline float atan01(float a) {  // a in [0,1]
  float off = 0.0f;
  if (a > 0.4142135679721832275390625f) {
    a = (a - 1.0f) / (a + 1.0f);
    off = 0.785398185253143310546875f;
  }
  const float z = a * a;
  float p = 8.05374449538e-2f;
  p = p * z - 1.38776856032e-1f;
  p = p * z + 1.99777106478e-1f;
  p = p * z - 3.33329491539e-1f;
  return off + (a + a * (z * p));
}
// Test switch (orc_set_libm_trig): the inverse functions and sincos_rad below call libm, as the reference's glm::asin /
// acosf / atan2 / glm::sin / glm::cos do, instead of the portable polynomials — used to show that the restatements that
// sit on top of them (compressed BTF) equal the reference's own code bit for bit once the trigonometry is the same.
// The product always uses the portable forms (identical on host and device).
static int g_libm_trig = 0;
static inline float atan2p(float y, float x) {
  if (g_libm_trig) return atan2f(y, x);
  if (y != y || x != x) return NAN;
  const float ax = fabsf(x), ay = fabsf(y);
  const float hi = ax > ay ? ax : ay, lo = ax > ay ? ay : ax;
  float r = hi > 0.0f ? atan01(lo / hi) : 0.0f;
  if (ay > ax) r = 1.57079637050628662109375f - r;
  if (x < 0.0f) r = 3.1415927410125732421875f - r;
  return y < 0.0f ? -r : r;
}
static inline float asinp(float v) { return g_libm_trig ? asinf(v) : atan2p(v, sqrtf((1.0f - v) * (1.0f + v))); }
static inline float acosp(float v) { return g_libm_trig ? acosf(v) : atan2p(sqrtf((1.0f - v) * (1.0f + v)), v); }
// sin / cos of an angle in radians through the 2*pi*u polynomial above
static inline void sincos_rad(float a, float* s, float* c) {
  if (g_libm_trig) {
    *s = sinf(a), *c = cosf(a);
    return;
  }
  sincos2pi(a * 0.15915493667125701904296875f, s, c);
}
static inline float radiansf(float d) { return d * 0.01745329238474369049072265625f; }
static inline float degreesf(float r) { return r * 57.295780181884765625f; }
static inline float modf_glm(float x, float y) { return x - y * floorf(x / y); }  // glm::mod

}  // namespace orc
