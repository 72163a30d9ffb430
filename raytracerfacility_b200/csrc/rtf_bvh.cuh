// rtf_bvh.cuh — data layout of the acceleration structure + the pure functions of its builder.
//
// This replaces the closed-source half of the reference path: optixAccelBuild
// (/root/reference/src/RayTracerFacility/RayTracer.cu:1528-1584 GAS, :1654-1707 IAS).  Algorithms:
// Karras HPG'12 (Morton + radix tree), Ylitie/Karras/Laine HPG'17 (80 B 8-wide compressed node).
//
// HBM layout (DESIGN.md "Data layout"):
//   wide node  = 80 B = 5 x 16 B, fetched as five 128-bit loads:
//       q0 = {p.x, p.y, p.z, e.x | e.y<<8 | e.z<<16 | imask<<24}
//       q1 = {child_base, tri_base, VALID (24 bits), imask | CNT << 8}   (node_meta_store below)
//       q2 = {qlo.x[0..3], qlo.x[4..7], qlo.y[0..3], qlo.y[4..7]}
//       q3 = {qlo.z[0..3], qlo.z[4..7], qhi.x[0..3], qhi.x[4..7]}
//       q4 = {qhi.y[0..3], qhi.y[4..7], qhi.z[0..3], qhi.z[4..7]}
//   triangle   = 48 B = 3 x 16 B: {v0.xyz, prim}, {v1.xyz, inst}, {v2.xyz, flags}
#pragma once
#include "rtf_math.cuh"

namespace rtf {

struct __align__(16) Node80 {
  float p[3];
  uint8_t e[3];
  uint8_t imask;
  uint32_t child_base;
  uint32_t tri_base;
  uint8_t meta[8];
  uint8_t qlo[3][8];
  uint8_t qhi[3][8];
};
static_assert(sizeof(Node80) == 80, "node must be 80 B");

struct __align__(16) Tri48 {
  float v0[3];
  uint32_t prim;
  float v1[3];
  uint32_t inst;
  float v2[3];
  uint32_t flags;
};
static_assert(sizeof(Tri48) == 48, "triangle must be 48 B");

// Node meta words: w0 bits 0..23 = VALID, bit 3s+j set iff slot s is a leaf slot holding more than j triangles;
// w1 = imask | CNT << 8, CNT = eight 2-bit fields, field s = triangles of leaf slot s (0 for inner / empty slots).
// A node's leaf triangles lie consecutively from tri_base in slot order: triangle j of slot s is
// tri_base + (sum of the CNT fields below s) + j.  The traversal keeps pending triangles as the mask "bit 3s+j" — a fixed
// position per slot, so a node visit needs no per-child variable shift (rtf_trace.cuh).
RTF_HD void node_meta_store(Node80& nd, uint32_t valid, uint32_t cnt16) {
  const uint32_t w0 = valid & 0xffffffu, w1 = (uint32_t)nd.imask | ((cnt16 & 0xffffu) << 8);
#pragma unroll
  for (int k = 0; k < 4; k++) nd.meta[k] = (uint8_t)(w0 >> (8 * k)), nd.meta[4 + k] = (uint8_t)(w1 >> (8 * k));
}

enum { TRI_FLAG_ALPHA = 1u };  // material may ignore hits (RayFunctions.cuh:5-23)

struct Box3 {
  float3 lo, hi;
};

RTF_HD uint32_t expand10(uint32_t v) {
  v &= 0x3ffu;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}
RTF_HD uint64_t expand21(uint64_t v) {
  v &= 0x1fffffull;
  v = (v | (v << 32)) & 0x1f00000000ffffull;
  v = (v | (v << 16)) & 0x1f0000ff0000ffull;
  v = (v | (v << 8)) & 0x100f00f00f00f00full;
  v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
  v = (v | (v << 2)) & 0x1249249249249249ull;
  return v;
}
RTF_HD uint32_t quantize_axis(float c, float lo, float inv, float scale, float maxq) {
  float nx = (c - lo) * inv;
  float q = fminf(fmaxf(nx * scale, 0.0f), maxq);
  return (uint32_t)q;
}
// 30-bit (10/axis) or 63-bit (21/axis) Morton code of a centroid; x owns the most significant bit.
RTF_HD uint64_t morton_code(float3 c, float3 lo, float3 inv, int bits) {
  if (bits == 30) {
    uint32_t x = quantize_axis(c.x, lo.x, inv.x, 1024.0f, 1023.0f);
    uint32_t y = quantize_axis(c.y, lo.y, inv.y, 1024.0f, 1023.0f);
    uint32_t z = quantize_axis(c.z, lo.z, inv.z, 1024.0f, 1023.0f);
    return (uint64_t)((expand10(x) << 2) | (expand10(y) << 1) | expand10(z));
  }
  uint64_t x = quantize_axis(c.x, lo.x, inv.x, 2097152.0f, 2097151.0f);
  uint64_t y = quantize_axis(c.y, lo.y, inv.y, 2097152.0f, 2097151.0f);
  uint64_t z = quantize_axis(c.z, lo.z, inv.z, 2097152.0f, 2097151.0f);
  return (expand21(x) << 2) | (expand21(y) << 1) | expand21(z);
}

RTF_HD float half_area(const Box3& b) {
  float3 d = b.hi - b.lo;
  return (d.x * d.y + d.y * d.z) + d.z * d.x;
}

// Biased exponent e with 2^(e-127) * 255 >= extent (integer bit operations only, no log2f).
RTF_HD uint32_t quant_exponent(float extent) {
  float f = extent / 255.0f;
  uint32_t bits = f2u(f);
  uint32_t e = (bits >> 23) & 0xffu;
  if (bits & 0x7fffffu) e += 1;
  if (e < 1) e = 1;
  if (u2f(e << 23) * 255.0f < extent) e += 1;
  if (e > 254) e = 254;
  return e;
}

// Conservative 8-bit quantisation of [lo,hi] against origin p and scale 2^(e-127).
RTF_HD void quantize_interval(float lo, float hi, float p, uint32_t e, uint8_t& qlo, uint8_t& qhi) {
  float inv_scale = u2f((254u - e) << 23);
  float scale = u2f(e << 23);
  float ql = floorf((lo - p) * inv_scale);
  float qh = ceilf((hi - p) * inv_scale);
  ql = fminf(fmaxf(ql, 0.0f), 255.0f);
  qh = fminf(fmaxf(qh, 0.0f), 255.0f);
  while (ql > 0.0f && p + ql * scale > lo) ql -= 1.0f;
  while (qh < 255.0f && p + qh * scale < hi) qh += 1.0f;
  qlo = (uint8_t)ql;
  qhi = (uint8_t)qh;
}

// ---- triangle references (pre-splitting) -----------------------------------------------------------------------
// Sorghum leaves are 2 mm x 45 mm slivers that run diagonally: a triangle's box is tens of times larger than the
// triangle, and a ray tests dozens of triangles it cannot hit (30 per illumination ray on the C2 field).  Before the
// Morton sort such a triangle is cut into k slabs along the longest axis of its box (early split clipping, Ernst &
// Greiner 2007); each slab becomes one REFERENCE whose box bounds the clipped triangle.  All references of a triangle
// carry the same 48 B triangle record, so hits do not change (equal t resolves to the lower primitive id).
constexpr float kSplitGain = 0.5f;  // k = floor(gain * box half-area / (2 * triangle area)), >= 2 to split
struct SplitPlan {
  int k, axis;
};
RTF_HD SplitPlan split_plan(float3 a, float3 b, float3 c, int max_split) {
  SplitPlan sp;
  sp.k = 1;
  const float3 lo = min3(min3(a, b), c), hi = max3(max3(a, b), c);
  const float3 e = hi - lo;
  sp.axis = (e.x >= e.y && e.x >= e.z) ? 0 : (e.y >= e.z ? 1 : 2);
  if (max_split <= 1) return sp;
  const float3 ab = b - a, ac = c - a;
  const float3 n = f3(ab.y * ac.z - ab.z * ac.y, ab.z * ac.x - ab.x * ac.z, ab.x * ac.y - ab.y * ac.x);
  const float area2 = sqrtf((n.x * n.x + n.y * n.y) + n.z * n.z);
  const float ha = (e.x * e.y + e.y * e.z) + e.z * e.x;
  if (!(area2 > 0.0f) || !(ha < INFINITY)) return sp;
  const float r = (ha / area2) * kSplitGain;
  if (!(r >= 2.0f)) return sp;
  sp.k = r >= (float)max_split ? max_split : (int)r;
  return sp;
}
// Box of triangle (a,b,c) clipped to slab j of k along `axis`.  Plane j = lo + j*(ext/k): the same float closes slab
// j-1 and opens slab j.  Clipped edge points are rounded, so the other two axes get 2^-21 of padding, clamped to the
// triangle's own box.
RTF_HD Box3 split_slab_box(float3 a, float3 b, float3 c, int axis, int k, int j) {
  const float A[3][3] = {{a.x, a.y, a.z}, {b.x, b.y, b.z}, {c.x, c.y, c.z}};
  float lo[3], hi[3];
#pragma unroll
  for (int d = 0; d < 3; d++) {
    lo[d] = fminf(fminf(A[0][d], A[1][d]), A[2][d]);
    hi[d] = fmaxf(fmaxf(A[0][d], A[1][d]), A[2][d]);
  }
  Box3 out;
  if (k <= 1) {
    out.lo = f3(lo[0], lo[1], lo[2]), out.hi = f3(hi[0], hi[1], hi[2]);
    return out;
  }
  const float lo_a = axis == 0 ? lo[0] : (axis == 1 ? lo[1] : lo[2]);
  const float hi_a = axis == 0 ? hi[0] : (axis == 1 ? hi[1] : hi[2]);
  const float w = (hi_a - lo_a) / (float)k;
  const float x0 = j == 0 ? lo_a : fminf(lo_a + (float)j * w, hi_a);
  const float x1 = j == k - 1 ? hi_a : fminf(lo_a + (float)(j + 1) * w, hi_a);
  float blo[3] = {INFINITY, INFINITY, INFINITY}, bhi[3] = {-INFINITY, -INFINITY, -INFINITY};
#pragma unroll
  for (int v = 0; v < 3; v++) {
    const float va = axis == 0 ? A[v][0] : (axis == 1 ? A[v][1] : A[v][2]);
    if (va >= x0 && va <= x1) {
#pragma unroll
      for (int d = 0; d < 3; d++) blo[d] = fminf(blo[d], A[v][d]), bhi[d] = fmaxf(bhi[d], A[v][d]);
    }
  }
#pragma unroll
  for (int ed = 0; ed < 3; ed++) {
    float P[3], Q[3];
#pragma unroll
    for (int d = 0; d < 3; d++) P[d] = A[ed][d], Q[d] = A[(ed + 1) % 3][d];
    float pa = axis == 0 ? P[0] : (axis == 1 ? P[1] : P[2]);
    float qa = axis == 0 ? Q[0] : (axis == 1 ? Q[1] : Q[2]);
    if (pa > qa) {
#pragma unroll
      for (int d = 0; d < 3; d++) {
        const float t = P[d];
        P[d] = Q[d], Q[d] = t;
      }
      const float t = pa;
      pa = qa, qa = t;
    }
#pragma unroll
    for (int s = 0; s < 2; s++) {
      const float xp = s ? x1 : x0;
      if (pa < xp && xp < qa) {
        const float t = (xp - pa) / (qa - pa);
#pragma unroll
        for (int d = 0; d < 3; d++) {
          const float v = fmaf(Q[d] - P[d], t, P[d]);
          blo[d] = fminf(blo[d], v), bhi[d] = fmaxf(bhi[d], v);
        }
      }
    }
  }
#pragma unroll
  for (int d = 0; d < 3; d++) {
    if (d == axis) {
      blo[d] = x0, bhi[d] = x1;
    } else {
      const float pad = fmaxf(fabsf(lo[d]), fabsf(hi[d])) * 4.76837158203125e-07f;  // 2^-21
      blo[d] = fmaxf(blo[d] - pad, lo[d]);
      bhi[d] = fminf(bhi[d] + pad, hi[d]);
      if (!(blo[d] <= bhi[d])) blo[d] = lo[d], bhi[d] = hi[d];
    }
  }
  out.lo = f3(blo[0], blo[1], blo[2]), out.hi = f3(bhi[0], bhi[1], bhi[2]);
  return out;
}

}  // namespace rtf
