// rtf_sss.cuh — the BSSRDF spatial probe of the reference (BSSDF.cuh:18-89, call site RayFunctions.cuh:97-133).
//
// At the first hit of a path on a material with subsurfaceFactor > 0 and subsurfaceRadius.x > 0 the reference shoots a
// "SpacialSampling" ray from a disk above the surface back into it and lets its any-hit program collect up to four
// intersections WITH THE SAME OBJECT (SBT handle) into SSPerRayData::m_records; one record is drawn at random and, if it
// lies within the radius of the entry point, a second radiance path continues from there.
//
// Which intersections the reference collects depends on the order in which the OptiX driver happens to report any-hit
// candidates (an accepted candidate shrinks the ray interval, BSSDF.cuh:18-37), so it is not a function of the scene.
// The order-independent statement used here (and restated by the CPU checker): the records are the (up to) four
// NEAREST same-handle intersections in [tmin, tmax], ascending (t, primitive id).  Pre-split references of one triangle
// give the same (t, prim) and are counted once.
//
// The probe is a short ray (length = radius), so it runs inside the shading kernel as a plain per-thread stack traversal
// of the same 8-wide compressed BVH — same conservative slab arithmetic as rtf_trace.cuh — instead of a queue of its own.
#pragma once
#include "rtf_shade.cuh"
#include "rtf_trace.cuh"

namespace rtf {

struct SSRecords {
  int n;
  float t[4], u[4], v[4];
  uint32_t prim[4];
};

// (t, prim) of a candidate is inserted if it is below the current 4th entry; duplicates (same triangle through another
// pre-split reference) are dropped.
RTF_D void ss_insert(SSRecords& r, float t, float u, float v, uint32_t prim) {
#pragma unroll
  for (int k = 0; k < 4; k++)
    if (k < r.n && r.prim[k] == prim) return;
  int pos = r.n;
#pragma unroll
  for (int k = 3; k >= 0; k--)
    if (k < r.n && (t < r.t[k] || (t == r.t[k] && prim < r.prim[k]))) pos = k;
  if (pos >= 4) return;
#pragma unroll
  for (int k = 3; k >= 1; k--)
    if (k > pos && k <= r.n) r.t[k] = r.t[k - 1], r.u[k] = r.u[k - 1], r.v[k] = r.v[k - 1], r.prim[k] = r.prim[k - 1];
#pragma unroll
  for (int k = 0; k < 4; k++)
    if (k == pos) r.t[k] = t, r.u[k] = u, r.v[k] = v, r.prim[k] = prim;
  if (r.n < 4) r.n++;
}

// Watertight ray/triangle test of rtf_trace.cuh (Woop/Benthin/Wald), as a function: true + (t,u,v) when the ray's line
// crosses the triangle; the interval test is the caller's.
RTF_D bool ss_tri_test(float3 o, float3 mx, float3 my, float3 mz, float4 t0, float4 t1, float4 t2, float& t, float& bu,
                       float& bv) {
  const float3 A = f3(t0.x - o.x, t0.y - o.y, t0.z - o.z);
  const float3 B = f3(t1.x - o.x, t1.y - o.y, t1.z - o.z);
  const float3 C = f3(t2.x - o.x, t2.y - o.y, t2.z - o.z);
  const float Ax = fmaf(mx.z, A.z, fmaf(mx.y, A.y, mx.x * A.x));
  const float Ay = fmaf(my.z, A.z, fmaf(my.y, A.y, my.x * A.x));
  const float Bx = fmaf(mx.z, B.z, fmaf(mx.y, B.y, mx.x * B.x));
  const float By = fmaf(my.z, B.z, fmaf(my.y, B.y, my.x * B.x));
  const float Cx = fmaf(mx.z, C.z, fmaf(mx.y, C.y, mx.x * C.x));
  const float Cy = fmaf(my.z, C.z, fmaf(my.y, C.y, my.x * C.x));
  float U = Cx * By - Cy * Bx;
  float V = Ax * Cy - Ay * Cx;
  float W = Bx * Ay - By * Ax;
  if (U == 0.0f || V == 0.0f || W == 0.0f) {
    U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
    V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
    W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
  }
  if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
  const float det = (U + V) + W;
  if (det == 0.0f) return false;
  const float Az = fmaf(mz.z, A.z, fmaf(mz.y, A.y, mz.x * A.x));
  const float Bz = fmaf(mz.z, B.z, fmaf(mz.y, B.y, mz.x * B.x));
  const float Cz = fmaf(mz.z, C.z, fmaf(mz.y, C.y, mz.x * C.x));
  const float T = fmaf(U, Az, fmaf(V, Bz, W * Cz));
  const float rcp = 1.0f / det;
  t = T * rcp, bu = V * rcp, bv = W * rcp;
  return true;
}

// The up-to-four nearest intersections of (o, d, [tmin, tmax]) with triangles whose instance carries `handle`.
RTF_D void ss_probe(const DevScene& sc, float3 o, float3 d, float tmin, float tmax, uint64_t handle, SSRecords& rec) {
  rec.n = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) rec.t[k] = 0.0f, rec.u[k] = 0.0f, rec.v[k] = 0.0f, rec.prim[k] = 0xFFFFFFFFu;
  // shear matrix (axis permutation folded in), as the pool refill of rtf_trace.cuh builds it
  const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
  const int kz = (ax >= ay && ax >= az) ? 0 : (ay >= az ? 1 : 2);
  const float dz = kz == 0 ? d.x : (kz == 1 ? d.y : d.z);
  const float dx = kz == 0 ? d.y : (kz == 1 ? d.z : d.x);
  const float dy = kz == 0 ? d.z : (kz == 1 ? d.x : d.y);
  const float Sz = 1.0f / dz;
  const float Sx = dx * Sz, Sy = dy * Sz;
  float3 mx, my, mz;
  if (kz == 0) {
    mx = f3(-Sx, 1.0f, 0.0f), my = f3(-Sy, 0.0f, 1.0f), mz = f3(Sz, 0.0f, 0.0f);
  } else if (kz == 1) {
    mx = f3(0.0f, -Sx, 1.0f), my = f3(1.0f, -Sy, 0.0f), mz = f3(0.0f, Sz, 0.0f);
  } else {
    mx = f3(1.0f, 0.0f, -Sx), my = f3(0.0f, 1.0f, -Sy), mz = f3(0.0f, 0.0f, Sz);
  }
  const float idx = 1.0f / safe_dir(d.x), idy = 1.0f / safe_dir(d.y), idz = 1.0f / safe_dir(d.z);
  const bool negx = idx < 0.0f, negy = idy < 0.0f, negz = idz < 0.0f;
  uint32_t stack[kMaxStackDepth * 7 + 8];  // every popped node pushes at most 8 children
  int sp = 0;
  stack[sp++] = 0u;
  while (sp > 0) {
    const uint32_t ni = stack[--sp];
    const uint4* np = sc.nodes + (size_t)ni * 5;
    const uint4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
    const float sx = (__uint_as_float((n0.w & 0xffu) << 23) * idx) * 32768.0f;
    const float sy = (__uint_as_float(((n0.w >> 8) & 0xffu) << 23) * idy) * 32768.0f;
    const float sz = (__uint_as_float(((n0.w >> 16) & 0xffu) << 23) * idz) * 32768.0f;
    const float ox0 = (__uint_as_float(n0.x) - o.x) * idx;
    const float oy0 = (__uint_as_float(n0.y) - o.y) * idy;
    const float oz0 = (__uint_as_float(n0.z) - o.z) * idz;
    const float mgx = fabsf(ox0) + fabsf(sx), mgy = fabsf(oy0) + fabsf(sy), mgz = fabsf(oz0) + fabsf(sz);
    const float ox = fmaf(mgx, -kFarSlack, ox0) - sx, oxf = fmaf(mgx, kFarSlack, ox0) - sx;
    const float oy = fmaf(mgy, -kFarSlack, oy0) - sy, oyf = fmaf(mgy, kFarSlack, oy0) - sy;
    const float oz = fmaf(mgz, -kFarSlack, oz0) - sz, ozf = fmaf(mgz, kFarSlack, oz0) - sz;
    const float bound = (rec.n == 4 ? rec.t[3] : tmax) * kBoxSlack;
    const uint32_t imask = n0.w >> 24;
    uint32_t first = n1.y;  // first triangle of the next leaf slot (node meta: rtf_bvh.cuh)
    for (int j = 0; j < 8; j++) {
      const bool inner = (imask >> j) & 1u;
      const uint32_t cnt = (n1.w >> (8 + 2 * j)) & 3u;  // triangles of leaf slot j
      const uint32_t slot_first = first;
      first += cnt;
      if (!inner && cnt == 0u) continue;
      const int sh = 8 * (j & 3);
      const uint32_t qlx = ((j < 4 ? n2.x : n2.y) >> sh) & 0xffu, qly = ((j < 4 ? n2.z : n2.w) >> sh) & 0xffu;
      const uint32_t qlz = ((j < 4 ? n3.x : n3.y) >> sh) & 0xffu, qhx = ((j < 4 ? n3.z : n3.w) >> sh) & 0xffu;
      const uint32_t qhy = ((j < 4 ? n4.x : n4.y) >> sh) & 0xffu, qhz = ((j < 4 ? n4.z : n4.w) >> sh) & 0xffu;
      const float tnx = fmaf(__uint_as_float(0x3F800000u | ((negx ? qhx : qlx) << 8)), sx, ox);
      const float tny = fmaf(__uint_as_float(0x3F800000u | ((negy ? qhy : qly) << 8)), sy, oy);
      const float tnz = fmaf(__uint_as_float(0x3F800000u | ((negz ? qhz : qlz) << 8)), sz, oz);
      const float tfx = fmaf(__uint_as_float(0x3F800000u | ((negx ? qlx : qhx) << 8)), sx, oxf);
      const float tfy = fmaf(__uint_as_float(0x3F800000u | ((negy ? qly : qhy) << 8)), sy, oyf);
      const float tfz = fmaf(__uint_as_float(0x3F800000u | ((negz ? qlz : qhz) << 8)), sz, ozf);
      const float tnear = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
      const float tfar = fminf(fminf(tfx, tfy), fminf(tfz, bound));
      if (!(tnear <= tfar)) continue;
      if (inner) {
        stack[sp++] = n1.x + (uint32_t)__popc(imask & ((1u << j) - 1u));
      } else {
        for (uint32_t b = 0; b < cnt; b++) {
          const float4* tp = sc.tris + (size_t)(slot_first + b) * 3;
          const float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
          float t, bu, bv;
          if (!ss_tri_test(o, mx, my, mz, t0, t1, t2, t, bu, bv)) continue;
          if (!(t >= tmin) || t > tmax) continue;
          if (sc.insts[__float_as_uint(t1.w)].handle != handle) continue;  // SSAnyHit, BSSDF.cuh:22-24
          ss_insert(rec, t, bu, bv, __float_as_uint(t0.w));
        }
      }
    }
  }
}

// glm::rotate(v, angle, normal) of <glm/gtx/rotate_vector.hpp> = mat3(glm::rotate(angle, normal)) * v
// (glm/ext/matrix_transform.inl), angle = 2*pi*u through the portable sincos2pi.
RTF_HD float3 rotate_about(float3 v, float u_turns, float3 axis_in) {
  float s, c;
  sincos2pi(u_turns, s, c);
  const float3 axis = normalize3(axis_in);
  const float3 temp = (1.0f - c) * axis;
  const float3 c0 = f3(c + temp.x * axis.x, temp.x * axis.y + s * axis.z, temp.x * axis.z - s * axis.y);
  const float3 c1 = f3(temp.y * axis.x - s * axis.z, c + temp.y * axis.y, temp.y * axis.z + s * axis.x);
  const float3 c2 = f3(temp.z * axis.x + s * axis.y, temp.z * axis.y - s * axis.x, c + temp.z * axis.z);
  return f3((c0.x * v.x + c1.x * v.y) + c2.x * v.z, (c0.y * v.x + c1.y * v.y) + c2.y * v.z,
            (c0.z * v.x + c1.z * v.y) + c2.z * v.z);
}

}  // namespace rtf
