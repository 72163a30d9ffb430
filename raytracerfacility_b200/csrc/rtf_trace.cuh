// rtf_trace.cuh — closest-hit / any-hit traversal of the 8-wide compressed BVH (device only).
//
// Replaces optixTrace (10 call sites in the reference: IlluminationEstimation.cu:66,103;
// PointCloudScanning.cu:58; CameraRendering.cu:87; RayFunctions.cuh:73,108,136,201,232; BSSDF.cuh:63)
// — BVH traversal + built-in triangle intersection, closed source inside the OptiX driver.
//
// * Traversal order: Ylitie et al. HPG'17 Alg. 1 — children of a node are visited in descending
//   (slot ^ octant) priority, a node's leaf triangles are tested right after the node.  The per-ray
//   visit sequence is a pure function of (BVH, ray); no warp-level postponing, so the CPU oracle
//   reproduces the node/triangle visit counts exactly.
// * Ray/triangle: Woop/Benthin/Wald JCGT'13 watertight test, without back-face culling; the axis
//   permutation is folded into a per-ray 3x3 shear matrix (no per-triangle selects).  The closest hit
//   resolves equal t to the lower primitive id, so the result does not depend on the BVH at all.
// * Traversal stack: entries [0, SMEM_STACK) live in shared memory, laid out [entry][thread] so a warp's
//   accesses hit 32 consecutive 8-byte words; deeper entries spill to local memory.
#pragma once
#include "rtf_bvh.cuh"

namespace rtf {

struct HitRec {
  float t, u, v;
  uint32_t prim;  // 0xFFFFFFFF = miss
};

struct AlphaCtx {          // stochastic alpha test = AnyHitFunc, RayFunctions.cuh:5-23
  const float* inst_alpha; // per instance: 1 - transmission (Default material), else 1
  uint32_t seed;
  uint64_t unit;
  uint32_t bounce;
};

constexpr int kSmemStack = 12;   // entries per thread in shared memory
constexpr int kLocalStack = 52;  // spill entries in local memory (kSmemStack + kLocalStack = kMaxStackDepth)
constexpr float kBoxSlack = 1.00000095367431640625f;  // 1 + 2^-20: absorbs rounding of the slab test

RTF_D float safe_dir(float d) {
  const float ooeps = __uint_as_float(0x17800000u);  // 2^-80
  return fabsf(d) < ooeps ? copysignf(ooeps, d) : d;
}
RTF_D uint32_t extract_byte(uint32_t w, int j) { return (w >> (8 * j)) & 0xffu; }

enum { TRACE_ANY_HIT = 1u, TRACE_ALPHA = 2u };

template <int SMEM_STACK, bool STATS>
RTF_D HitRec traverse(const uint4* __restrict__ nodes, const float4* __restrict__ tris, const float3 o, const float3 d,
                      const float tmin, const float tmax, uint2* __restrict__ sstack, const int sstride,
                      const uint32_t flags, const AlphaCtx& actx, uint32_t& nvis, uint32_t& tvis) {
  HitRec best;
  best.t = tmax, best.u = 0.0f, best.v = 0.0f, best.prim = 0xFFFFFFFFu;

  // ---- per-ray constants -------------------------------------------------------------------
  const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
  const int kz = (ax >= ay && ax >= az) ? 0 : (ay >= az ? 1 : 2);
  // shear matrix rows (mx, my, mz): kx = kz+1, ky = kz+2 (mod 3)
  float3 mx, my, mz;
  {
    const float dz = kz == 0 ? d.x : (kz == 1 ? d.y : d.z);
    const float dx = kz == 0 ? d.y : (kz == 1 ? d.z : d.x);
    const float dy = kz == 0 ? d.z : (kz == 1 ? d.x : d.y);
    const float Sz = 1.0f / dz;
    const float Sx = dx * Sz, Sy = dy * Sz;
    if (kz == 0) {  // kx = y, ky = z
      mx = f3(-Sx, 1.0f, 0.0f), my = f3(-Sy, 0.0f, 1.0f), mz = f3(Sz, 0.0f, 0.0f);
    } else if (kz == 1) {  // kx = z, ky = x
      mx = f3(0.0f, -Sx, 1.0f), my = f3(1.0f, -Sy, 0.0f), mz = f3(0.0f, Sz, 0.0f);
    } else {  // kx = x, ky = y
      mx = f3(1.0f, 0.0f, -Sx), my = f3(0.0f, 1.0f, -Sy), mz = f3(0.0f, 0.0f, Sz);
    }
  }
  const float sdx = safe_dir(d.x), sdy = safe_dir(d.y), sdz = safe_dir(d.z);
  const float idx = 1.0f / sdx, idy = 1.0f / sdy, idz = 1.0f / sdz;
  const bool negx = sdx < 0.0f, negy = sdy < 0.0f, negz = sdz < 0.0f;
  const uint32_t oct_inv = (negx ? 0u : 4u) | (negy ? 0u : 2u) | (negz ? 0u : 1u);
  const uint32_t oct_inv4 = oct_inv * 0x01010101u;

  int sp = 0;
  uint2 lstack[kLocalStack];
  uint2 cur = make_uint2(0u, 0x80000000u);

  for (;;) {
    uint2 tg;
    {
      const uint32_t hits = cur.y;
      const uint32_t off = 31u - (uint32_t)__clz((int)hits);
      cur.y &= ~(1u << off);
      if (cur.y & 0xff000000u) {
        if (sp < SMEM_STACK)
          sstack[sp * sstride] = cur;
        else
          lstack[sp - SMEM_STACK] = cur;
        sp++;
      }
      const uint32_t slot = (off - 24u) ^ oct_inv;
      const uint32_t rel = (uint32_t)__popc(hits & ~(0xffffffffu << slot));
      const uint4* np = nodes + (size_t)(cur.x + rel) * 5;
      const uint4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
      if (STATS) nvis++;
      const float sx = __uint_as_float((n0.w & 0xffu) << 23) * idx;
      const float sy = __uint_as_float(((n0.w >> 8) & 0xffu) << 23) * idy;
      const float sz = __uint_as_float(((n0.w >> 16) & 0xffu) << 23) * idz;
      const float ox = (__uint_as_float(n0.x) - o.x) * idx;
      const float oy = (__uint_as_float(n0.y) - o.y) * idy;
      const float oz = (__uint_as_float(n0.z) - o.z) * idz;
      uint32_t hitmask = 0;
#pragma unroll
      for (int half = 0; half < 2; half++) {
        const uint32_t meta4 = half ? n1.w : n1.z;
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xffu;
        const uint32_t bit_index4 = (meta4 ^ (oct_inv4 & inner_mask4)) & 0x1f1f1f1fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t qlox = half ? n2.y : n2.x, qloy = half ? n2.w : n2.z, qloz = half ? n3.y : n3.x;
        const uint32_t qhix = half ? n3.w : n3.z, qhiy = half ? n4.y : n4.x, qhiz = half ? n4.w : n4.z;
        const uint32_t nx = negx ? qhix : qlox, fx = negx ? qlox : qhix;
        const uint32_t ny = negy ? qhiy : qloy, fy = negy ? qloy : qhiy;
        const uint32_t nz = negz ? qhiz : qloz, fz = negz ? qloz : qhiz;
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const float tnx = fmaf((float)extract_byte(nx, j), sx, ox);
          const float tny = fmaf((float)extract_byte(ny, j), sy, oy);
          const float tnz = fmaf((float)extract_byte(nz, j), sz, oz);
          const float tfx = fmaf((float)extract_byte(fx, j), sx, ox);
          const float tfy = fmaf((float)extract_byte(fy, j), sy, oy);
          const float tfz = fmaf((float)extract_byte(fz, j), sz, oz);
          const float tnear = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
          const float tfar = fminf(fminf(tfx, tfy), fminf(tfz, best.t));
          if (tnear <= tfar * kBoxSlack) {
            const uint32_t cb = extract_byte(child_bits4, j);
            const uint32_t bi = extract_byte(bit_index4, j);
            hitmask |= cb << bi;
          }
        }
      }
      cur.x = n1.x;
      cur.y = (hitmask & 0xff000000u) | (n0.w >> 24);
      tg.x = n1.y;
      tg.y = hitmask & 0x00ffffffu;
    }
    while (tg.y) {
      const uint32_t ti = 31u - (uint32_t)__clz((int)tg.y);
      tg.y &= ~(1u << ti);
      const float4* tp = tris + (size_t)(tg.x + ti) * 3;
      const float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
      if (STATS) tvis++;
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
      if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) continue;
      const float det = (U + V) + W;
      if (det == 0.0f) continue;
      const float Az = fmaf(mz.z, A.z, fmaf(mz.y, A.y, mz.x * A.x));
      const float Bz = fmaf(mz.z, B.z, fmaf(mz.y, B.y, mz.x * B.x));
      const float Cz = fmaf(mz.z, C.z, fmaf(mz.y, C.y, mz.x * C.x));
      const float T = fmaf(U, Az, fmaf(V, Bz, W * Cz));
      const float rcp = 1.0f / det;
      const float t = T * rcp;
      const uint32_t prim = __float_as_uint(t0.w);
      if (!(t >= tmin)) continue;
      if (t > best.t) continue;
      if (t == best.t && prim >= best.prim) continue;
      if (flags & TRACE_ALPHA) {
        if (__float_as_uint(t2.w) & TRI_FLAG_ALPHA) {
          const float alpha = actx.inst_alpha[__float_as_uint(t1.w)];
          const float rnd = path_rng(actx.seed, STREAM_ALPHA, actx.unit, actx.bounce, prim).x;
          if (alpha <= rnd) continue;
        }
      }
      best.t = t, best.u = V * rcp, best.v = W * rcp, best.prim = prim;
      if (flags & TRACE_ANY_HIT) return best;
    }
    if ((cur.y & 0xff000000u) == 0) {
      if (sp == 0) break;
      sp--;
      cur = sp < SMEM_STACK ? sstack[sp * sstride] : lstack[sp - SMEM_STACK];
    }
  }
  return best;
}

}  // namespace rtf
