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
// * Execution model ("persistent lanes"): a warp owns 32 traversal slots.  A lane whose ray has
//   finished takes the next ray from the warp's pool in shared memory in the SAME loop iteration, so
//   the node/triangle code always runs with (almost) all 32 lanes busy — incoherent rays have a
//   heavy-tailed visit count (mean 11, max > 100 nodes) and a warp that waited for its slowest lane
//   ran at 11 % SIMT utilisation (profiles/r1_trace_baseline.md).
// * Ray pool: rays are claimed in tiles of 32 from a global cursor.  Lane 0 brings the NEXT tile into
//   shared memory with a TMA bulk copy (cp.async.bulk + mbarrier complete_tx) while the current one is
//   traversed; when the pool runs dry the whole warp — converged, 32 lanes — turns the raw tile into
//   traversal-ready records (reciprocal direction, shear constants) and goes back to traversing.
// * Traversal stack: entries [0, kSmemStack) live in shared memory, laid out [entry][thread] so a warp's
//   accesses hit 32 consecutive 8-byte words; deeper entries spill to local memory.
#pragma once
#include "rtf_bvh.cuh"

namespace rtf {

struct HitRec {
  float t, u, v;
  uint32_t prim;  // 0xFFFFFFFF = miss
};

#ifndef RTF_SMEM_STACK
#define RTF_SMEM_STACK 8
#endif
constexpr int kSmemStack = RTF_SMEM_STACK;        // entries per thread in shared memory
constexpr int kLocalStack = 64 - RTF_SMEM_STACK;  // spill entries in local memory (kSmemStack + kLocalStack = kMaxStackDepth)
constexpr float kBoxSlack = 1.00000095367431640625f;  // 1 + 2^-20 on the current best t
constexpr float kFarSlack = 2.384185791015625e-07f;   // 2^-22 of the slab operands: near planes pulled in, far pushed out
constexpr int kTraceBlock = 128;
constexpr int kTraceWarps = kTraceBlock / 32;
constexpr int kRawTileBytes = 1024;  // 32 rays x 32 B
#ifndef RTF_NODE_QUORUM
#define RTF_NODE_QUORUM 24
#endif
constexpr int kNodeQuorum = RTF_NODE_QUORUM;  // lanes that must be ready before the warp runs a node visit
#ifndef RTF_REFILL_MIN
#define RTF_REFILL_MIN 1
#endif
constexpr int kRefillMin = RTF_REFILL_MIN;  // idle lanes that must have gathered before the warp goes to the pool
#ifndef RTF_FFMA2
#define RTF_FFMA2 0  // 1: plane distances with packed fma.rn.f32x2 (Blackwell FFMA2); needs more than 64 registers to pay
#endif

RTF_D float safe_dir(float d) {
  const float ooeps = __uint_as_float(0x17800000u);  // 2^-80
  return fabsf(d) < ooeps ? copysignf(ooeps, d) : d;
}
RTF_D uint32_t extract_byte(uint32_t w, int j) { return (w >> (8 * j)) & 0xffu; }
// Byte j of w as the float 1 + b*2^-15 (bits 0x3F80bb00), exact, ONE PRMT on the full-rate ALU pipe.  The plane
// distance q*s + o is evaluated as (1 + q*2^-15) * (s*2^15) + (o - s*2^15): the bias is folded into the per-node
// constants, so a child costs 6 PRMT + 6 FFMA.  (History: I2F.U8 is a quarter-rate XU instruction — 48 per node made
// XU the busiest pipe of the first version; PRMT to 2^23 + b needed an FADD per byte to remove the bias.)
RTF_D float byte_to_float(uint32_t w, int j, uint32_t magic /* 0x3F800000, kept in a register */) {
  return __uint_as_float(__byte_perm(w, magic, 0x7604u + ((uint32_t)j << 4)));
}
// The bias constant must stay in ONE register so that the selector becomes PRMT's immediate; given a literal, ptxas
// encodes the constant as the immediate and re-materialises the selector with a MOV per PRMT (48 per node).  It is
// therefore derived from a run-time value the compiler cannot fold (n < 2^31 always).
RTF_D uint32_t prmt_magic(uint32_t n) { return 0x3F800000u | (n >> 31); }

// ---- TMA 1-D bulk copy + mbarrier (PTX; SASS: UBLKCP / SYNCS) -------------------------------------------
RTF_D uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
RTF_D void mbar_init(uint32_t mbar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
RTF_D void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
RTF_D void tma_load_1d(uint32_t dst_smem, const void* src_gmem, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
               "l"(src_gmem), "r"(bytes), "r"(mbar)
               : "memory");
}
RTF_D void mbar_wait(uint32_t mbar, uint32_t parity) {
  uint32_t ok = 0;
  do {
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(mbar), "r"(parity)
        : "memory");
  } while (!ok);
}

// Shared memory of one trace block (static, 23 KB): per warp a raw TMA tile, a pool of prepared rays and
// an mbarrier; per thread kSmemStack stack entries.
struct __align__(128) TraceSmem {
  unsigned char raw[kTraceWarps][kRawTileBytes];
  float4 pool[kTraceWarps][5][32];
  uint2 stack[kSmemStack * kTraceBlock];
  unsigned long long mbar[kTraceWarps];
};

// A ray source/sink ("Src") provides:
//   static constexpr bool kTma     — tiles arrive by TMA bulk copy (issue()) or are read/generated in fetch()
//   static constexpr bool kAnyHit  — rays may be first-hit rays (any_hit(tag))
//   static constexpr bool kAlpha   — stochastic alpha test may reject candidates (alpha_pass(tag, prim, inst, u, v));
//                                    all_candidates(): call it for every candidate, not only flagged triangles
//   uint32_t n                     — rays in this launch
//   void issue(base, cnt, raw_smem_u32, mbar_u32)            (lane 0; kTma only)
//   void fetch(raw, base, slot, a, b, tmin, tag)              a = {o.xyz, *}, b = {d.xyz, tmax}
//   void store(index, tag, hit, node_visits, tri_visits)
template <bool STATS, class Src>
RTF_D void trace_engine(const uint4* __restrict__ nodes, const float4* __restrict__ tris, const Src& src,
                        uint32_t* __restrict__ work, TraceSmem& sm) {
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const uint32_t n = src.n;
  const unsigned char* raw = sm.raw[warp];
  float4(*pool)[32] = sm.pool[warp];
  uint2* sstack = sm.stack + threadIdx.x;
  const uint32_t raw_s = smem_u32(raw), mbar_s = smem_u32(&sm.mbar[warp]);
  uint32_t parity = 0;
  const uint32_t magic = prmt_magic(src.n);

  if constexpr (Src::kTma) {
    if (lane == 0) mbar_init(mbar_s, 1);
    __syncwarp();
  }
  uint32_t next_base = 0;  // tile in flight (>= n: none left)
  if (lane == 0) next_base = atomicAdd(work, 32u);
  next_base = __shfl_sync(0xffffffffu, next_base, 0);
  if constexpr (Src::kTma) {
    if (lane == 0 && next_base < n) src.issue(next_base, min(32u, n - next_base), raw_s, mbar_s);
  }
  uint32_t pool_base = 0, pool_pos = 0, pool_cnt = 0;

  // ---- per-lane traversal slot ----
  bool have = false;
  float3 o = f3(0, 0, 0), mx = f3(0, 0, 0), my = f3(0, 0, 0), mz = f3(0, 0, 0);
  float tmin = 0, idx = 0, idy = 0, idz = 0;
  uint32_t oct_inv = 0, pmask = 0, tag = 0, ridx = 0, nvis = 0, tvis = 0;
  HitRec best;
  best.t = 0, best.u = 0, best.v = 0, best.prim = 0xFFFFFFFFu;
  int sp = 0;
  uint2 lstack[kLocalStack];
  uint2 cur = make_uint2(0u, 0u);
  uint2 tg = make_uint2(0u, 0u);  // pending leaf triangles of the node visited last: {base, hit bits}

  for (;;) {
    unsigned idle = __ballot_sync(0xffffffffu, !have);
    if (kRefillMin > 1 ? (__popc(idle) >= kRefillMin) : (idle != 0u)) {
      for (;;) {
        if (pool_pos >= pool_cnt) {
          if (next_base >= n) break;
          // ---- whole warp, converged: turn tile `next_base` into traversal-ready records ----
          const uint32_t base = next_base, cnt = min(32u, n - base);
          if constexpr (Src::kTma) {
            mbar_wait(mbar_s, parity);
            parity ^= 1u;
          }
          float4 a = make_float4(0, 0, 0, 0), b = make_float4(0, 0, 1, 0);
          float r_tmin = 0;
          uint32_t r_tag = 0;
          if ((uint32_t)lane < cnt) src.fetch(raw, base, (uint32_t)lane, a, b, r_tmin, r_tag);
          __syncwarp();  // raw tile and old pool records are dead from here on
          if (lane == 0) next_base = atomicAdd(work, 32u);
          next_base = __shfl_sync(0xffffffffu, next_base, 0);
          if constexpr (Src::kTma) {
    if (lane == 0 && next_base < n) src.issue(next_base, min(32u, n - next_base), raw_s, mbar_s);
  }
          if ((uint32_t)lane < cnt) {
            const float3 d = f3(b.x, b.y, b.z);
            const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
            const int kz = (ax >= ay && ax >= az) ? 0 : (ay >= az ? 1 : 2);
            const float dz = kz == 0 ? d.x : (kz == 1 ? d.y : d.z);
            const float dx = kz == 0 ? d.y : (kz == 1 ? d.z : d.x);
            const float dy = kz == 0 ? d.z : (kz == 1 ? d.x : d.y);
            const float Sz = 1.0f / dz;
            const float Sx = dx * Sz, Sy = dy * Sz;
            const float sdx = safe_dir(d.x), sdy = safe_dir(d.y), sdz = safe_dir(d.z);
            // shear matrix rows (mx, my, mz): kx = kz+1, ky = kz+2 (mod 3) — built here, at 32/32 lanes
            float3 rx, ry, rz;
            if (kz == 0) {  // kx = y, ky = z
              rx = f3(-Sx, 1.0f, 0.0f), ry = f3(-Sy, 0.0f, 1.0f), rz = f3(Sz, 0.0f, 0.0f);
            } else if (kz == 1) {  // kx = z, ky = x
              rx = f3(0.0f, -Sx, 1.0f), ry = f3(1.0f, -Sy, 0.0f), rz = f3(0.0f, Sz, 0.0f);
            } else {  // kx = x, ky = y
              rx = f3(1.0f, 0.0f, -Sx), ry = f3(0.0f, 1.0f, -Sy), rz = f3(0.0f, 0.0f, Sz);
            }
            pool[0][lane] = make_float4(a.x, a.y, a.z, r_tmin);
            pool[1][lane] = make_float4(1.0f / sdx, 1.0f / sdy, 1.0f / sdz, b.w);
            pool[2][lane] = make_float4(rx.x, rx.y, rx.z, ry.x);
            pool[3][lane] = make_float4(ry.y, ry.z, rz.x, rz.y);
            const bool ngx = sdx < 0.0f, ngy = sdy < 0.0f, ngz = sdz < 0.0f;  // sign of 1/sd = sign of sd
            const uint32_t r_oct = (ngx ? 0u : 4u) | (ngy ? 0u : 2u) | (ngz ? 0u : 1u);
            const uint32_t r_pm = (ngz ? 0u : 0x55u) | (ngy ? 0u : 0x33u << 10) | (ngx ? 0u : 0x0Fu << 20);
            pool[4][lane] = make_float4(rz.z, __uint_as_float(r_tag), __uint_as_float(r_oct), __uint_as_float(r_pm));
          }
          __syncwarp();
          pool_base = base, pool_pos = 0, pool_cnt = cnt;
        }
        const uint32_t slot = pool_pos + (uint32_t)__popc(idle & ((1u << lane) - 1u));
        if (!have && slot < pool_cnt) {
          const float4 q0 = pool[0][slot], q1 = pool[1][slot], q2 = pool[2][slot], q3 = pool[3][slot], q4 = pool[4][slot];
          tag = __float_as_uint(q4.y);
          ridx = pool_base + slot;
          o = f3(q0.x, q0.y, q0.z), tmin = q0.w;
          idx = q1.x, idy = q1.y, idz = q1.z;
          best.t = q1.w, best.u = 0.0f, best.v = 0.0f, best.prim = 0xFFFFFFFFu;
          mx = f3(q2.x, q2.y, q2.z), my = f3(q2.w, q3.x, q3.y), mz = f3(q3.z, q3.w, q4.x);
          oct_inv = __float_as_uint(q4.z), pmask = __float_as_uint(q4.w);  // octant words, built at 32 lanes above
          cur = make_uint2(0u, 0x80000000u);
          tg = make_uint2(0u, 0u);
          sp = 0, nvis = 0, tvis = 0;
          have = true;
        }
        pool_pos = min(pool_pos + (uint32_t)__popc(idle), pool_cnt);
        idle = __ballot_sync(0xffffffffu, !have);
        if (!idle) break;
      }
      if (idle == 0xffffffffu) break;  // pool dry, no tile left, nobody traversing
    }
    // One iteration = (a) a node visit for the lanes with no pending triangle, then (b) at most ONE triangle test per
    // lane.  A lane whose node produced several leaf hits keeps them in `tg` and skips (a) until they are done, so its
    // own sequence (node, all its triangles, next node) is unchanged — but the warp no longer waits, at 8 % utilisation,
    // for the one lane that hit ten triangles.  (a) is the expensive phase (~300 instructions): it only runs when at
    // least kNodeQuorum lanes want it or nobody has a triangle to test, so it executes with a well-filled warp while
    // the triangle-bound lanes catch up (profiles/r1_trace_persistent.md).
    const unsigned want_node = __ballot_sync(0xffffffffu, have && tg.y == 0u);
    const unsigned want_tri = __ballot_sync(0xffffffffu, have && tg.y != 0u);
    const bool run_node = __popc(want_node) >= kNodeQuorum || want_tri == 0u;
    if (have) {
      bool finished = false;
      const bool did_node = run_node && tg.y == 0u;
      if (did_node) {
        const uint32_t hits = cur.y;
        const uint32_t off = 31u - (uint32_t)__clz((int)hits);
        cur.y &= ~(1u << off);
        if (cur.y & 0xff000000u) {
          if (sp < kSmemStack)
            sstack[sp * kTraceBlock] = cur;
          else
            lstack[sp - kSmemStack] = cur;
          sp++;
        }
        const uint32_t cslot = (off - 24u) ^ oct_inv;
        const uint32_t rel = (uint32_t)__popc(hits & ~(0xffffffffu << cslot));
        const uint4* np = nodes + (size_t)(cur.x + rel) * 5;
        const uint4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
        if (STATS) nvis++;
        const bool negx = idx < 0.0f, negy = idy < 0.0f, negz = idz < 0.0f;  // from the sign bits: no flag registers
        // s = 2^e / d scaled by 2^15, o = (p - origin) / d with the bias s*2^15 taken off (see byte_to_float)
        const float sx = (__uint_as_float((n0.w & 0xffu) << 23) * idx) * 32768.0f;
        const float sy = (__uint_as_float(((n0.w >> 8) & 0xffu) << 23) * idy) * 32768.0f;
        const float sz = (__uint_as_float(((n0.w >> 16) & 0xffu) << 23) * idz) * 32768.0f;
        const float ox0 = (__uint_as_float(n0.x) - o.x) * idx;
        const float oy0 = (__uint_as_float(n0.y) - o.y) * idy;
        const float oz0 = (__uint_as_float(n0.z) - o.z) * idz;
        // Near planes are pulled in and far planes pushed out by 2^-22 of the operands that cancel (|o| + 2^15|s|):
        // this bounds the rounding error of the slab arithmetic (the box test is conservative, also against the
        // current best t), and it keeps a ray that runs exactly inside a box face (axis-parallel, origin on the
        // face) from being culled.  2^-22 * 2^15 |s| is 1/128 of a quantisation step.
        const float mgx = fabsf(ox0) + fabsf(sx), mgy = fabsf(oy0) + fabsf(sy), mgz = fabsf(oz0) + fabsf(sz);
        const float ox = fmaf(mgx, -kFarSlack, ox0) - sx, oxf = fmaf(mgx, kFarSlack, ox0) - sx;
        const float oy = fmaf(mgy, -kFarSlack, oy0) - sy, oyf = fmaf(mgy, kFarSlack, oy0) - sy;
        const float oz = fmaf(mgz, -kFarSlack, oz0) - sz, ozf = fmaf(mgz, kFarSlack, oz0) - sz;
        const float best_s = best.t * kBoxSlack;  // equal-t candidates behind a rounded box entry stay reachable
        // A child is entered iff none of (tfar - tnear), (best_s - tnear), (tfar - tmin) is negative: the slab test
        // max(tnear, tmin) <= min(tfar, best_s) as three subtractions on the FMA pipe whose sign bits are OR-ed (one LOP3)
        // and shifted into an 8-bit miss mask (one funnel shift).  Children run from slot 7 down, so slot s ends at bit s.
#if RTF_FFMA2
        // packed FFMA2 / FADD2 (fma.rn.f32x2, Blackwell): the near and the far plane of one axis in one instruction —
        // {tn, tf} = {qn, qf} * s + {o, of}, s broadcast — and the three differences of TWO children per instruction.
        // Every half is the same IEEE operation as in the scalar form, so results do not change.
        const float2 o_x = make_float2(ox, oxf), o_y = make_float2(oy, oyf), o_z = make_float2(oz, ozf);
        const float2 s_x = make_float2(sx, sx), s_y = make_float2(sy, sy), s_z = make_float2(sz, sz);
        const float2 m1 = make_float2(-1.0f, -1.0f), bs2 = make_float2(best_s, best_s), ntm2 = make_float2(-tmin, -tmin);
#endif
        uint32_t acc = 0;
#pragma unroll
        for (int half = 1; half >= 0; half--) {
          const uint32_t qlox = half ? n2.y : n2.x, qloy = half ? n2.w : n2.z, qloz = half ? n3.y : n3.x;
          const uint32_t qhix = half ? n3.w : n3.z, qhiy = half ? n4.y : n4.x, qhiz = half ? n4.w : n4.z;
          const uint32_t nx = negx ? qhix : qlox, fx = negx ? qlox : qhix;
          const uint32_t ny = negy ? qhiy : qloy, fy = negy ? qloy : qhiy;
          const uint32_t nz = negz ? qhiz : qloz, fz = negz ? qloz : qhiz;
#if RTF_FFMA2
#pragma unroll
          for (int j = 3; j >= 1; j -= 2) {
            float tn[2], tf[2];
#pragma unroll
            for (int c = 0; c < 2; c++) {  // children j and j-1
              const float2 px = __ffma2_rn(make_float2(byte_to_float(nx, j - c, magic), byte_to_float(fx, j - c, magic)), s_x, o_x);
              const float2 py = __ffma2_rn(make_float2(byte_to_float(ny, j - c, magic), byte_to_float(fy, j - c, magic)), s_y, o_y);
              const float2 pz = __ffma2_rn(make_float2(byte_to_float(nz, j - c, magic), byte_to_float(fz, j - c, magic)), s_z, o_z);
              tn[c] = fmaxf(fmaxf(px.x, py.x), pz.x);
              tf[c] = fminf(fminf(px.y, py.y), pz.y);
            }
            const float2 tn2 = make_float2(tn[0], tn[1]), tf2 = make_float2(tf[0], tf[1]);
            const float2 d1 = __ffma2_rn(tn2, m1, tf2);    // tfar - tnear (the product with -1 is exact)
            const float2 d2 = __ffma2_rn(tn2, m1, bs2);    // best_s - tnear
            const float2 d3 = __fadd2_rn(tf2, ntm2);       // tfar - tmin
            acc = __funnelshift_l(__float_as_uint(d1.x) | __float_as_uint(d2.x) | __float_as_uint(d3.x), acc, 1);
            acc = __funnelshift_l(__float_as_uint(d1.y) | __float_as_uint(d2.y) | __float_as_uint(d3.y), acc, 1);
          }
#else
#pragma unroll
          for (int j = 3; j >= 0; j--) {
            const float tnx = fmaf(byte_to_float(nx, j, magic), sx, ox);
            const float tny = fmaf(byte_to_float(ny, j, magic), sy, oy);
            const float tnz = fmaf(byte_to_float(nz, j, magic), sz, oz);
            const float tfx = fmaf(byte_to_float(fx, j, magic), sx, oxf);
            const float tfy = fmaf(byte_to_float(fy, j, magic), sy, oyf);
            const float tfz = fmaf(byte_to_float(fz, j, magic), sz, ozf);
            const float tnear = fmaxf(fmaxf(tnx, tny), tnz);
            const float tfar = fminf(fminf(tfx, tfy), tfz);
            const uint32_t sgn = __float_as_uint(tfar - tnear) | __float_as_uint(best_s - tnear) | __float_as_uint(tfar - tmin);
            acc = __funnelshift_l(sgn, acc, 1);
          }
#endif
        }
        // n1.z = VALID (bit 3s+j: leaf slot s holds more than j triangles), n1.w = imask | CNT << 8 (rtf_bvh.cuh)
        const uint32_t hit8 = ~acc & 0xffu;
        // leaf slots: bit s -> bits 3s..3s+2 (three "part by two" steps, the shifts as multiplies on the FMA pipe)
        uint32_t sp3 = (hit8 * 0x101u) & 0x00F00Fu;
        sp3 = (sp3 * 0x11u) & 0x0C30C3u;
        sp3 = (sp3 * 0x5u) & 0x249249u;
        // inner children: the hit bit of slot s moves to priority s ^ oct_inv — three conditional swap stages (adjacent
        // bits, pairs, nibbles); the stage masks of this ray sit in ONE register at bits 0, 10 and 20 and the byte moves along
        uint32_t pi = ~acc & n1.w & 0xffu;
        uint32_t sw = ((pi >> 1) ^ pi) & pmask;
        pi = (pi ^ (sw * 3u)) << 10;
        sw = ((pi >> 2) ^ pi) & pmask;
        pi = (pi ^ (sw * 5u)) << 10;
        sw = ((pi >> 4) ^ pi) & pmask;
        pi = pi ^ (sw * 17u);
        cur.x = n1.x;
        cur.y = (pi << 4) | n1.w;
        tg.x = n1.y;
        tg.y = (sp3 * 7u) & n1.z;
      }
      bool tested = false;
      do {
        if (tg.y == 0u) break;
        tested = true;
        const uint32_t pos = 31u - (uint32_t)__clz((int)tg.y);  // bit 3s+j = triangle j of leaf slot s
        tg.y &= ~(1u << pos);
        const uint32_t ts = (pos * 11u) >> 5;  // pos / 3 for pos < 24
        const uint32_t below = (cur.y >> 8) & ~(0xffffffffu << (2u * ts));  // 2-bit triangle counts of the slots below s
        const uint32_t ti = (pos - 3u * ts) + (uint32_t)__popc(below) + (uint32_t)__popc(below & 0xAAAAu);
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
        if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) break;
        const float det = (U + V) + W;
        if (det == 0.0f) break;
        const float Az = fmaf(mz.z, A.z, fmaf(mz.y, A.y, mz.x * A.x));
        const float Bz = fmaf(mz.z, B.z, fmaf(mz.y, B.y, mz.x * B.x));
        const float Cz = fmaf(mz.z, C.z, fmaf(mz.y, C.y, mz.x * C.x));
        const float T = fmaf(U, Az, fmaf(V, Bz, W * Cz));
        const float rcp = 1.0f / det;
        const float t = T * rcp;
        const uint32_t prim = __float_as_uint(t0.w);
        if (!(t >= tmin)) break;
        if (t > best.t) break;
        if (t == best.t && prim >= best.prim) break;
        const float bu = V * rcp, bv = W * rcp;
        if constexpr (Src::kAlpha) {
          if ((__float_as_uint(t2.w) & TRI_FLAG_ALPHA) || src.all_candidates())
            if (!src.alpha_pass(tag, prim, __float_as_uint(t1.w), bu, bv)) break;
        }
        best.t = t, best.u = bu, best.v = bv, best.prim = prim;
        if constexpr (Src::kAnyHit) {
          if (src.any_hit(tag)) finished = true;
        }
      } while (false);
      if (!finished && tg.y == 0u && (cur.y & 0xff000000u) == 0 && (did_node || tested)) {
        if (sp == 0) {
          finished = true;
        } else {
          sp--;
          cur = sp < kSmemStack ? sstack[sp * kTraceBlock] : lstack[sp - kSmemStack];
        }
      }
      if (finished) {
        src.store(ridx, tag, best, nvis, tvis);
        have = false;
      }
    }
  }
}

}  // namespace rtf
