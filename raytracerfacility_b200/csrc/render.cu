// render.cu — trace kernels and the wavefront path tracer behind the three entry points.
//
// Reference programs re-expressed as wavefront kernels (paths relative to /root/reference/src/RayTracerFacility/ptx/
// and /root/reference/include/RayTracerFacility/RayTracer/):
//   __raygen__IE   IlluminationEstimation.cu:30-129  -> k_gen_illum + k_trace_queue/k_shade loop + k_reduce_illum
//   __raygen__CR   CameraRendering.cu:32-147         -> k_gen_camera + loop + k_resolve_camera
//   __raygen__PCS  PointCloudScanning.cu:44-75       -> k_scan_grid / k_scan_samples (generate + trace + write)
//   ClosestHitFunc RayFunctions.cuh:25-172 (non-BTF), MissFunc :260-272, AnyHitFunc :5-23 -> k_shade / traverse()
// The reference runs one thread per triangle/pixel with a serial loop over samples and recursive
// optixTrace; here one lane owns one path, rays live in SoA float4 queues that are compacted with
// warp ballots between bounces, and per-triangle sums use a fixed reduction tree (no float atomics).
#include <cstdlib>

#include "rtf_internal.h"
#include "rtf_shade.cuh"
#include "rtf_trace.cuh"
#include "rtf_sss.cuh"
#include "rtf_sort.cuh"

namespace rtf {

static_assert(kSmemStack + kLocalStack == kMaxStackDepth, "stack split must cover kMaxStackDepth");

#define LAUNCH(ctx, kernel, grid, block, smem, ...)                  \
  do {                                                               \
    kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__); \
    (ctx)->launches++;                                               \
  } while (0)

#define LAUNCH_P(ctx, cat, kernel, grid, block, smem, ...)            \
  do {                                                               \
    (ctx)->prof.begin((cat), (ctx)->stream);                         \
    kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__); \
    (ctx)->launches++;                                               \
    (ctx)->prof.end((ctx)->stream);                                  \
  } while (0)

static inline unsigned div_up(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

constexpr uint32_t kShadowBit = 0x80000000u;   // first-hit ray (TERMINATE_ON_FIRST_HIT | DISABLE_ANYHIT | DISABLE_CLOSESTHIT)
constexpr uint32_t kNoAnyHitBit = 0x40000000u; // closest-hit ray traced with OPTIX_RAY_FLAG_DISABLE_ANYHIT (BTF bounce, RF.cuh:239)
constexpr uint32_t kPathMask = ~(kShadowBit | kNoAnyHitBit);  // path index inside the batch (batch_paths <= 2^28)
// resident blocks per SM the register allocator must allow: 8 x 128 threads = 64 registers (measured 3118 vs 3017
// Mrays/s at 80 registers / 6 blocks); the queue kernel with stochastic alpha carries the path map and RNG keys: 80 registers / 6 blocks (spilling at 72 cost 12 %).
#ifndef RTF_TRACE_MIN_BLOCKS
#define RTF_TRACE_MIN_BLOCKS 8
#endif
#ifndef RTF_QUEUE_MIN_BLOCKS
#define RTF_QUEUE_MIN_BLOCKS 6
#endif

// ============================ raw closest-hit (M1 metric) =========================================
// Persistent lanes (rtf_trace.cuh): rays are the interleaved n x 8 float layout {o.xyz, tmin, d.xyz, tmax};
// a 32-ray tile is one contiguous 1 KB TMA bulk copy.
struct SrcRays {
  static constexpr bool kTma = true, kAnyHit = false, kAlpha = false;
  const float4* rays;
  float4* hits;
  uint32_t* nv;
  uint32_t* tv;
  uint32_t n;
  RTF_D void issue(uint32_t base, uint32_t cnt, uint32_t raw_s, uint32_t mbar_s) const {
    mbar_expect_tx(mbar_s, cnt * 32u);
    tma_load_1d(raw_s, rays + 2 * (size_t)base, cnt * 32u, mbar_s);
  }
  RTF_D void fetch(const unsigned char* raw, uint32_t, uint32_t slot, float4& a, float4& b, float& tmin, uint32_t& tag) const {
    const float4* r = (const float4*)raw;
    a = r[2 * slot], b = r[2 * slot + 1];
    tmin = a.w, tag = 0u;
  }
  RTF_D bool any_hit(uint32_t) const { return false; }
  RTF_D bool alpha_pass(uint32_t, uint32_t, uint32_t, float, float) const { return true; }
  RTF_D void store(uint32_t i, uint32_t, const HitRec& h, uint32_t nvis, uint32_t tvis) const {
    hits[i] = make_float4(h.t, h.u, h.v, __uint_as_float(h.prim));
    if (nv) nv[i] = nvis;
    if (tv) tv[i] = tvis;
  }
};
template <bool STATS>
__global__ void __launch_bounds__(kTraceBlock, RTF_TRACE_MIN_BLOCKS) k_trace_rays(DevScene sc, SrcRays src, uint32_t* work) {
  __shared__ TraceSmem sm;
  trace_engine<STATS>(sc.nodes, sc.tris, src, work, sm);
}

// Persistent-lane kernels run one resident wave: SMs x the blocks of THIS kernel variant that fit on an SM (its register
// count and launch bounds differ per variant), asked of the runtime at the variant's first launch.
enum { TK_RAYS = 0, TK_RAYS_STATS, TK_QUEUE, TK_QUEUE_ALPHA, TK_GRID, TK_SAMPLES };
template <class Kernel>
static unsigned trace_grid(rtf_ctx* ctx, int which, Kernel kernel, size_t n) {
  if (ctx->trace_blocks[which] == 0) {
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, kTraceBlock, 0) != cudaSuccess || nb < 1) nb = 1;
    ctx->trace_blocks[which] = nb;
  }
  const unsigned full = (unsigned)ctx->sm_count * (unsigned)ctx->trace_blocks[which];
  const unsigned need = div_up(n, kTraceBlock);
  return need < full ? need : full;
}

int trace_rays_device(rtf_ctx* ctx, const float4* rays, size_t n, float4* hits, uint32_t* nv, uint32_t* tv) {
  if (!ctx->ready) {
    set_error("no acceleration structure (call rtf_commit)");
    return RTF_ERR_NOT_READY;
  }
  if (n == 0) return RTF_OK;
  if (((uintptr_t)rays & 15u) || ((uintptr_t)hits & 15u)) {
    set_error("ray and hit buffers must be 16-byte aligned");
    return RTF_ERR_INVALID;
  }
  int rc;
  if ((rc = ctx->ws.counters.reserve(4096))) return rc;
  uint32_t* work = ctx->ws.counters.as<uint32_t>();
  const size_t chunk = 1u << 30;
  for (size_t off = 0; off < n; off += chunk) {
    const uint32_t m = (uint32_t)((n - off) < chunk ? (n - off) : chunk);
    RTF_CUDA(cudaMemsetAsync(work, 0, 4, ctx->stream));
    SrcRays src{rays + 2 * off, hits + off, nv ? nv + off : nullptr, tv ? tv + off : nullptr, m};
    if (nv || tv)
      LAUNCH(ctx, k_trace_rays<true>, trace_grid(ctx, TK_RAYS_STATS, k_trace_rays<true>, m), kTraceBlock, 0, ctx->scene, src, work);
    else
      LAUNCH(ctx, k_trace_rays<false>, trace_grid(ctx, TK_RAYS, k_trace_rays<false>, m), kTraceBlock, 0, ctx->scene, src, work);
  }
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}

// ============================ wavefront machinery ==================================================
// How a path index of the current batch maps to its RNG unit (and probe / pixel).
// Local pixel L of a camera job -> pixel of the frame.  band_px == 0: the job is the contiguous range from `first`.
// Otherwise the job is this rank's interleaved row bands (multi-GPU): local band L / band_px is band
// (L / band_px) * stride + rank of the frame, band_px = band rows * width pixels each — all bands of a rank run as ONE
// job, batched like a full frame (one job per band cost a host round trip and five short wavefronts each).
// A PixelMap describes one BATCH: k is the pixel's index inside the batch (32-bit arithmetic only: a 64-bit division per
// path doubled the time of the generation kernel).
struct PixelMap {
  uint64_t first;                  // contiguous: first pixel of the batch
  uint32_t band_px, stride, rank;  // interleaved: pixels per band (< 2^31), number of ranks, this rank
  uint64_t blk0;                   // interleaved: local band and offset inside it of the batch's first pixel
  uint32_t off0;
};
__device__ __forceinline__ uint64_t pixel_of(const PixelMap& m, uint32_t k) {
  if (m.band_px == 0u) return m.first + k;
  const uint32_t t = m.off0 + k, b = t / m.band_px, off = t - b * m.band_px;
  return ((m.blk0 + b) * m.stride + m.rank) * (uint64_t)m.band_px + off;
}
struct PathMap {
  int mode;            // 0 = illumination, 1 = camera
  uint64_t unit_base;  // camera, contiguous job: first pixel of the batch * spp
  uint64_t tri_begin;  // illumination: first probe of the batch
  uint32_t per_probe;  // 2 * sc
  uint32_t sc;         // samples in this call
  uint32_t s0;         // first sample
  uint32_t s_total;    // RayProperties::m_samples
  PixelMap px;         // camera, interleaved job (px.band_px != 0): pixel map of the batch, spp
  uint32_t spp;
};
__device__ __forceinline__ uint64_t unit_of(const PathMap& pm, uint32_t p) {
  if (pm.mode == 1) {
    if (pm.px.band_px == 0u) return pm.unit_base + p;
    const uint32_t k = p / pm.spp;
    return pixel_of(pm.px, k) * (uint64_t)pm.spp + (p - k * pm.spp);
  }
  const uint32_t tl = p / pm.per_probe, rem = p - tl * pm.per_probe;
  const uint32_t face = rem / pm.sc, s = pm.s0 + (rem - face * pm.sc);
  return ((pm.tri_begin + tl) * 2ull + face) * (uint64_t)pm.s_total + s;
}

struct Queues {
  float4* ray_a[2];  // {origin.xyz, path index | kShadowBit (bits)}
  float4* ray_b[2];  // {direction.xyz, tmax}
  float4* hits;
  uint32_t* count;  // count[k] = rays in wavefront k
  uint32_t* work;   // work[k]  = dynamic-fetch cursor of wavefront k
};
struct PathState {
  float4* thr;    // throughput.xyz, w = hit count (bits) | emission (shadow pending)
  float4* acc;    // accumulated radiance.xyz
  float4* dir0w;  // illumination: first direction.xyz, w = |N.dir|
  float4* aux;    // shadow pending: albedo.xyz, w = N.L
  float4* first_n;  // camera: first-hit normal
  float4* first_a;  // camera: first-hit albedo (or miss colour)
  float4* first_p;  // camera: first-hit position, w = 1 if there was a first hit
  // BSSRDF (scenes with a subsurface material only): the bounce ray that waits for the subsurface path to end
  float4* pend_o;  // origin.xyz, w = 1 while pending
  float4* pend_d;  // direction.xyz
  float4* pend_t;  // throughput.xyz of the bounce path
};

// warp-aggregated append: one atomicAdd per warp (ballot + popc), slots keep lane order.
__device__ __forceinline__ uint32_t warp_append(bool emit, uint32_t* counter) {
  const unsigned m = __ballot_sync(0xffffffffu, emit);
  if (m == 0) return 0;
  const int lane = threadIdx.x & 31;
  const int leader = __ffs(m) - 1;
  uint32_t base = 0;
  if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  return base + (uint32_t)__popc(m & ((1u << lane) - 1u));
}

// ---- trace one wavefront -----------------------------------------------------------------------------
// Queue rays are SoA: a 32-ray tile is two 512 B TMA bulk copies (origins|path, directions|tmax).  tmin is
// uniform per wavefront (camera primary rays 0, everything else 1e-3: CameraRendering.cu:89, IlluminationEstimation.cu:69,
// RayFunctions.cuh:77,140), so the .w of the origin stream carries the path index instead.
template <bool ALPHA>
struct SrcQueue {
  static constexpr bool kTma = true, kAnyHit = true, kAlpha = ALPHA;
  const float4* ra;
  const float4* rb;
  float4* hits;
  uint32_t n;
  float tmin;
  const float* inst_alpha;  // nullptr: no material can ignore hits
  DevScene sc;              // alpha from an albedo texture needs the hit's texture coordinate (ALPHA variant only)
  PathMap pm;
  uint32_t seed, bounce;
  uint32_t* lcg;  // option "rng" = reference: per-path LCG state, drawn from in traversal order (else nullptr)
  // The reference's any-hit program draws at EVERY candidate of a default-material surface, also where alpha = 1 and the
  // draw cannot reject (RF.cuh:19): a sequential stream sees those draws, so that mode visits every candidate.
  RTF_D bool all_candidates() const { return ALPHA && lcg != nullptr; }
  RTF_D void issue(uint32_t base, uint32_t cnt, uint32_t raw_s, uint32_t mbar_s) const {
    mbar_expect_tx(mbar_s, cnt * 32u);
    tma_load_1d(raw_s, ra + base, cnt * 16u, mbar_s);
    tma_load_1d(raw_s + 512u, rb + base, cnt * 16u, mbar_s);
  }
  RTF_D void fetch(const unsigned char* raw, uint32_t, uint32_t slot, float4& a, float4& b, float& tmin_, uint32_t& tag) const {
    const float4* r = (const float4*)raw;
    a = r[slot], b = r[32 + slot];
    tmin_ = tmin, tag = __float_as_uint(a.w);
  }
  RTF_D bool any_hit(uint32_t tag) const { return (tag & kShadowBit) != 0u; }  // TERMINATE_ON_FIRST_HIT, RayFunctions.cuh:81-83
  // AnyHitFunc, RayFunctions.cuh:5-23: ignore the candidate if alpha <= rnd; the draw is keyed by (path, bounce, primitive)
  // so it does not depend on traversal order.  Shadow rays and BTF bounce rays run with DISABLE_ANYHIT.
  RTF_D bool alpha_pass(uint32_t tag, uint32_t prim, uint32_t inst, float bu, float bv) const {
    if (inst_alpha == nullptr || (tag & (kShadowBit | kNoAnyHitBit))) return true;
    if (lcg != nullptr) {  // reference stream: one draw per any-hit invocation (RF.cuh:5-23)
      if (sc.mats[sc.insts[inst].mat].m.material_type != RTF_MAT_DEFAULT) return true;
      float al = inst_alpha[inst];
      if (al < 0.0f) {
        const HitAttrs at = hit_attrs(sc, prim, bu, bv, inst);
        al = tex2d(sc.mats[sc.insts[inst].mat].tex[MAT_TEX_ALBEDO], at.u, at.v).w;
      }
      uint32_t st = lcg[tag & kPathMask];
      const float rnd = lcg_next(st);
      lcg[tag & kPathMask] = st;
      return !(al <= rnd);
    }
    float alpha = inst_alpha[inst];
    if (alpha < 0.0f) {  // SurfaceMaterial::GetAlbedo(texCoord).w, RDD.hpp:248-254
      const HitAttrs at = hit_attrs(sc, prim, bu, bv, inst);
      alpha = tex2d(sc.mats[sc.insts[inst].mat].tex[MAT_TEX_ALBEDO], at.u, at.v).w;
    }
    const float rnd = path_rng(seed, STREAM_ALPHA, unit_of(pm, tag & kPathMask), bounce, prim).x;
    return !(alpha <= rnd);
  }
  RTF_D void store(uint32_t i, uint32_t, const HitRec& h, uint32_t, uint32_t) const {
    hits[i] = make_float4(h.t, h.u, h.v, __uint_as_float(h.prim));
  }
};
// ALPHA = some material can ignore hits (transmission > 0): only then does the kernel carry the path map and RNG keys.
template <bool ALPHA>
__global__ void __launch_bounds__(kTraceBlock, ALPHA ? RTF_QUEUE_MIN_BLOCKS : RTF_QUEUE_MIN_BLOCKS + 1)
    k_trace_queue(DevScene sc, Queues q, int qi, int wave, float tmin, PathMap pm, uint32_t seed, unsigned long long* stats,
                  uint32_t* lcg) {
  __shared__ TraceSmem sm;
  const uint32_t n = q.count[wave];
  if (blockIdx.x == 0 && threadIdx.x == 0 && stats) atomicAdd(stats, (unsigned long long)n);
  SrcQueue<ALPHA> src{q.ray_a[qi], q.ray_b[qi], q.hits, n, tmin, ALPHA ? sc.inst_alpha : nullptr, sc, pm, seed, (uint32_t)wave,
                      ALPHA ? lcg : nullptr};
  trace_engine<false>(sc.nodes, sc.tris, src, q.work + wave, sm);
}

// ---- shade one wavefront: ClosestHitFunc (RayFunctions.cuh:25-172) / MissFunc (:260-272), unrolled ------
struct ShadeParams {
  int bounces;
  uint32_t seed;
  int camera;
  int brdf;  // RTF_BRDF_REFERENCE / RTF_BRDF_LAMBERT_GGX
  uint32_t* lcg;  // option "rng" = reference: per-path LCG state (else nullptr: keyed Philox draws)
};
// The bounce draws of a shading event: keyed Philox numbers, or the next `n` numbers of the path's reference stream.
__device__ __forceinline__ float4 shade_draws(const ShadeParams& sp, uint32_t p, uint64_t unit, uint32_t hit_count, int n) {
  if (sp.lcg == nullptr) return path_rng(sp.seed, STREAM_PATH, unit, hit_count, 0u);
  uint32_t st = sp.lcg[p];
  float4 r = make_float4(0, 0, 0, 0);
  r.x = lcg_next(st);
  if (n > 1) r.y = lcg_next(st);
  if (n > 2) r.z = lcg_next(st);
  sp.lcg[p] = st;
  return r;
}
// SSS = some material has subsurfaceFactor > 0: the kernel carries the BSSRDF spatial probe (rtf_sss.cuh) and the
// pending-bounce state.
#ifndef RTF_SHADE_MIN_BLOCKS
#define RTF_SHADE_MIN_BLOCKS 3
#endif
// BTF = some material is a compressed BTF: the kernel carries the MLVQ evaluator in its wide (level-by-level) form, which
// wants registers for the loads it keeps in flight; scenes without one run the leaner variant.  Measured on C5: wide at
// 2 blocks/SM (96 registers) 156 ms of shading, wide at 3 blocks (80 registers) 134.7 ms, depth-first at 3 blocks 135.1 ms.
#ifndef RTF_SHADE_BTF_MIN_BLOCKS
#define RTF_SHADE_BTF_MIN_BLOCKS 3
#endif
#ifndef RTF_SHADE_SSS_MIN_BLOCKS
#define RTF_SHADE_SSS_MIN_BLOCKS 3  // C2 scene, every plant subsurface: shading 23.6 / 15.0 / 13.3 ms at 1 / 2 / 3 blocks
#endif
#ifndef RTF_BTF_WIDE
#define RTF_BTF_WIDE 1
#endif
template <bool SSS, bool BTF>
__global__ void __launch_bounds__(256, SSS ? RTF_SHADE_SSS_MIN_BLOCKS : (BTF ? RTF_SHADE_BTF_MIN_BLOCKS : RTF_SHADE_MIN_BLOCKS))
    k_shade(DevScene sc, rtf_env env, ShadeParams sp, PathMap pm, Queues q, int qi, int wave, PathState ps) {
  const uint32_t n = q.count[wave];
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (blockIdx.x * blockDim.x >= n) return;  // whole block idle (uniform exit keeps warp_append safe)
  bool emit = false;
  float4 na = make_float4(0, 0, 0, 0), nb = make_float4(0, 0, 0, 0);
  if (i < n) {
    const float4 a = q.ray_a[qi][i], b = q.ray_b[qi][i];
    const uint32_t pi = __float_as_uint(a.w);
    const uint32_t p = pi & kPathMask;
    const float4 h = q.hits[i];
    const uint32_t prim = __float_as_uint(h.w);
    const float3 o = f3(a.x, a.y, a.z), d = f3(b.x, b.y, b.z);
    const float4 thr4 = ps.thr[p];
    float3 T = f3(thr4.x, thr4.y, thr4.z);
    const float4 acc4 = ps.acc[p];
    float3 acc = f3(acc4.x, acc4.y, acc4.z);
    uint32_t count_now = __float_as_uint(thr4.w);  // PerRayData::m_hitCount after this event
    if (pi & kShadowBit) {  // shadow ray result, RayFunctions.cuh:70-92 (BTF: :193-217)
      const float4 ax = ps.aux[p];
      const float3 albedo = f3(ax.x, ax.y, ax.z);
      if (prim == 0xFFFFFFFFu) {
        const float3 L = env_light(o, d, env, sc.env_faces);
        acc = acc + T * ((L * ax.w) * albedo);
      }
      acc = acc + T * (thr4.w * albedo);
      ps.acc[p] = make_float4(acc.x, acc.y, acc.z, 0.0f);
    } else if (prim == 0xFFFFFFFFu) {  // MissFunc
      const float3 L = env_light(o, d, env, sc.env_faces);
      acc = acc + T * L;
      ps.acc[p] = make_float4(acc.x, acc.y, acc.z, 0.0f);
      if (sp.camera && __float_as_uint(thr4.w) == 0u) ps.first_a[p] = make_float4(L.x, L.y, L.z, 0.0f);
    } else {
      const uint32_t hit_count = __float_as_uint(thr4.w) + 1u;
      count_now = hit_count;
      const SurfaceHit sh = surface_hit(sc, prim, h.y, h.z, d);
      const DevMaterial& dm = sc.mats[sc.insts[sh.inst].mat];
      const rtf_material& m = dm.m;
      HitAttrs at;
      at.u = 0.0f, at.v = 0.0f, at.tangent = f3(0, 0, 0), at.color = f3(0, 0, 0);
      if (dm.rich) at = hit_attrs(sc, prim, h.y, h.z, sh.inst);
      const uint64_t unit = unit_of(pm, p);
      bool is_btf = false;
      if constexpr (BTF) is_btf = m.material_type == RTF_MAT_COMPRESSED_BTF;
      if (is_btf) {  // RayFunctions.cuh:172-258
        const DevBTF& btf = sc.btfs[dm.btf];
        float3 btf_color = f3(0, 0, 0);  // the reference leaves it uninitialised past the bounce limit: 0 here
        if ((int)hit_count <= sp.bounces) {
          const float4 r = shade_draws(sp, p, unit, hit_count, 2);
          if (env.lighting_type == RTF_ENV_SINGLE_LIGHT_SOURCE) {  // :180-218
            const float3 sun = f3(env.sun_direction[0], env.sun_direction[1], env.sun_direction[2]);
            const float3 nd = sample_lobe(r.x, r.y, sun, 1.0f - env.light_size);
            btf_color = btf_get_value<RTF_BTF_WIDE != 0>(btf, at.u, at.v, d, nd, sh.normal, at.tangent);
            const float3 col = f3(env.color[0], env.color[1], env.color[2]);
            acc = acc + T * ((col * env.ambient_light_intensity) * btf_color);
            const float ndl = dot3(sh.normal, nd);
            if (ndl > 0.0f) {
              const float3 so = sh.position + sh.normal * 1e-3f;
              emit = true;
              na = make_float4(so.x, so.y, so.z, __uint_as_float(p | kShadowBit));
              nb = make_float4(nd.x, nd.y, nd.z, 1e20f);
              ps.thr[p] = make_float4(T.x, T.y, T.z, 0.0f);  // no emission term on this branch
              ps.aux[p] = make_float4(btf_color.x, btf_color.y, btf_color.z, ndl);
            }
            ps.acc[p] = make_float4(acc.x, acc.y, acc.z, 0.0f);
          } else {  // :219-246
            const float3 refl = reflect3(d, sh.normal);
            const float3 nd = sample_lobe(r.x, r.y, refl, 0.0f);
            btf_color = btf_get_value<RTF_BTF_WIDE != 0>(btf, at.u, at.v, d, nd, sh.normal, at.tangent);
            T = T * btf_color;
            const float3 so = sh.position + sh.normal * 1e-3f;
            ps.thr[p] = make_float4(T.x, T.y, T.z, __uint_as_float(hit_count));
            emit = true;
            na = make_float4(so.x, so.y, so.z, __uint_as_float(p | kNoAnyHitBit));
            nb = make_float4(nd.x, nd.y, nd.z, 1e20f);
          }
        }
        if (sp.camera && hit_count == 1u) {
          ps.first_n[p] = make_float4(sh.normal.x, sh.normal.y, sh.normal.z, 0.0f);
          ps.first_a[p] = make_float4(btf_color.x, btf_color.y, btf_color.z, 0.0f);
          ps.first_p[p] = make_float4(sh.position.x, sh.position.y, sh.position.z, 1.0f);
        }
      } else {  // surface materials, RayFunctions.cuh:44-172
        float3 normal = sh.normal;
        if (dm.tex[MAT_TEX_NORMAL].texels) {  // ApplyNormalTexture, RDD.hpp:268-281
          const float4 t = tex2d(dm.tex[MAT_TEX_NORMAL], at.u, at.v);
          const float3 B = cross3(normal, at.tangent);
          const float3 tn = f3(t.x * 2.0f - 1.0f, t.y * 2.0f - 1.0f, t.z * 2.0f - 1.0f);
          normal = normalize3((at.tangent * tn.x + B * tn.y) + normal * tn.z);
        }
        float metallic = m.metallic, roughness = m.roughness;
        if (dm.tex[MAT_TEX_METALLIC].texels) metallic = tex2d(dm.tex[MAT_TEX_METALLIC], at.u, at.v).x;
        if (dm.tex[MAT_TEX_ROUGHNESS].texels) roughness = tex2d(dm.tex[MAT_TEX_ROUGHNESS], at.u, at.v).x;
        float3 albedo = f3(m.albedo_color[0], m.albedo_color[1], m.albedo_color[2]);
        if (m.material_type != RTF_MAT_DEFAULT) {
          albedo = at.color;  // VertexColor: hitInfo.m_color, RayFunctions.cuh:55-57
        } else if (dm.tex[MAT_TEX_ALBEDO].texels) {
          const float4 t = tex2d(dm.tex[MAT_TEX_ALBEDO], at.u, at.v);
          albedo = f3(t.x, t.y, t.z);
        }
        if (sp.camera && hit_count == 1u) {
          ps.first_n[p] = make_float4(normal.x, normal.y, normal.z, 0.0f);
          ps.first_a[p] = make_float4(albedo.x, albedo.y, albedo.z, 0.0f);
          ps.first_p[p] = make_float4(sh.position.x, sh.position.y, sh.position.z, 1.0f);
        }
        float f = 1.0f;
        if (metallic >= 0.0f) f = (metallic + 2) / (metallic + 1);
        const float3 emis = m.emission * albedo;
        if (env.lighting_type == RTF_ENV_SINGLE_LIGHT_SOURCE) {  // RayFunctions.cuh:61-92
          const float4 r = shade_draws(sp, p, unit, hit_count, 2);
          const float3 sun = f3(env.sun_direction[0], env.sun_direction[1], env.sun_direction[2]);
          const float3 nd = sample_lobe(r.x, r.y, sun, 1.0f - env.light_size);
          const float3 col = f3(env.color[0], env.color[1], env.color[2]);
          acc = acc + T * ((col * env.ambient_light_intensity) * albedo);
          const float ndl = dot3(normal, nd);
          if (ndl > 0.0f) {
            emit = true;
            na = make_float4(sh.position.x, sh.position.y, sh.position.z, __uint_as_float(p | kShadowBit));
            nb = make_float4(nd.x, nd.y, nd.z, 1e20f);
            ps.thr[p] = make_float4(T.x, T.y, T.z, m.emission);
            ps.aux[p] = make_float4(albedo.x, albedo.y, albedo.z, ndl);
          } else {
            acc = acc + T * emis;
          }
          ps.acc[p] = make_float4(acc.x, acc.y, acc.z, 0.0f);
        } else if ((int)hit_count <= sp.bounces) {  // RayFunctions.cuh:93-161
          bool need_sample = false;  // BSSRDF, RayFunctions.cuh:97-133 + BSSDF.cuh:42-89
          float3 ss_o = f3(0, 0, 0), ss_d = f3(0, 0, 0), ss_w = f3(0, 0, 0);
          if constexpr (SSS) {
            if (hit_count <= 1u && m.subsurface_factor > 0.0f && m.subsurface_radius[0] > 0.0f) {
              const float radius = m.subsurface_radius[0];
              const float4 rs = path_rng(sp.seed, STREAM_SSS, unit, hit_count, 0u);
              const float4 rb = path_rng(sp.seed, STREAM_SSS, unit, hit_count, 1u);
              const float3 disk_center = sh.position + (radius * normal) * 0.5f;
              const float disk_radius = radius * sqrtf(rs.x);
              const float distance = sqrtf(radius * radius - disk_radius * disk_radius);
              const float3 sample_pos =
                  disk_center + disk_radius * rotate_about(f3(normal.y, normal.z, normal.x), rs.y, normal);
              SSRecords rec;
              ss_probe(sc, sample_pos, neg3(normal), distance, radius + distance, sc.insts[sh.inst].handle, rec);
              if (rec.n > 0) {
                // the reference picks the record with a COPY of its generator (BSSDF.cuh:58,77): the pick is the same
                // number as the exit lobe's first draw (BSSDF.cuh:83)
                int index = (int)(rb.x * (float)rec.n);
                index = index < 0 ? 0 : (index > rec.n - 1 ? rec.n - 1 : index);
                float rt = 0, ru = 0, rv = 0;
                uint32_t rprim = 0;
#pragma unroll
                for (int k = 0; k < 4; k++)
                  if (k == index) rt = rec.t[k], ru = rec.u[k], rv = rec.v[k], rprim = rec.prim[k];
                (void)rt;
                const SurfaceHit so = surface_hit(sc, rprim, ru, rv, f3(0, 0, 0));  // GetHitInfo(rayDirection, false): no flip
                float3 out_n = so.normal;
                const DevMaterial& om = sc.mats[sc.insts[so.inst].mat];
                if (om.tex[MAT_TEX_NORMAL].texels) {  // ApplyNormalTexture of the record's own material, BSSDF.cuh:31-32
                  const HitAttrs oa = hit_attrs(sc, rprim, ru, rv, so.inst);
                  const float4 t = tex2d(om.tex[MAT_TEX_NORMAL], oa.u, oa.v);
                  const float3 B = cross3(out_n, oa.tangent);
                  const float3 tn = f3(t.x * 2.0f - 1.0f, t.y * 2.0f - 1.0f, t.z * 2.0f - 1.0f);
                  out_n = normalize3((oa.tangent * tn.x + B * tn.y) + out_n * tn.z);
                }
                if (length3(sh.position - so.position) <= radius) {
                  ss_o = so.position + out_n * 0.01f;
                  ss_d = sample_lobe(rb.x, rb.y, reflect3(neg3(out_n), out_n), metallic);  // BRDF(metallic, rnd, -n, n)
                  const float k2 = clampf(fabsf(dot3(out_n, ss_d)) * roughness + (1.0f - roughness) * f, 0.0f, 1.0f);
                  ss_w = (m.subsurface_factor * f3(m.subsurface_color[0], m.subsurface_color[1], m.subsurface_color[2])) * k2;
                  need_sample = true;
                }
              }
            }
          }
          const float4 r = shade_draws(sp, p, unit, hit_count, sp.brdf == RTF_BRDF_LAMBERT_GGX ? 3 : 2);
          float3 nd, wgt;
          if (sp.brdf == RTF_BRDF_LAMBERT_GGX) {
            wgt = sample_lambert_ggx(albedo, metallic, roughness, normal, neg3(d), r.x, r.y, r.z, nd);
          } else {
            const float3 refl = reflect3(d, normal);
            nd = sample_lobe(r.x, r.y, refl, metallic);  // BRDF(), BSDF.cuh:6-13
            const float k = clampf(fabsf(dot3(normal, nd)) * roughness + (1.0f - roughness) * f, 0.0f, 1.0f);
            wgt = ((1.0f - m.subsurface_factor) * albedo) * k;
          }
          acc = acc + T * emis;
          ps.acc[p] = make_float4(acc.x, acc.y, acc.z, 0.0f);
          const bool absorbed = wgt.x == 0.0f && wgt.y == 0.0f && wgt.z == 0.0f;  // (only the GGX lobe can absorb a sample)
          if (SSS && need_sample) {  // the subsurface path runs first; the bounce waits for it (shared m_hitCount)
            if (!absorbed) {
              const float3 tb = T * wgt;
              ps.pend_o[p] = make_float4(sh.position.x, sh.position.y, sh.position.z, 1.0f);
              ps.pend_d[p] = make_float4(nd.x, nd.y, nd.z, 0.0f);
              ps.pend_t[p] = make_float4(tb.x, tb.y, tb.z, 0.0f);
            }
            T = T * ss_w;
            ps.thr[p] = make_float4(T.x, T.y, T.z, __uint_as_float(hit_count));
            emit = true;
            na = make_float4(ss_o.x, ss_o.y, ss_o.z, __uint_as_float(p));
            nb = make_float4(ss_d.x, ss_d.y, ss_d.z, 1e20f);
          } else if (!absorbed) {
            T = T * wgt;
            ps.thr[p] = make_float4(T.x, T.y, T.z, __uint_as_float(hit_count));
            emit = true;
            na = make_float4(sh.position.x, sh.position.y, sh.position.z, __uint_as_float(p));  // origin not offset, RayFunctions.cuh:138
            nb = make_float4(nd.x, nd.y, nd.z, 1e20f);
          }
        } else {
          acc = acc + T * emis;
          ps.acc[p] = make_float4(acc.x, acc.y, acc.z, 0.0f);
        }
      }
    }
    if constexpr (SSS) {  // a path that ended here resumes with its waiting bounce ray, hit count carried over
      if (!emit && !(pi & kShadowBit)) {
        const float4 po = ps.pend_o[p];
        if (po.w != 0.0f) {
          const float4 pd = ps.pend_d[p], pt = ps.pend_t[p];
          ps.pend_o[p] = make_float4(0, 0, 0, 0);
          ps.thr[p] = make_float4(pt.x, pt.y, pt.z, __uint_as_float(count_now));
          emit = true;
          na = make_float4(po.x, po.y, po.z, __uint_as_float(p));
          nb = make_float4(pd.x, pd.y, pd.z, 1e20f);
        }
      }
    }
  }
  const uint32_t slot = warp_append(emit, q.count + wave + 1);
  if (emit) {
    q.ray_a[qi ^ 1][slot] = na;
    q.ray_b[qi ^ 1][slot] = nb;
  }
}

// ---- __raygen__IE path start (IlluminationEstimation.cu:47-80) --------------------------------------------
struct IllumGen {
  uint64_t tri_begin;
  uint32_t n_tris;  // probes in this batch
  uint32_t sc, s0, s_total;
  uint32_t seed;
  float push;
  int front, back;                   // mesh mode
  const rtf_illum_sampler* probes;   // probe mode (AoS, device) or nullptr
  uint64_t probe_base;               // index of probes[0] in the caller's array (= RNG probe id)
};
__global__ void __launch_bounds__(256) k_gen_illum(DevScene sc, IllumGen g, Queues q, PathState ps) {
  const uint32_t per_probe = 2u * g.sc;
  const uint64_t total = (uint64_t)g.n_tris * per_probe;
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  bool emit = false;
  float4 na = make_float4(0, 0, 0, 0), nb = make_float4(0, 0, 0, 0);
  if (p < total) {
    const uint32_t tl = p / per_probe, rem = p - tl * per_probe;
    const uint32_t face = rem / g.sc, s = g.s0 + (rem - face * g.sc);
    const uint64_t probe = g.tri_begin + tl;
    float3 P0, P1, P2, N0, N1, N2;
    bool enabled;
    if (g.probes) {
      const rtf_illum_sampler& pr = g.probes[probe - g.probe_base];
      P0 = f3(pr.a.position[0], pr.a.position[1], pr.a.position[2]);
      P1 = f3(pr.b.position[0], pr.b.position[1], pr.b.position[2]);
      P2 = f3(pr.c.position[0], pr.c.position[1], pr.c.position[2]);
      N0 = f3(pr.a.normal[0], pr.a.normal[1], pr.a.normal[2]);
      N1 = f3(pr.b.normal[0], pr.b.normal[1], pr.b.normal[2]);
      N2 = f3(pr.c.normal[0], pr.c.normal[1], pr.c.normal[2]);
      enabled = face == 0 ? pr.front_face != 0 : pr.back_face != 0;
    } else {
      const float* pp = sc.pos9 + 9 * probe;
      const float* nn = sc.nrm9 + 9 * probe;
      P0 = f3(pp[0], pp[1], pp[2]), P1 = f3(pp[3], pp[4], pp[5]), P2 = f3(pp[6], pp[7], pp[8]);
      N0 = f3(nn[0], nn[1], nn[2]), N1 = f3(nn[3], nn[4], nn[5]), N2 = f3(nn[6], nn[7], nn[8]);
      enabled = face == 0 ? g.front != 0 : g.back != 0;
    }
    float4 d0w = make_float4(0, 0, 0, 0);
    if (enabled) {
      const uint64_t unit = (probe * 2ull + face) * (uint64_t)g.s_total + s;
      const float4 r = path_rng(g.seed, STREAM_PATH, unit, 0u, 0u);
      const float a = r.x, b = r.y;
      const float w = (1.0f - a) - b;  // not folded back when a+b > 1 (IlluminationEstimation.cu:56-59)
      const float3 position = (w * P0 + a * P1) + b * P2;
      float3 normal = (w * N0 + a * N1) + b * N2;  // un-normalised
      if (face == 1) normal = neg3(normal);
      const float3 dir = sample_hemisphere(r.z, r.w, normal);
      const float3 origin = position + normal * g.push;
      emit = true;
      na = make_float4(origin.x, origin.y, origin.z, __uint_as_float(p));
      nb = make_float4(dir.x, dir.y, dir.z, 1e20f);
      d0w = make_float4(dir.x, dir.y, dir.z, fabsf(dot3(normal, dir)));
    }
    ps.thr[p] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(0u));
    ps.acc[p] = make_float4(0, 0, 0, 0);
    ps.dir0w[p] = d0w;
  }
  const uint32_t slot = warp_append(emit, q.count + 0);
  if (emit) {
    q.ray_a[0][slot] = na;
    q.ray_b[0][slot] = nb;
  }
}

// ---- per-probe reduction: 32-lane strided partial sums + xor butterfly (fixed tree, no atomics) ----------
// IlluminationEstimation.cu:81-84: energy = L*|N.dir|; pointEnergy += energy; pointDirection += dir*|energy|.
__global__ void __launch_bounds__(256)
    k_reduce_illum(uint32_t n_tris, uint32_t sc_, PathState ps, float* __restrict__ accum6 /* 6 per probe of batch */) {
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= n_tris) return;
  float tot[6] = {0, 0, 0, 0, 0, 0};
  for (uint32_t face = 0; face < 2; face++) {
    float v[6] = {0, 0, 0, 0, 0, 0};
    const uint32_t base = warp * 2u * sc_ + face * sc_;
    for (uint32_t s = lane; s < sc_; s += 32) {
      const float4 a = ps.acc[base + s], dw = ps.dir0w[base + s];
      const float3 e = f3(a.x, a.y, a.z) * dw.w;
      const float le = length3(e);
      v[0] = v[0] + e.x, v[1] = v[1] + e.y, v[2] = v[2] + e.z;
      v[3] = v[3] + dw.x * le, v[4] = v[4] + dw.y * le, v[5] = v[5] + dw.z * le;
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1)
#pragma unroll
      for (int k = 0; k < 6; k++) v[k] = v[k] + __shfl_xor_sync(0xffffffffu, v[k], off);
#pragma unroll
    for (int k = 0; k < 6; k++) tot[k] = tot[k] + v[k];
  }
  if (lane < 6) {
    float x = tot[0];
#pragma unroll
    for (int k = 1; k < 6; k++)
      if (lane == k) x = tot[k];
    float* o = accum6 + 6 * (size_t)warp + lane;
    *o = *o + x;
  }
}

// IlluminationEstimation.cu:125-128
__global__ void k_finalize_illum(float* accum6, uint64_t n, int faces, int samples, rtf_illum_sampler* probes) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  int nf = faces;
  if (probes) nf = (probes[t].front_face ? 1 : 0) + (probes[t].back_face ? 1 : 0);
  const int cnt = nf * samples;
  if (cnt == 0) return;
  float* v = accum6 + 6 * t;
  const float fn = (float)cnt;
  const float3 e = f3(v[0] / fn, v[1] / fn, v[2] / fn);
  const float3 d = normalize3(f3(v[3], v[4], v[5]));
  v[0] = e.x, v[1] = e.y, v[2] = e.z, v[3] = d.x, v[4] = d.y, v[5] = d.z;
  if (probes) {
    probes[t].energy[0] = e.x, probes[t].energy[1] = e.y, probes[t].energy[2] = e.z;
    probes[t].direction[0] = d.x, probes[t].direction[1] = d.y, probes[t].direction[2] = d.z;
  }
}

// ---- ray sort between wavefronts (option "sort_rays"): key = direction octant (3 b) | Morton code of the origin's cell in a
//      64^3 grid over the scene box (18 b); empty queue slots get bit 21 and sort behind every ray.  The sorted order only
//      changes which lanes trace which rays: per-path state is addressed by the path index the ray carries, so results are
//      bit-identical with the option on or off. ---------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_ray_sort_keys(const float4* __restrict__ ra, const float4* __restrict__ rb,
                                                       const uint32_t* __restrict__ count, uint32_t n_max,
                                                       const float* __restrict__ fb, uint32_t* __restrict__ keys) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_max) return;
  uint32_t key = 1u << 21;
  if (i < *count) {
    const float4 a = ra[i], b = rb[i];
    const float3 lo = f3(fb[0], fb[1], fb[2]), hi = f3(fb[3], fb[4], fb[5]);
    const float3 ext = hi - lo;
    const float3 inv = f3(ext.x > 0 ? 64.0f / ext.x : 0.0f, ext.y > 0 ? 64.0f / ext.y : 0.0f, ext.z > 0 ? 64.0f / ext.z : 0.0f);
    const uint32_t cx = (uint32_t)fminf(fmaxf((a.x - lo.x) * inv.x, 0.0f), 63.0f);
    const uint32_t cy = (uint32_t)fminf(fmaxf((a.y - lo.y) * inv.y, 0.0f), 63.0f);
    const uint32_t cz = (uint32_t)fminf(fmaxf((a.z - lo.z) * inv.z, 0.0f), 63.0f);
    const uint32_t oct = (b.x < 0.0f ? 4u : 0u) | (b.y < 0.0f ? 2u : 0u) | (b.z < 0.0f ? 1u : 0u);
    key = (oct << 18) | (expand10(cx) << 2) | (expand10(cy) << 1) | expand10(cz);
  }
  keys[i] = key;
}
__global__ void __launch_bounds__(256) k_ray_gather(const float4* __restrict__ ra, const float4* __restrict__ rb,
                                                    const uint32_t* __restrict__ perm, const uint32_t* __restrict__ count,
                                                    uint32_t n_max, float4* __restrict__ oa, float4* __restrict__ ob) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_max || i >= *count) return;
  const uint32_t src = perm[i];
  oa[i] = ra[src], ob[i] = rb[src];
}
// Sorts the rays of wavefront `wave` (in queue qi) into queue qi^1, which is free at that moment.
static int sort_wavefront(rtf_ctx* ctx, Queues& q, int qi, int wave, size_t paths) {
  Workspace& w = ctx->ws;
  int rc;
  const uint32_t n = (uint32_t)paths;
  if ((rc = w.sort_kin.reserve(4 * paths)) || (rc = w.sort_keys_alt.reserve(4 * paths)) || (rc = w.sort_tmp.reserve(4 * paths)) ||
      (rc = w.sort_vals.reserve(4 * paths)) || (rc = w.sort_vals_alt.reserve(4 * paths)))
    return rc;
  LAUNCH(ctx, k_ray_sort_keys, div_up(paths, 256), 256, 0, q.ray_a[qi], q.ray_b[qi], q.count + wave, n,
         ctx->ba.scene_bounds.as<float>() + 8, w.sort_kin.as<uint32_t>());
  if ((rc = radix_sort_pairs<uint32_t>(ctx, w.sort_kin.as<uint32_t>(), w.sort_keys_alt.as<uint32_t>(), w.sort_tmp.as<uint32_t>(),
                                       nullptr, w.sort_vals.as<uint32_t>(), w.sort_vals_alt.as<uint32_t>(), n, 0, 22,
                                       ctx->ba.sort_hist, ctx->ba.sort_status)))
    return rc;
  LAUNCH(ctx, k_ray_gather, div_up(paths, 256), 256, 0, q.ray_a[qi], q.ray_b[qi], w.sort_vals.as<uint32_t>(), q.count + wave, n,
         q.ray_a[qi ^ 1], q.ray_b[qi ^ 1]);
  return RTF_OK;
}

static size_t batch_capacity(const rtf_ctx* ctx) { return ctx->batch_paths < 1024 ? 1024 : ctx->batch_paths; }

static int ensure_workspace(rtf_ctx* ctx, size_t cap, bool camera) {
  Workspace& w = ctx->ws;
  int rc;
  for (int k = 0; k < 2; k++) {
    if ((rc = w.ray_a[k].reserve(16 * cap))) return rc;
    if ((rc = w.ray_b[k].reserve(16 * cap))) return rc;
  }
  if ((rc = w.hits.reserve(16 * cap))) return rc;
  if ((rc = w.counters.reserve(4096))) return rc;
  if ((rc = w.thr.reserve(16 * cap))) return rc;
  if ((rc = w.acc.reserve(16 * cap))) return rc;
  if ((rc = w.dir0w.reserve(16 * cap))) return rc;
  if ((rc = w.first_n.reserve(16 * cap))) return rc;  // doubles as `aux` outside camera mode
  if (camera) {
    if ((rc = w.first_a.reserve(16 * cap))) return rc;
    if ((rc = w.first_p.reserve(16 * cap))) return rc;
    if ((rc = w.sort_keys.reserve(16 * cap))) return rc;  // aux for camera + single light
  }
  if (ctx->scene.has_sss) {
    if ((rc = w.pend_o.reserve(16 * cap))) return rc;
    if ((rc = w.pend_d.reserve(16 * cap))) return rc;
    if ((rc = w.pend_t.reserve(16 * cap))) return rc;
  }
  if ((rc = ctx->d_stats.reserve(64))) return rc;
  w.capacity = cap;
  return RTF_OK;
}

static void bind(rtf_ctx* ctx, Queues& q, PathState& ps, bool camera) {
  Workspace& w = ctx->ws;
  for (int k = 0; k < 2; k++) {
    q.ray_a[k] = w.ray_a[k].as<float4>(), q.ray_b[k] = w.ray_b[k].as<float4>();
  }
  q.hits = w.hits.as<float4>();
  q.count = w.counters.as<uint32_t>() + 16;
  q.work = w.counters.as<uint32_t>() + 16 + 64;
  ps.thr = w.thr.as<float4>(), ps.acc = w.acc.as<float4>(), ps.dir0w = w.dir0w.as<float4>();
  ps.first_n = w.first_n.as<float4>(), ps.first_a = w.first_a.as<float4>(), ps.first_p = w.first_p.as<float4>();
  ps.aux = camera ? w.sort_keys.as<float4>() : w.first_n.as<float4>();
  ps.pend_o = w.pend_o.as<float4>(), ps.pend_d = w.pend_d.as<float4>(), ps.pend_t = w.pend_t.as<float4>();
}

// Runs the trace/shade loop for the wavefronts of one batch (no host synchronisation inside).
static int run_wavefronts(rtf_ctx* ctx, const rtf_env* env, int bounces, uint32_t seed, const PathMap& pm, size_t paths,
                          Queues& q, PathState& ps, bool camera, uint32_t* lcg = nullptr) {
  // BSSRDF: the subsurface path and the waiting bounce path share one hit counter, so together they add at most one ray
  // to the longest chain (primary + `bounces` hits + the resumed ray).
  const bool sss = ctx->scene.has_sss && env->lighting_type != RTF_ENV_SINGLE_LIGHT_SOURCE;
  const int waves = env->lighting_type == RTF_ENV_SINGLE_LIGHT_SOURCE ? 2 : bounces + 1 + (sss ? 1 : 0);
  if (sss) RTF_CUDA(cudaMemsetAsync(ps.pend_o, 0, 16 * paths, ctx->stream));
  if (waves + 1 > 60) {
    set_error("too many bounces");
    return RTF_ERR_INVALID;
  }
  ShadeParams sp{bounces, seed, camera ? 1 : 0, ctx->brdf, lcg};
  unsigned long long* stats = ctx->d_stats.as<unsigned long long>();
  const bool alpha_kernel = ctx->scene.has_alpha || lcg != nullptr;  // the reference stream draws at every any-hit candidate
  const unsigned tgrid = alpha_kernel ? trace_grid(ctx, TK_QUEUE_ALPHA, k_trace_queue<true>, paths)
                                      : trace_grid(ctx, TK_QUEUE, k_trace_queue<false>, paths);
  int qi = 0;
  for (int w = 0; w < waves; w++) {
    const float tmin = (camera && w == 0) ? 0.0f : 1e-3f;
    if (ctx->sort_rays && w >= 1) {  // bounce rays leave the surfaces in random directions: restore some coherence
      int rc;
      ctx->prof.begin(PROF_SORT, ctx->stream);
      rc = sort_wavefront(ctx, q, qi, w, paths);
      ctx->prof.end(ctx->stream);
      if (rc) return rc;
      qi ^= 1;
    }
    if (alpha_kernel)
      LAUNCH_P(ctx, PROF_TRACE, k_trace_queue<true>, tgrid, kTraceBlock, 0, ctx->scene, q, qi, w, tmin, pm, seed, stats, lcg);
    else
      LAUNCH_P(ctx, PROF_TRACE, k_trace_queue<false>, tgrid, kTraceBlock, 0, ctx->scene, q, qi, w, tmin, pm, seed, stats, lcg);
    if (sss && ctx->scene.has_btf)
      LAUNCH_P(ctx, PROF_SHADE, (k_shade<true, true>), div_up(paths, 256), 256, 0, ctx->scene, *env, sp, pm, q, qi, w, ps);
    else if (sss)
      LAUNCH_P(ctx, PROF_SHADE, (k_shade<true, false>), div_up(paths, 256), 256, 0, ctx->scene, *env, sp, pm, q, qi, w, ps);
    else if (ctx->scene.has_btf)
      LAUNCH_P(ctx, PROF_SHADE, (k_shade<false, true>), div_up(paths, 256), 256, 0, ctx->scene, *env, sp, pm, q, qi, w, ps);
    else
      LAUNCH_P(ctx, PROF_SHADE, (k_shade<false, false>), div_up(paths, 256), 256, 0, ctx->scene, *env, sp, pm, q, qi, w, ps);
    qi ^= 1;
  }
  return RTF_OK;
}

int illum_mesh_device(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push, int front,
                      int back, uint64_t tri_begin, uint64_t tri_end, int s0, int s1, float* d_accum6,
                      const rtf_illum_sampler* d_probes, uint64_t probe_base) {
  if (!ctx->ready) {
    set_error("no acceleration structure (call rtf_commit)");
    return RTF_ERR_NOT_READY;
  }
  if (env->environmental_map_id != 0 && env->lighting_type == RTF_ENV_SCENE && ctx->scene.env_faces == nullptr) {
    set_error("environmental_map_id != 0 but no cube map is registered (rtf_set_environment_map, then rtf_commit)");
    return RTF_ERR_INVALID;
  }
  if (tri_end <= tri_begin || s1 <= s0) return RTF_OK;
  const uint32_t sc_ = (uint32_t)(s1 - s0);
  const size_t per_probe = 2 * (size_t)sc_;
  size_t cap = batch_capacity(ctx);
  if (cap < per_probe) cap = per_probe;
  int rc;
  if ((rc = ensure_workspace(ctx, cap, false))) return rc;
  Queues q;
  PathState ps;
  bind(ctx, q, ps, false);
  const uint64_t tris_per_batch = cap / per_probe;
  if (!ctx->stats_hold) RTF_CUDA(cudaMemsetAsync(ctx->d_stats.p, 0, 16, ctx->stream));
  ctx->stats_cached = false;
  uint64_t paths_total = ctx->stats_hold ? ctx->last_paths : 0;
  for (uint64_t t0 = tri_begin; t0 < tri_end; t0 += tris_per_batch) {
    const uint32_t nt = (uint32_t)((tri_end - t0) < tris_per_batch ? (tri_end - t0) : tris_per_batch);
    const size_t paths = (size_t)nt * per_probe;
    RTF_CUDA(cudaMemsetAsync(ctx->ws.counters.p, 0, 4096, ctx->stream));
    IllumGen g{t0, nt, sc_, (uint32_t)s0, (uint32_t)rp->samples, seed, push, front, back, d_probes, probe_base};
    LAUNCH_P(ctx, PROF_GEN, k_gen_illum, div_up(paths, 256), 256, 0, ctx->scene, g, q, ps);
    PathMap pm{0, 0ull, t0, (uint32_t)per_probe, sc_, (uint32_t)s0, (uint32_t)rp->samples, PixelMap{0ull, 0u, 0u, 0u, 0ull, 0u}, 0u};
    if ((rc = run_wavefronts(ctx, env, rp->bounces, seed, pm, paths, q, ps, false))) return rc;
    LAUNCH_P(ctx, PROF_REDUCE, k_reduce_illum, div_up((size_t)nt * 32, 256), 256, 0, nt, sc_, ps, d_accum6 + 6 * (t0 - tri_begin));
    paths_total += paths;
  }
  RTF_CUDA(cudaGetLastError());
  ctx->last_paths = paths_total;
  return RTF_OK;
}

// ---- option "rng" = RTF_RNG_REFERENCE_LCG: __raygen__IE with the reference's own random stream ---------------------
// One LCG per probe, seeded (probe index, seed) (IE.cu:47), consumed in the reference's order: front-face samples one
// after the other, then the back face; per sample the two barycentric draws, the two of RandomSampleHemisphere, and
// whatever the path draws on its way (one per any-hit candidate, two per bounce).  The length of a path decides where the
// next sample's draws start, so the samples of a probe run in sequence: one single-path wavefront set per (face, sample)
// over all probes of the batch.  pointEnergy / pointDirection accumulate in sample order as IE.cu:81-84 does.
__global__ void k_lcg_init(uint32_t n, uint64_t first_probe, uint32_t seed, uint32_t* __restrict__ lcg) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) lcg[i] = lcg_seed((uint32_t)(first_probe + i), seed);
}
__global__ void __launch_bounds__(256) k_gen_illum_lcg(const rtf_illum_sampler* __restrict__ probes, uint32_t n, int face, float push,
                                                       uint32_t* __restrict__ lcg, Queues q, PathState ps) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  bool emit = false;
  float4 na = make_float4(0, 0, 0, 0), nb = make_float4(0, 0, 0, 0);
  if (p < n) {
    const rtf_illum_sampler& pr = probes[p];
    float4 d0w = make_float4(0, 0, 0, 0);
    if (face == 0 ? pr.front_face != 0 : pr.back_face != 0) {
      const float3 P0 = f3(pr.a.position[0], pr.a.position[1], pr.a.position[2]);
      const float3 P1 = f3(pr.b.position[0], pr.b.position[1], pr.b.position[2]);
      const float3 P2 = f3(pr.c.position[0], pr.c.position[1], pr.c.position[2]);
      const float3 N0 = f3(pr.a.normal[0], pr.a.normal[1], pr.a.normal[2]);
      const float3 N1 = f3(pr.b.normal[0], pr.b.normal[1], pr.b.normal[2]);
      const float3 N2 = f3(pr.c.normal[0], pr.c.normal[1], pr.c.normal[2]);
      uint32_t st = lcg[p];
      const float a = lcg_next(st), b = lcg_next(st), u0 = lcg_next(st), u1 = lcg_next(st);
      lcg[p] = st;
      const float w = (1.0f - a) - b;
      const float3 position = (w * P0 + a * P1) + b * P2;
      float3 normal = (w * N0 + a * N1) + b * N2;
      if (face == 1) normal = neg3(normal);
      const float3 dir = sample_hemisphere(u0, u1, normal);
      const float3 origin = position + normal * push;
      emit = true;
      na = make_float4(origin.x, origin.y, origin.z, __uint_as_float(p));
      nb = make_float4(dir.x, dir.y, dir.z, 1e20f);
      d0w = make_float4(dir.x, dir.y, dir.z, fabsf(dot3(normal, dir)));
    }
    ps.thr[p] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(0u));
    ps.acc[p] = make_float4(0, 0, 0, 0);
    ps.dir0w[p] = d0w;
  }
  const uint32_t slot = warp_append(emit, q.count + 0);
  if (emit) {
    q.ray_a[0][slot] = na;
    q.ray_b[0][slot] = nb;
  }
}
__global__ void k_accum_illum_lcg(const rtf_illum_sampler* __restrict__ probes, uint32_t n, int face, PathState ps,
                                  float* __restrict__ accum6) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  if (!(face == 0 ? probes[p].front_face != 0 : probes[p].back_face != 0)) return;
  const float4 a = ps.acc[p], dw = ps.dir0w[p];
  const float3 e = f3(a.x, a.y, a.z) * dw.w;
  const float le = length3(e);
  float* o = accum6 + 6 * (size_t)p;
  o[0] = o[0] + e.x, o[1] = o[1] + e.y, o[2] = o[2] + e.z;
  o[3] = o[3] + dw.x * le, o[4] = o[4] + dw.y * le, o[5] = o[5] + dw.z * le;
}
// d_probes: the n probes of this call on the device (index 0 = probe `first_probe` of the caller's array); d_accum6: n x 6
// zeroed floats.  Subsurface materials are refused: their probe collects records in an order of its own.
int illum_probes_lcg_device(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push,
                            const rtf_illum_sampler* d_probes, uint64_t n, uint64_t first_probe, float* d_accum6) {
  if (!ctx->ready) {
    set_error("no acceleration structure (call rtf_commit)");
    return RTF_ERR_NOT_READY;
  }
  if (ctx->scene.has_sss) {
    set_error("option rng = reference: scenes with a subsurface material are not supported (order-dependent probe)");
    return RTF_ERR_INVALID;
  }
  if (env->environmental_map_id != 0 && env->lighting_type == RTF_ENV_SCENE && ctx->scene.env_faces == nullptr) {
    set_error("environmental_map_id != 0 but no cube map is registered (rtf_set_environment_map, then rtf_commit)");
    return RTF_ERR_INVALID;
  }
  if (n == 0 || rp->samples <= 0) return RTF_OK;
  const size_t cap = batch_capacity(ctx);
  int rc;
  if ((rc = ensure_workspace(ctx, cap, false))) return rc;
  if ((rc = ctx->ws.lcg.reserve(4 * cap))) return rc;
  uint32_t* lcg = ctx->ws.lcg.as<uint32_t>();
  Queues q;
  PathState ps;
  bind(ctx, q, ps, false);
  if (!ctx->stats_hold) RTF_CUDA(cudaMemsetAsync(ctx->d_stats.p, 0, 16, ctx->stream));
  ctx->stats_cached = false;
  uint64_t paths_total = ctx->stats_hold ? ctx->last_paths : 0;
  const PathMap pm{0, 0ull, 0ull, 1u, 1u, 0u, 1u, PixelMap{0ull, 0u, 0u, 0u, 0ull, 0u}, 0u};  // unused by the LCG draws
  for (uint64_t b0 = 0; b0 < n; b0 += cap) {
    const uint32_t nb = (uint32_t)((n - b0) < cap ? (n - b0) : cap);
    LAUNCH(ctx, k_lcg_init, div_up(nb, 256), 256, 0, nb, first_probe + b0, seed, lcg);
    for (int face = 0; face < 2; face++)
      for (int s = 0; s < rp->samples; s++) {
        RTF_CUDA(cudaMemsetAsync(ctx->ws.counters.p, 0, 4096, ctx->stream));
        LAUNCH_P(ctx, PROF_GEN, k_gen_illum_lcg, div_up(nb, 256), 256, 0, d_probes + b0, nb, face, push, lcg, q, ps);
        if ((rc = run_wavefronts(ctx, env, rp->bounces, seed, pm, nb, q, ps, false, lcg))) return rc;
        LAUNCH_P(ctx, PROF_REDUCE, k_accum_illum_lcg, div_up(nb, 256), 256, 0, d_probes + b0, nb, face, ps, d_accum6 + 6 * b0);
      }
    paths_total += 2ull * (uint64_t)rp->samples * nb;
  }
  RTF_CUDA(cudaGetLastError());
  ctx->last_paths = paths_total;
  return RTF_OK;
}

int illum_finalize_device(rtf_ctx* ctx, float* d_accum6, uint64_t n, int faces, int samples, rtf_illum_sampler* d_probes) {
  if (n == 0) return RTF_OK;
  LAUNCH_P(ctx, PROF_REDUCE, k_finalize_illum, div_up(n, 256), 256, 0, d_accum6, n, faces, samples, d_probes);
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}

// ============================ LiDAR =================================================================
// PointCloudScanner::Scan ray grid (PointCloudScanner.cpp:70-92) generated in the kernel, traced
// (tmin 1e-3, tmax 1e20, any-hit disabled: PointCloudScanning.cu:58-64) and written as SoA.
struct SrcGrid {
  static constexpr bool kTma = false, kAnyHit = false, kAlpha = false;
  rtf_scan_grid g;
  uint32_t row;
  int column_start, row_start;
  float3 dir;
  uint64_t ray_begin;
  uint32_t n;
  float* t_out;
  uint32_t* prim_out;
  uint32_t* inst_out;
  const uint32_t* tri_inst;
  // Interleaved shards (multi-GPU, il_stride = number of ranks, 0 = one contiguous range): the launch's ray j is ray
  // (j mod B) of block (j / B) * il_stride + il_rank of the grid, B = 2^il_shift — contiguous ranges of a plant field
  // differ several-fold in cost (bare ground at the rim, canopy in the middle).  out_global: results go to the ray's
  // own index of full-size arrays instead of the launch-local one.
  uint32_t il_shift, il_stride, il_rank, out_global;
  RTF_D uint64_t ray_index(uint32_t j) const {
    if (il_stride == 0u) return ray_begin + j;
    const uint64_t blk = (uint64_t)(j >> il_shift) * il_stride + il_rank;
    return (blk << il_shift) | (uint64_t)(j & ((1u << il_shift) - 1u));
  }
  RTF_D void fetch(const unsigned char*, uint32_t base, uint32_t slot, float4& a, float4& b, float& tmin, uint32_t& tag) const {
    const uint64_t i = ray_index(base + slot);
    const int ci = (int)(i / row), ri = (int)(i % row);
    const float3 center = f3(g.center[0], g.center[1], g.center[2]);
    const float3 left = f3(g.left[0], g.left[1], g.left[2]), up = f3(g.up[0], g.up[1], g.up[2]);
    const float3 pos = (center + (left * (float)(column_start + ci)) * g.distance[0]) + (up * (float)(row_start + ri)) * g.distance[1];
    a = make_float4(pos.x, pos.y, pos.z, 0.0f);
    b = make_float4(dir.x, dir.y, dir.z, 1e20f);
    tmin = 1e-3f, tag = 0u;
  }
  RTF_D bool any_hit(uint32_t) const { return false; }
  RTF_D bool alpha_pass(uint32_t, uint32_t, uint32_t, float, float) const { return true; }
  RTF_D void store(uint32_t j, uint32_t, const HitRec& h, uint32_t, uint32_t) const {
    const uint64_t k = out_global ? ray_index(j) : (uint64_t)j;
    if (t_out) t_out[k] = h.t;
    if (prim_out) prim_out[k] = h.prim;
    if (inst_out) inst_out[k] = h.prim == 0xFFFFFFFFu ? 0xFFFFFFFFu : __ldg(tri_inst + h.prim);
  }
};
#ifndef RTF_SCAN_MIN_BLOCKS
#define RTF_SCAN_MIN_BLOCKS 8
#endif
__global__ void __launch_bounds__(kTraceBlock, RTF_SCAN_MIN_BLOCKS) k_scan_grid(DevScene sc, SrcGrid src, uint32_t* work) {
  __shared__ TraceSmem sm;
  trace_engine<false>(sc.nodes, sc.tris, src, work, sm);
}
// CudaModule::SamplePointCloud's AoS samples (112 B in-out): read in the converged prepare phase, written back with
// the full HitInfo of __closesthit__PCS_R (PointCloudScanning.cu:15-28).
struct SrcSamples {
  static constexpr bool kTma = false, kAnyHit = false, kAlpha = false;
  DevScene sc;
  rtf_pc_sample* samples;
  uint32_t* prim_out;
  uint32_t n;
  RTF_D void fetch(const unsigned char*, uint32_t base, uint32_t slot, float4& a, float4& b, float& tmin, uint32_t& tag) const {
    const rtf_pc_sample& s = samples[base + slot];
    a = make_float4(s.start[0], s.start[1], s.start[2], 0.0f);
    b = make_float4(s.direction[0], s.direction[1], s.direction[2], 1e20f);
    tmin = 1e-3f, tag = 0u;
  }
  RTF_D bool any_hit(uint32_t) const { return false; }
  RTF_D bool alpha_pass(uint32_t, uint32_t, uint32_t, float, float) const { return true; }
  RTF_D void store(uint32_t k, uint32_t, const HitRec& h, uint32_t, uint32_t) const {
    rtf_pc_sample& s = samples[k];
    if (prim_out) prim_out[k] = h.prim;
    rtf_hit_info hi;
    if (h.prim == 0xFFFFFFFFu) {  // __miss__PCS_R (PointCloudScanning.cu:36-40) + HitInfo() defaults
      s.hit = 0, s.handle = 0;
      for (int c = 0; c < 3; c++) hi.position[c] = hi.normal[c] = hi.tangent[c] = 0.0f;
      for (int c = 0; c < 4; c++) hi.color[c] = 1.0f, hi.data[c] = 0.0f;
      hi.tex_coord[0] = hi.tex_coord[1] = 0.0f;
    } else {
      uint64_t handle;
      full_hit_info(sc, h.prim, h.u, h.v, f3(s.direction[0], s.direction[1], s.direction[2]), hi, handle);
      s.hit = 1, s.handle = handle;
    }
    s.hit_info = hi;
  }
};
__global__ void __launch_bounds__(kTraceBlock) k_scan_samples(DevScene sc, SrcSamples src, uint32_t* work) {
  __shared__ TraceSmem sm;
  trace_engine<false>(sc.nodes, sc.tris, src, work, sm);
}

int scan_grid_device(rtf_ctx* ctx, const rtf_scan_grid* grid, uint64_t ray_begin, uint64_t ray_end, float* d_t,
                     uint32_t* d_prim, uint32_t* d_inst) {
  if (!ctx->ready) {
    set_error("no acceleration structure (call rtf_commit)");
    return RTF_ERR_NOT_READY;
  }
  uint32_t column = 0, row = 0;
  int rc;
  if ((rc = rtf_scan_grid_dims(grid, &column, &row))) return rc;
  if (ray_end > (uint64_t)column * row || ray_end < ray_begin) {
    set_error("ray range outside the scan grid");
    return RTF_ERR_INVALID;
  }
  if (ray_end == ray_begin) return RTF_OK;
  if ((rc = ctx->ws.counters.reserve(4096))) return rc;
  uint32_t* work = ctx->ws.counters.as<uint32_t>();
  const uint64_t chunk = 1ull << 30;
  for (uint64_t off = ray_begin; off < ray_end; off += chunk) {
    const uint32_t m = (uint32_t)((ray_end - off) < chunk ? (ray_end - off) : chunk);
    RTF_CUDA(cudaMemsetAsync(work, 0, 4, ctx->stream));
    const uint64_t o = off - ray_begin;
    SrcGrid src{*grid, row, -(int)(column / 2), -(int)(row / 2),
                normalize3(f3(grid->direction[0], grid->direction[1], grid->direction[2])), off, m,
                d_t ? d_t + o : nullptr, d_prim ? d_prim + o : nullptr, d_inst ? d_inst + o : nullptr, ctx->scene.tri_inst,
                0u, 0u, 0u, 0u};
    LAUNCH_P(ctx, PROF_TRACE, k_scan_grid, trace_grid(ctx, TK_GRID, k_scan_grid, m), kTraceBlock, 0, ctx->scene, src, work);
  }
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}
// Block layout of the interleaved LiDAR shards: about 64 blocks per rank, 2^12 .. 2^24 rays each (at 16 per rank the ranks
// of C4 on 8 GPUs still differed by 15 %: 4.06 .. 4.73 ms).
void scan_interleave_layout(uint64_t total, int nranks, int rank, uint32_t* shift, uint64_t* n_blocks, uint64_t* n_local) {
  uint32_t sh = 12;
  while (sh < 24 && (total >> sh) > 64ull * (uint64_t)nranks) sh++;
  const uint64_t B = 1ull << sh, nb = (total + B - 1) / B;
  uint64_t mine = (uint64_t)rank < nb ? (nb - (uint64_t)rank + (uint64_t)nranks - 1) / (uint64_t)nranks : 0;
  uint64_t rays = mine * B;
  if (mine && (nb - 1) % (uint64_t)nranks == (uint64_t)rank) rays -= nb * B - total;  // the grid's last block is partial
  *shift = sh, *n_blocks = nb, *n_local = rays;
}
// The rank's interleaved blocks of the grid in one launch.  global_out: d_* are full-size arrays indexed by ray;
// otherwise they hold the rank's rays densely, block after block.
int scan_grid_interleaved_device(rtf_ctx* ctx, const rtf_scan_grid* grid, int nranks, int rank, bool global_out, float* d_t,
                                 uint32_t* d_prim, uint32_t* d_inst) {
  if (!ctx->ready) {
    set_error("no acceleration structure (call rtf_commit)");
    return RTF_ERR_NOT_READY;
  }
  uint32_t column = 0, row = 0, shift = 0;
  int rc;
  if ((rc = rtf_scan_grid_dims(grid, &column, &row))) return rc;
  uint64_t nb = 0, n_local = 0;
  scan_interleave_layout((uint64_t)column * row, nranks, rank, &shift, &nb, &n_local);
  if (n_local == 0) return RTF_OK;
  if (n_local >= (1ull << 32)) {
    set_error("interleaved scan: more than 2^32 rays per rank");
    return RTF_ERR_LIMIT;
  }
  if ((rc = ctx->ws.counters.reserve(4096))) return rc;
  uint32_t* work = ctx->ws.counters.as<uint32_t>();
  RTF_CUDA(cudaMemsetAsync(work, 0, 4, ctx->stream));
  SrcGrid src{*grid, row, -(int)(column / 2), -(int)(row / 2),
              normalize3(f3(grid->direction[0], grid->direction[1], grid->direction[2])), 0, (uint32_t)n_local,
              d_t, d_prim, d_inst, ctx->scene.tri_inst, shift, (uint32_t)nranks, (uint32_t)rank, global_out ? 1u : 0u};
  LAUNCH_P(ctx, PROF_TRACE, k_scan_grid, trace_grid(ctx, TK_GRID, k_scan_grid, n_local), kTraceBlock, 0, ctx->scene, src, work);
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}
int scan_samples_device(rtf_ctx* ctx, rtf_pc_sample* d_samples, size_t n, uint32_t* d_prim) {
  if (!ctx->ready) {
    set_error("no acceleration structure (call rtf_commit)");
    return RTF_ERR_NOT_READY;
  }
  if (n == 0) return RTF_OK;
  int rc;
  if ((rc = ctx->ws.counters.reserve(4096))) return rc;
  uint32_t* work = ctx->ws.counters.as<uint32_t>();
  RTF_CUDA(cudaMemsetAsync(work, 0, 4, ctx->stream));
  SrcSamples src{ctx->scene, d_samples, d_prim, (uint32_t)n};
  LAUNCH_P(ctx, PROF_TRACE, k_scan_samples, trace_grid(ctx, TK_SAMPLES, k_scan_samples, n), kTraceBlock, 0, ctx->scene, src, work);
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}

// ============================ camera ================================================================
// __raygen__CR path start, CameraRendering.cu:63-85.
__global__ void __launch_bounds__(256)
    k_gen_camera(rtf_camera_props cam, PixelMap pxm, uint32_t n_pixels, uint32_t spp, Queues q, PathState ps) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t total = (uint64_t)n_pixels * spp;
  bool emit = false;
  float4 na = make_float4(0, 0, 0, 0), nb = make_float4(0, 0, 0, 0);
  if (p < total) {
    const uint32_t lk = p / spp;
    const uint64_t pixel = pixel_of(pxm, lk);
    const uint32_t ix = (uint32_t)(pixel % (uint64_t)cam.width), iy = (uint32_t)(pixel / (uint64_t)cam.width);
    const uint64_t unit = pixel * spp + (p - lk * spp);
    const float4 r = path_rng((uint32_t)cam.frame_id, STREAM_CAMERA, unit, 0u, 0u);
    const float halfX = cam.width / 2.0f, halfY = cam.height / 2.0f;
    const float sx = (((float)ix + r.x) - halfX) / halfX;
    const float sy = (((float)iy + r.y) - halfY) / halfY;
    const float* m = cam.inverse_projection_view;
    float st[4], en[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      st[k] = (m[k] * sx + m[4 + k] * sy) + (m[8 + k] * -1.0f + m[12 + k] * 1.0f);
      en[k] = (m[k] * sx + m[4 + k] * sy) + (m[8 + k] * 1.0f + m[12 + k] * 1.0f);
    }
    const float3 start = f3(st[0] / st[3], st[1] / st[3], st[2] / st[3]);
    const float3 end = f3(en[0] / en[3], en[1] / en[3], en[2] / en[3]);
    const float3 pdir = normalize3(end - start);
    const float3 conv = start + pdir * cam.focal_length;
    float sn, cs;
    sincos2pi(r.z, sn, cs);
    const float3 hor = f3(cam.horizontal[0], cam.horizontal[1], cam.horizontal[2]);
    const float3 ver = f3(cam.vertical[0], cam.vertical[1], cam.vertical[2]);
    const float3 ap = start + cam.aperture * (hor * sn + ver * cs);
    const float3 dir = normalize3(conv - ap);
    emit = true;
    na = make_float4(ap.x, ap.y, ap.z, __uint_as_float(p));  // tmin 0 (CameraRendering.cu:89) is the wavefront's
    nb = make_float4(dir.x, dir.y, dir.z, 1e20f);
    ps.thr[p] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(0u));
    ps.acc[p] = make_float4(0, 0, 0, 0);
    ps.first_n[p] = make_float4(0, 0, 0, 0);
    ps.first_a[p] = make_float4(0, 0, 0, 0);
    ps.first_p[p] = make_float4(0, 0, 0, 0);
  }
  const uint32_t slot = warp_append(emit, q.count + 0);
  if (emit) {
    q.ray_a[0][slot] = na;
    q.ray_b[0][slot] = nb;
  }
}
// CameraRendering.cu:112-146: progressive accumulation in gamma space and the three output writes of a pixel.
__device__ __forceinline__ void camera_write(const rtf_camera_props& cam, uint64_t pixel, float3 pc, float3 pn, float3 pa, float3 pp,
                                             float4* __restrict__ color, float4* __restrict__ normal, float4* __restrict__ albedo) {
  if (cam.accumulate && cam.frame_id > 1) {
    const float4 prev = color[pixel];
    const float3 accd = f3(powf(prev.x, cam.gamma), powf(prev.y, cam.gamma), powf(prev.z, cam.gamma));
    const float fid = (float)cam.frame_id;
    pc = pc + fid * accd;
    const float den = (float)(cam.frame_id + 1);
    pc = f3(pc.x / den, pc.y / den, pc.z / den);
  }
  const float ig = (float)(1.0 / (double)cam.gamma);
  color[pixel] = make_float4(powf(pc.x, ig), powf(pc.y, ig), powf(pc.z, ig), 1.0f);
  if (cam.output_type == RTF_OUT_DEPTH) {
    const float3 from = f3(cam.from[0], cam.from[1], cam.from[2]);
    const float dd = clampf(length3(from - pp) / cam.max_distance, 0.0f, 1.0f);
    albedo[pixel] = make_float4(dd, dd, dd, 1.0f);
  } else {
    albedo[pixel] = make_float4(pa.x, pa.y, pa.z, 1.0f);
  }
  normal[pixel] = make_float4(pn.x, pn.y, pn.z, 1.0f);
}
// CameraRendering.cu:101-146: per-pixel sums in sample order, progressive accumulation in gamma space.
__global__ void __launch_bounds__(256)
    k_resolve_camera(rtf_camera_props cam, PixelMap pxm, uint32_t n_pixels, uint32_t spp, PathState ps,
                     float4* __restrict__ color, float4* __restrict__ normal, float4* __restrict__ albedo) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_pixels) return;
  const uint64_t pixel = pixel_of(pxm, k);
  const float nsf = (float)spp;
  float3 pc = f3(0, 0, 0), pn = f3(0, 0, 0), pa = f3(0, 0, 0), pp = f3(0, 0, 0);
  for (uint32_t s = 0; s < spp; s++) {
    const size_t p = (size_t)k * spp + s;
    const float4 e = ps.acc[p], n4 = ps.first_n[p], a4 = ps.first_a[p], p4 = ps.first_p[p];
    float3 pos = f3(p4.x, p4.y, p4.z);
    if (p4.w == 0.0f) pos = s == 0 ? f3(999999.0f, 999999.0f, 999999.0f) : f3(0, 0, 0);  // CameraRendering.cu:48,108
    pc = pc + f3(e.x / nsf, e.y / nsf, e.z / nsf);
    pn = pn + f3(n4.x / nsf, n4.y / nsf, n4.z / nsf);
    pa = pa + f3(a4.x / nsf, a4.y / nsf, a4.z / nsf);
    pp = pp + f3(pos.x / nsf, pos.y / nsf, pos.z / nsf);
  }
  camera_write(cam, pixel, pc, pn, pa, pp, color, normal, albedo);
}
// ---- option "rng" = RTF_RNG_REFERENCE_LCG for the camera: __raygen__CR with the reference's own random stream --------
// One LCG per pixel seeded (pixel index computed in float as CR.cu:42-44 does, frame id); per sample the two jitters — the
// second drawn first: CR.cu:60-61 draws both inside one constructor call whose argument order the language leaves open, and
// the host build of the reference's program that made the fixture takes the second first — then the aperture angle, then
// the path's draws.  Samples of a pixel run in sequence; the four pixel sums accumulate in sample order (CR.cu:101-110).
__global__ void k_lcg_init_camera(rtf_camera_props cam, uint64_t first_pixel, uint32_t n, uint32_t* __restrict__ lcg,
                                  float4* __restrict__ sums) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const uint64_t pixel = first_pixel + k;
  const uint32_t ix = (uint32_t)(pixel % (uint64_t)cam.width), iy = (uint32_t)(pixel / (uint64_t)cam.width);
  lcg[k] = lcg_seed((uint32_t)((float)ix + (float)cam.width * (float)iy), (uint32_t)cam.frame_id);
#pragma unroll
  for (int j = 0; j < 4; j++) sums[4 * (size_t)k + j] = make_float4(0, 0, 0, 0);
}
__global__ void __launch_bounds__(256) k_gen_camera_lcg(rtf_camera_props cam, uint64_t first_pixel, uint32_t n,
                                                        uint32_t* __restrict__ lcg, Queues q, PathState ps) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  bool emit = false;
  float4 na = make_float4(0, 0, 0, 0), nb = make_float4(0, 0, 0, 0);
  if (p < n) {
    const uint64_t pixel = first_pixel + p;
    const uint32_t ix = (uint32_t)(pixel % (uint64_t)cam.width), iy = (uint32_t)(pixel / (uint64_t)cam.width);
    uint32_t st_ = lcg[p];
    const float r1 = lcg_next(st_), r0 = lcg_next(st_), r2 = lcg_next(st_);
    lcg[p] = st_;
    const float halfX = cam.width / 2.0f, halfY = cam.height / 2.0f;
    const float sx = (((float)ix + r0) - halfX) / halfX;
    const float sy = (((float)iy + r1) - halfY) / halfY;
    const float* m = cam.inverse_projection_view;
    float st[4], en[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      st[k] = (m[k] * sx + m[4 + k] * sy) + (m[8 + k] * -1.0f + m[12 + k] * 1.0f);
      en[k] = (m[k] * sx + m[4 + k] * sy) + (m[8 + k] * 1.0f + m[12 + k] * 1.0f);
    }
    const float3 start = f3(st[0] / st[3], st[1] / st[3], st[2] / st[3]);
    const float3 end = f3(en[0] / en[3], en[1] / en[3], en[2] / en[3]);
    const float3 pdir = normalize3(end - start);
    const float3 conv = start + pdir * cam.focal_length;
    float sn, cs;
    sincos2pi(r2, sn, cs);
    const float3 hor = f3(cam.horizontal[0], cam.horizontal[1], cam.horizontal[2]);
    const float3 ver = f3(cam.vertical[0], cam.vertical[1], cam.vertical[2]);
    const float3 ap = start + cam.aperture * (hor * sn + ver * cs);
    const float3 dir = normalize3(conv - ap);
    emit = true;
    na = make_float4(ap.x, ap.y, ap.z, __uint_as_float(p));
    nb = make_float4(dir.x, dir.y, dir.z, 1e20f);
    ps.thr[p] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(0u));
    ps.acc[p] = make_float4(0, 0, 0, 0);
    ps.first_n[p] = make_float4(0, 0, 0, 0);
    ps.first_a[p] = make_float4(0, 0, 0, 0);
    ps.first_p[p] = make_float4(0, 0, 0, 0);
  }
  const uint32_t slot = warp_append(emit, q.count + 0);
  if (emit) {
    q.ray_a[0][slot] = na;
    q.ray_b[0][slot] = nb;
  }
}
__global__ void k_accum_camera_lcg(uint32_t n, uint32_t sample, uint32_t spp, PathState ps, float4* __restrict__ sums) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const float nsf = (float)spp;
  const float4 e = ps.acc[p], n4 = ps.first_n[p], a4 = ps.first_a[p], p4 = ps.first_p[p];
  float3 pos = f3(p4.x, p4.y, p4.z);
  if (p4.w == 0.0f) pos = sample == 0u ? f3(999999.0f, 999999.0f, 999999.0f) : f3(0, 0, 0);  // CameraRendering.cu:48,108
  float4* s4 = sums + 4 * (size_t)p;
  const float4 c = s4[0], nn = s4[1], aa = s4[2], pp = s4[3];
  s4[0] = make_float4(c.x + e.x / nsf, c.y + e.y / nsf, c.z + e.z / nsf, 0.0f);
  s4[1] = make_float4(nn.x + n4.x / nsf, nn.y + n4.y / nsf, nn.z + n4.z / nsf, 0.0f);
  s4[2] = make_float4(aa.x + a4.x / nsf, aa.y + a4.y / nsf, aa.z + a4.z / nsf, 0.0f);
  s4[3] = make_float4(pp.x + pos.x / nsf, pp.y + pos.y / nsf, pp.z + pos.z / nsf, 0.0f);
}
__global__ void k_resolve_camera_lcg(rtf_camera_props cam, uint64_t first_pixel, uint32_t n, const float4* __restrict__ sums,
                                     float4* __restrict__ color, float4* __restrict__ normal, float4* __restrict__ albedo) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const float4* s4 = sums + 4 * (size_t)k;
  camera_write(cam, first_pixel + k, f3(s4[0].x, s4[0].y, s4[0].z), f3(s4[1].x, s4[1].y, s4[1].z), f3(s4[2].x, s4[2].y, s4[2].z),
               f3(s4[3].x, s4[3].y, s4[3].z), color, normal, albedo);
}
static int render_camera_lcg(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, uint64_t px_begin,
                             uint64_t n_pixels) {
  const rtf_camera_props& cp = cam->props;
  if (ctx->scene.has_sss) {
    set_error("option rng = reference: scenes with a subsurface material are not supported (order-dependent probe)");
    return RTF_ERR_INVALID;
  }
  const size_t cap = batch_capacity(ctx);
  int rc;
  if ((rc = ensure_workspace(ctx, cap, true))) return rc;
  if ((rc = ctx->ws.lcg.reserve(4 * cap + 64 * cap))) return rc;
  uint32_t* lcg = ctx->ws.lcg.as<uint32_t>();
  float4* sums = (float4*)(lcg + cap);
  Queues q;
  PathState ps;
  bind(ctx, q, ps, true);
  if (!ctx->stats_hold) RTF_CUDA(cudaMemsetAsync(ctx->d_stats.p, 0, 16, ctx->stream));
  ctx->stats_cached = false;
  const PathMap pm{1, 0ull, 0ull, 0u, 0u, 0u, 0u, PixelMap{0ull, 0u, 0u, 0u, 0ull, 0u}, 1u};  // unused by the LCG draws
  for (uint64_t b0 = 0; b0 < n_pixels; b0 += cap) {
    const uint32_t nb = (uint32_t)((n_pixels - b0) < cap ? (n_pixels - b0) : cap);
    LAUNCH(ctx, k_lcg_init_camera, div_up(nb, 256), 256, 0, cp, px_begin + b0, nb, lcg, sums);
    for (int s = 0; s < rp->samples; s++) {
      RTF_CUDA(cudaMemsetAsync(ctx->ws.counters.p, 0, 4096, ctx->stream));
      LAUNCH_P(ctx, PROF_GEN, k_gen_camera_lcg, div_up(nb, 256), 256, 0, cp, px_begin + b0, nb, lcg, q, ps);
      if ((rc = run_wavefronts(ctx, env, rp->bounces, (uint32_t)cp.frame_id, pm, nb, q, ps, true, lcg))) return rc;
      LAUNCH_P(ctx, PROF_REDUCE, k_accum_camera_lcg, div_up(nb, 256), 256, 0, nb, (uint32_t)s, (uint32_t)rp->samples, ps, sums);
    }
    LAUNCH_P(ctx, PROF_REDUCE, k_resolve_camera_lcg, div_up(nb, 256), 256, 0, cp, px_begin + b0, nb, sums, cam->color.as<float4>(),
             cam->normal.as<float4>(), cam->albedo.as<float4>());
  }
  RTF_CUDA(cudaGetLastError());
  ctx->last_paths = n_pixels * (uint64_t)rp->samples;
  return RTF_OK;
}

// One camera job over `n_local` pixels addressed through `pxm`, in batches of the wavefront capacity.
static int render_camera_mapped(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, const PixelMap& job,
                                uint64_t n_local) {
  const rtf_camera_props& cp = cam->props;
  ctx->last_paths = 0;
  if (n_local == 0 || rp->samples <= 0) return RTF_OK;
  if (env->environmental_map_id != 0 && env->lighting_type == RTF_ENV_SCENE && ctx->scene.env_faces == nullptr) {
    set_error("environmental_map_id != 0 but no cube map is registered (rtf_set_environment_map, then rtf_commit)");
    return RTF_ERR_INVALID;
  }
  if (ctx->rng == RTF_RNG_REFERENCE_LCG) {
    if (job.band_px != 0u) {
      set_error("option rng = reference: not available for sharded camera jobs");
      return RTF_ERR_INVALID;
    }
    return render_camera_lcg(ctx, env, cam, rp, job.first, n_local);
  }
  const uint32_t spp = (uint32_t)rp->samples;
  size_t cap = batch_capacity(ctx);
  if (cap < spp) cap = spp;
  int rc;
  if ((rc = ensure_workspace(ctx, cap, true))) return rc;
  Queues q;
  PathState ps;
  bind(ctx, q, ps, true);
  const uint64_t px_per_batch = cap / spp;
  if (!ctx->stats_hold) RTF_CUDA(cudaMemsetAsync(ctx->d_stats.p, 0, 16, ctx->stream));
  ctx->stats_cached = false;
  for (uint64_t l0 = 0; l0 < n_local; l0 += px_per_batch) {
    const uint32_t np = (uint32_t)((n_local - l0) < px_per_batch ? (n_local - l0) : px_per_batch);
    const size_t paths = (size_t)np * spp;
    RTF_CUDA(cudaMemsetAsync(ctx->ws.counters.p, 0, 4096, ctx->stream));
    PixelMap pxm = job;  // the batch's view of the job
    pxm.first = job.first + l0;
    if (job.band_px) pxm.blk0 = l0 / job.band_px, pxm.off0 = (uint32_t)(l0 % job.band_px);
    LAUNCH_P(ctx, PROF_GEN, k_gen_camera, div_up(paths, 256), 256, 0, cp, pxm, np, spp, q, ps);
    PathMap pm{1, pxm.first * spp, 0ull, 0u, 0u, 0u, 0u, pxm, spp};
    if ((rc = run_wavefronts(ctx, env, rp->bounces, (uint32_t)cp.frame_id, pm, paths, q, ps, true))) return rc;
    LAUNCH_P(ctx, PROF_REDUCE, k_resolve_camera, div_up(np, 256), 256, 0, cp, pxm, np, spp, ps, cam->color.as<float4>(),
             cam->normal.as<float4>(), cam->albedo.as<float4>());
  }
  RTF_CUDA(cudaGetLastError());
  ctx->last_paths = n_local * spp;
  return RTF_OK;
}
int render_camera_device(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, uint64_t px_begin,
                         uint64_t px_end) {
  const uint64_t all_pixels = (uint64_t)cam->props.width * (uint64_t)cam->props.height;
  if (px_end > all_pixels) px_end = all_pixels;
  ctx->last_paths = 0;
  if (px_begin >= px_end) return RTF_OK;
  return render_camera_mapped(ctx, env, cam, rp, PixelMap{px_begin, 0u, 0u, 0u, 0ull, 0u}, px_end - px_begin);
}
// This rank's interleaved bands of `band_rows` rows (band k belongs to rank k mod nranks) as one job.
int render_camera_interleaved_device(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, int band_rows,
                                     int nranks, int rank) {
  const int h = cam->props.height, w = cam->props.width;
  ctx->last_paths = 0;
  if (h <= 0 || w <= 0 || band_rows <= 0) return RTF_OK;
  const int n_bands = (h + band_rows - 1) / band_rows;
  uint64_t rows = 0;
  for (int k = rank; k < n_bands; k += nranks) rows += (uint64_t)(((k + 1) * band_rows < h ? (k + 1) * band_rows : h) - k * band_rows);
  if (nranks <= 1) return render_camera_mapped(ctx, env, cam, rp, PixelMap{0ull, 0u, 0u, 0u, 0ull, 0u}, (uint64_t)h * (uint64_t)w);
  if ((uint64_t)band_rows * (uint64_t)w >= (1ull << 31) || batch_capacity(ctx) >= (1ull << 31)) {
    set_error("interleaved camera bands: band or batch larger than 2^31 pixels");
    return RTF_ERR_LIMIT;
  }
  return render_camera_mapped(ctx, env, cam, rp,
                              PixelMap{0ull, (uint32_t)band_rows * (uint32_t)w, (uint32_t)nranks, (uint32_t)rank, 0ull, 0u},
                              rows * (uint64_t)w);
}

// ---- test hook: the BSSRDF spatial probe on its own (rtf_sss.cuh) ----------------------------------------------------
__global__ void k_ss_probe_test(DevScene sc, const float4* __restrict__ rays, const unsigned long long* __restrict__ handles,
                                uint32_t n, uint32_t* __restrict__ out17) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 a = rays[2 * i], b = rays[2 * i + 1];
  SSRecords rec;
  ss_probe(sc, f3(a.x, a.y, a.z), f3(b.x, b.y, b.z), a.w, b.w, handles[i], rec);
  uint32_t* o = out17 + 17 * (size_t)i;
  o[0] = (uint32_t)rec.n;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    o[1 + k] = rec.prim[k], o[5 + k] = __float_as_uint(rec.t[k]);
    o[9 + k] = __float_as_uint(rec.u[k]), o[13 + k] = __float_as_uint(rec.v[k]);
  }
}
int ss_probe_device(rtf_ctx* ctx, const float4* d_rays, const unsigned long long* d_handles, size_t n, uint32_t* d_out17) {
  if (n == 0) return RTF_OK;
  LAUNCH(ctx, k_ss_probe_test, div_up(n, 128), 128, 0, ctx->scene, d_rays, d_handles, (uint32_t)n, d_out17);
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}

// ============================ roofline denominators ===================================================
__global__ void k_fp32_peak(float* out, int iters) {
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.0f, a2 = a0 + 2.0f, a3 = a0 + 3.0f;
  float a4 = a0 + 4.0f, a5 = a0 + 5.0f, a6 = a0 + 6.0f, a7 = a0 + 7.0f;
  const float b = 1.000001f, c = 1e-7f;
  for (int i = 0; i < iters; i++) {
    a0 = fmaf(a0, b, c), a1 = fmaf(a1, b, c), a2 = fmaf(a2, b, c), a3 = fmaf(a3, b, c);
    a4 = fmaf(a4, b, c), a5 = fmaf(a5, b, c), a6 = fmaf(a6, b, c), a7 = fmaf(a7, b, c);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
}
int microbench_fp32(rtf_ctx* ctx, double* tflops) {
  DevBuf out;
  const int blocks = ctx->sm_count * 16, threads = 256, iters = 1 << 14;
  int rc;
  if ((rc = out.reserve((size_t)blocks * threads * 4))) return rc;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  LAUNCH(ctx, k_fp32_peak, blocks, threads, 0, out.as<float>(), iters);
  cudaEventRecord(e0, ctx->stream);
  LAUNCH(ctx, k_fp32_peak, blocks, threads, 0, out.as<float>(), iters);
  cudaEventRecord(e1, ctx->stream);
  cudaError_t e = cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0), cudaEventDestroy(e1);
  out.release();
  if (e != cudaSuccess) return cuda_fail(e, "fp32 microbench", __FILE__, __LINE__);
  *tflops = (double)blocks * threads * (double)iters * 8.0 * 2.0 / (ms * 1e-3) / 1e12;
  return RTF_OK;
}
// L2 -> SM read bandwidth.  Every repetition rotates the index space by a large odd offset, so the lines a block reads
// in repetition r were last touched by OTHER SMs: the 228 KB L1 of an SM never holds them and each 128-bit load is an
// L2 hit (footprint <= ~96 MB of the 126 MB L2).
__global__ void k_l2_read(const uint4* __restrict__ src, size_t n_vec, int reps, uint4* sink) {
  uint4 acc = make_uint4(0, 0, 0, 0);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; r++) {
    const size_t shift = ((size_t)r * 2654435761ull) % n_vec;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride) {
      size_t j = i + shift;
      if (j >= n_vec) j -= n_vec;
      const uint4 v = __ldg(src + j);
      acc.x ^= v.x, acc.y ^= v.y, acc.z ^= v.z, acc.w ^= v.w;
    }
  }
  if (acc.x == 0x12345678u && acc.y == 0x9abcdef0u) *sink = acc;
}
int microbench_l2(rtf_ctx* ctx, double bytes, double* gbs) {
  DevBuf buf, sink;
  const size_t n_vec = (size_t)(bytes / 16.0);
  int rc;
  if ((rc = buf.reserve(n_vec * 16))) return rc;
  if ((rc = sink.reserve(64))) return rc;
  cudaMemsetAsync(buf.p, 1, n_vec * 16, ctx->stream);
  const int reps = 20;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  LAUNCH(ctx, k_l2_read, ctx->sm_count * 8, 512, 0, buf.as<uint4>(), n_vec, 2, sink.as<uint4>());
  cudaEventRecord(e0, ctx->stream);
  LAUNCH(ctx, k_l2_read, ctx->sm_count * 8, 512, 0, buf.as<uint4>(), n_vec, reps, sink.as<uint4>());
  cudaEventRecord(e1, ctx->stream);
  cudaError_t e = cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0), cudaEventDestroy(e1);
  buf.release(), sink.release();
  if (e != cudaSuccess) return cuda_fail(e, "l2 microbench", __FILE__, __LINE__);
  *gbs = (double)n_vec * 16.0 * reps / (ms * 1e-3) / 1e9;
  return RTF_OK;
}

}  // namespace rtf
