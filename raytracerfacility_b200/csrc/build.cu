// build.cu — device LBVH builder: flatten -> bounds -> pre-split references -> Morton 30/63 -> radix sort
// (rtf_sort.cuh) -> Karras radix tree ->
// refit -> breadth-first greedy surface-area collapse to 8-wide -> quantised 80 B node encode.
//
// Replaces RayTracedGeometry::BuildGAS + RayTracer::BuildIAS of the reference
// (/root/reference/src/RayTracerFacility/RayTracer.cu:1251-1598, :1618-1715: optixAccelBuild +
// optixAccelCompact, three device syncs and one D2H per geometry) and the instanced pre-bake kernel
// CopyVerticesInstancedKernel (RayTracer.cu:1148-1175) with a single world-space triangle soup.
//
// Every stage is an integer or exactly-rounded float computation, so tests compare each array bit for
// bit with the CPU oracle (Morton codes, sorted order, radix-tree links, boxes, wide nodes, triangles).
#include <algorithm>

#include "rtf_internal.h"
#include "rtf_sort.cuh"

namespace rtf {

#define LAUNCH(ctx, kernel, grid, block, smem, ...)                      \
  do {                                                                   \
    kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);     \
    (ctx)->launches++;                                                   \
  } while (0)

static inline unsigned div_up(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

// ---- 1. flatten instances to world space (TriangleIlluminationEstimator.cpp:110-118 semantics:
//         position = M*(p,1), normal = M*(n,0) NOT renormalised) ---------------------------------
__global__ void k_flatten(const InstDev* __restrict__ insts, uint32_t n_insts, uint32_t n_tris, float* __restrict__ pos9,
                          float* __restrict__ nrm9, uint32_t* __restrict__ tri_inst) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tris) return;
  uint32_t lo = 0, hi = n_insts;  // last instance with first_prim <= t
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (insts[mid].first_prim <= t)
      lo = mid;
    else
      hi = mid;
  }
  const InstDev& in = insts[lo];
  const uint32_t lt = t - in.first_prim;
  float m[16];
#pragma unroll
  for (int k = 0; k < 16; k++) m[k] = in.xform[k];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const rtf_vertex& v = in.verts[in.tris[3 * lt + k]];
    const float3 p = mat4_mul(m, f3(v.position[0], v.position[1], v.position[2]), 1.0f);
    const float3 n = mat4_mul(m, f3(v.normal[0], v.normal[1], v.normal[2]), 0.0f);
    float* po = pos9 + 9 * (size_t)t + 3 * k;
    float* no = nrm9 + 9 * (size_t)t + 3 * k;
    po[0] = p.x, po[1] = p.y, po[2] = p.z;
    no[0] = n.x, no[1] = n.y, no[2] = n.z;
  }
  tri_inst[t] = lo;
}

// ---- 2. scene bounds: ordered-int atomics (min/max are exact, any order gives the same bits) ------
__device__ __forceinline__ int float_to_ordered(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float ordered_to_float(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__device__ __forceinline__ Box3 tri_box(const float* __restrict__ pos9, uint32_t t) {
  const float* p = pos9 + 9 * (size_t)t;
  const float3 a = f3(p[0], p[1], p[2]), b = f3(p[3], p[4], p[5]), c = f3(p[6], p[7], p[8]);
  Box3 bx;
  bx.lo = min3(min3(a, b), c);
  bx.hi = max3(max3(a, b), c);
  return bx;
}

// ---- exclusive prefix sum of u32, single pass with decoupled look-back (Merrill & Garland 2016) -----------------
// Tiles of 2048 elements are taken in arrival order (atomic ticket), so a tile only ever waits for tiles that are
// already running.  Tile state = one 64-bit word {flag, value}: flag 1 = tile aggregate, 2 = inclusive prefix.
// out[n] receives the grand total (callers size `out` for n + 1 elements).
constexpr int kScanBlock = 256, kScanItems = 8, kScanTile = kScanBlock * kScanItems;
__global__ void __launch_bounds__(kScanBlock) k_scan_u32(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                         uint32_t n, unsigned long long* state /* [0]=ticket, [1+tile] */) {
  __shared__ uint32_t s_tile, s_warp[kScanBlock / 32], s_prefix;
  if (threadIdx.x == 0) s_tile = (uint32_t)atomicAdd(state, 1ull);
  __syncthreads();
  const uint32_t tile = s_tile;
  volatile unsigned long long* desc = state + 1;
  const uint32_t base = tile * kScanTile + threadIdx.x * kScanItems;
  uint32_t v[kScanItems], sum = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; k++) {
    v[k] = base + k < n ? in[base + k] : 0u;
    sum += v[k];
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = sum;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t warp_off = 0, total = 0;
#pragma unroll
  for (int w = 0; w < kScanBlock / 32; w++) {
    if (w < warp) warp_off += s_warp[w];
    total += s_warp[w];
  }
  if (warp == 0) {
    uint32_t prefix = 0;
    if (tile == 0) {
      if (lane == 0) desc[0] = (2ull << 32) | total;
    } else {
      if (lane == 0) desc[tile] = (1ull << 32) | total;
      int look = (int)tile - 1;
      for (;;) {
        const int idx = look - lane;
        unsigned long long d = idx >= 0 ? desc[idx] : (2ull << 32);
        while (__any_sync(0xffffffffu, (d >> 32) == 0ull)) d = idx >= 0 ? desc[idx] : (2ull << 32);
        const unsigned done = __ballot_sync(0xffffffffu, (d >> 32) == 2ull);
        const int first_done = done ? __ffs((int)done) - 1 : 32;  // nearest predecessor holding an inclusive prefix
        uint32_t contrib = lane <= first_done ? (uint32_t)d : 0u;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, off);
        prefix += contrib;
        if (done) break;
        look -= 32;
      }
      if (lane == 0) desc[tile] = (2ull << 32) | (unsigned long long)(prefix + total);
    }
    if (lane == 0) s_prefix = prefix;
  }
  __syncthreads();
  uint32_t run = s_prefix + warp_off + (incl - sum);
#pragma unroll
  for (int k = 0; k < kScanItems; k++) {
    if (base + k < n) out[base + k] = run;
    run += v[k];
    if (base + k + 1 == n) out[n] = run;
  }
}
static int exclusive_scan_u32(rtf_ctx* ctx, const uint32_t* in, uint32_t* out, uint32_t n, DevBuf& state) {
  const uint32_t tiles = n ? (n + kScanTile - 1) / kScanTile : 0;
  if (tiles == 0) {
    RTF_CUDA(cudaMemsetAsync(out, 0, 4, ctx->stream));
    return RTF_OK;
  }
  int rc;
  if ((rc = state.reserve(8 * ((size_t)tiles + 1)))) return rc;
  RTF_CUDA(cudaMemsetAsync(state.p, 0, 8 * ((size_t)tiles + 1), ctx->stream));
  k_scan_u32<<<tiles, kScanBlock, 0, ctx->stream>>>(in, out, n, state.as<unsigned long long>());
  ctx->launches++;
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}

// ---- 2b. triangle references: count, (scan), emit ------------------------------------------------------------------
__device__ __forceinline__ void load_tri(const float* __restrict__ pos9, uint32_t t, float3& a, float3& b, float3& c) {
  const float* p = pos9 + 9 * (size_t)t;
  a = f3(p[0], p[1], p[2]), b = f3(p[3], p[4], p[5]), c = f3(p[6], p[7], p[8]);
}
__global__ void k_split_count(const float* __restrict__ pos9, uint32_t n, int max_split, uint32_t* __restrict__ cnt) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  float3 a, b, c;
  load_tri(pos9, t, a, b, c);
  cnt[t] = (uint32_t)split_plan(a, b, c, max_split).k;
}
// One thread per REFERENCE: a block owns 256 consecutive triangles, i.e. the contiguous run of references
// [first[T0], first[T0 + 256)), and its threads stride over that run — every lane clips one slab per step and the stores
// are coalesced.  (One thread per triangle looping over its 1..8 slabs ran at the lane count of the longest loop in the
// warp and scattered 28-byte stores: 161 us for the C2 scene.)  The triangle of a reference is found by a binary search
// over the block's 257 offsets in shared memory; the slab axis is the longest axis of the triangle's box, the slab count
// the one k_split_count planned (build) or the last build kept (refit: ref_tri == nullptr).
__global__ void __launch_bounds__(256) k_split_emit(const float* __restrict__ pos9, uint32_t n, const uint32_t* __restrict__ first,
                                                    uint32_t* __restrict__ ref_tri, float* __restrict__ ref_box) {
  __shared__ uint32_t s_first[257];
  const uint32_t T0 = blockIdx.x * 256u;
  const uint32_t nt = min(256u, n - T0);
  for (uint32_t k = threadIdx.x; k <= nt; k += 256u) s_first[k] = first[T0 + k];
  __syncthreads();
  const uint32_t R0 = s_first[0], R1 = s_first[nt];
  for (uint32_t r = R0 + threadIdx.x; r < R1; r += 256u) {
    uint32_t lo = 0, hi = nt;  // s_first[lo] <= r < s_first[hi]
    while (hi - lo > 1u) {
      const uint32_t mid = (lo + hi) >> 1;
      if (s_first[mid] <= r) lo = mid; else hi = mid;
    }
    const uint32_t t = T0 + lo;
    const int k = (int)(s_first[lo + 1] - s_first[lo]), j = (int)(r - s_first[lo]);
    float3 a, b, c;
    load_tri(pos9, t, a, b, c);
    const int axis = split_plan(a, b, c, 1).axis;
    const Box3 bx = split_slab_box(a, b, c, axis, k, j);
    if (ref_tri) ref_tri[r] = t;
    float2* o = (float2*)(ref_box + 6 * (size_t)r);
    o[0] = make_float2(bx.lo.x, bx.lo.y), o[1] = make_float2(bx.lo.z, bx.hi.x), o[2] = make_float2(bx.hi.y, bx.hi.z);
  }
}
__device__ __forceinline__ Box3 ref_box_load(const float* __restrict__ ref_box, uint32_t r) {
  const float* b = ref_box + 6 * (size_t)r;
  Box3 bx;
  bx.lo = f3(b[0], b[1], b[2]), bx.hi = f3(b[3], b[4], b[5]);
  return bx;
}

__global__ void k_bounds_init(int* ob) {
  if (threadIdx.x < 3) ob[threadIdx.x] = float_to_ordered(INFINITY);
  else if (threadIdx.x < 6) ob[threadIdx.x] = float_to_ordered(-INFINITY);
}
__global__ void k_bounds(const float* __restrict__ pos9, uint32_t n, int* __restrict__ ob) {
  float3 lo = f3(INFINITY, INFINITY, INFINITY), hi = f3(-INFINITY, -INFINITY, -INFINITY);
  for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
    const Box3 b = tri_box(pos9, t);
    lo = min3(lo, b.lo), hi = max3(hi, b.hi);
  }
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    lo.x = fminf(lo.x, __shfl_xor_sync(0xffffffffu, lo.x, off));
    lo.y = fminf(lo.y, __shfl_xor_sync(0xffffffffu, lo.y, off));
    lo.z = fminf(lo.z, __shfl_xor_sync(0xffffffffu, lo.z, off));
    hi.x = fmaxf(hi.x, __shfl_xor_sync(0xffffffffu, hi.x, off));
    hi.y = fmaxf(hi.y, __shfl_xor_sync(0xffffffffu, hi.y, off));
    hi.z = fmaxf(hi.z, __shfl_xor_sync(0xffffffffu, hi.z, off));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(ob + 0, float_to_ordered(lo.x)), atomicMin(ob + 1, float_to_ordered(lo.y));
    atomicMin(ob + 2, float_to_ordered(lo.z)), atomicMax(ob + 3, float_to_ordered(hi.x));
    atomicMax(ob + 4, float_to_ordered(hi.y)), atomicMax(ob + 5, float_to_ordered(hi.z));
  }
}
__global__ void k_bounds_finish(const int* ob, float* fb) {
  if (threadIdx.x < 6) fb[threadIdx.x] = ordered_to_float(ob[threadIdx.x]);
}

// ---- 3. Morton codes ---------------------------------------------------------------------------
__global__ void k_morton(const float* __restrict__ ref_box, uint32_t n, const float* __restrict__ fb, int bits,
                         uint64_t* __restrict__ keys) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const float3 lo = f3(fb[0], fb[1], fb[2]), hi = f3(fb[3], fb[4], fb[5]);
  const float3 ext = hi - lo;
  const float3 inv = f3(ext.x > 0 ? 1.0f / ext.x : 0.0f, ext.y > 0 ? 1.0f / ext.y : 0.0f, ext.z > 0 ? 1.0f / ext.z : 0.0f);
  const Box3 b = ref_box_load(ref_box, t);
  const float3 c = (b.lo + b.hi) * 0.5f;
  keys[t] = morton_code(c, lo, inv, bits);
}

// ---- 5. Karras 2012 radix tree -------------------------------------------------------------------
__device__ __forceinline__ int karras_delta(const uint64_t* __restrict__ keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const uint64_t a = keys[i], b = keys[j];
  if (a == b) return 64 + __clz((int)((uint32_t)i ^ (uint32_t)j));
  return __clzll((long long)(a ^ b));
}
__global__ void k_karras(const uint64_t* __restrict__ keys, int n, int2* __restrict__ children, uint2* __restrict__ range) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int ni = n - 1;
  if (i >= ni) return;
  const int d = (karras_delta(keys, n, i, i + 1) - karras_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
  const int dmin = karras_delta(keys, n, i, i - d);
  int lmax = 2;
  while (karras_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int t = lmax / 2; t >= 1; t /= 2)
    if (karras_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = karras_delta(keys, n, i, j);
  int s = 0, t = l;
  do {
    t = (t + 1) >> 1;
    if (karras_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  const int gamma = i + s * d + min(d, 0);
  const int lo = min(i, j), hi = max(i, j);
  const int left = (lo == gamma) ? ~gamma : gamma;
  const int right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
  children[i] = make_int2(left, right);
  range[i] = make_uint2((uint32_t)lo, (uint32_t)hi);
}

// ---- 6. node boxes without a bottom-up walk --------------------------------------------------------------------------
// A radix-tree node covers a contiguous range [lo, hi] of the sorted references, so its box is a range union over the
// sorted leaf boxes.  A segment tree over those (level j, entry i = union of leaves [i*2^j, (i+1)*2^j)) is built by plain
// reductions (8 levels per launch: 5 by warp shuffles, 3 through shared memory), then every node unions the O(log len)
// aligned blocks that tile its range.  min/max are exact, so the boxes are bit-identical to a bottom-up merge, and
// nothing waits on anything: the earlier atomic-flag climb (one __threadfence per level per thread) took 1.05 ms at
// 1.3 M references, fence-throughput bound.
struct SegLevels {
  uint32_t off[32];  // first entry of level j (j >= 1) in the float4 pair array
  uint32_t cnt[32];  // entries of level j (cnt[0] = n)
  int levels;        // highest level (cnt[levels] == 1)
};
__device__ __forceinline__ Box3 box_union(const Box3& a, const Box3& b) {
  Box3 r;
  r.lo = min3(a.lo, b.lo), r.hi = max3(a.hi, b.hi);
  return r;
}
__device__ __forceinline__ Box3 box_empty() {
  Box3 r;
  r.lo = f3(INFINITY, INFINITY, INFINITY), r.hi = f3(-INFINITY, -INFINITY, -INFINITY);
  return r;
}
__device__ __forceinline__ Box3 box_shfl_down(const Box3& b, int d) {
  Box3 r;
  r.lo.x = __shfl_down_sync(0xffffffffu, b.lo.x, d), r.lo.y = __shfl_down_sync(0xffffffffu, b.lo.y, d);
  r.lo.z = __shfl_down_sync(0xffffffffu, b.lo.z, d), r.hi.x = __shfl_down_sync(0xffffffffu, b.hi.x, d);
  r.hi.y = __shfl_down_sync(0xffffffffu, b.hi.y, d), r.hi.z = __shfl_down_sync(0xffffffffu, b.hi.z, d);
  return r;
}
__device__ __forceinline__ void seg_store(float4* __restrict__ seg, const SegLevels& sl, int j, uint32_t i, const Box3& b) {
  if (j > sl.levels || i >= sl.cnt[j]) return;
  float4* o = seg + 2 * ((size_t)sl.off[j] + i);
  o[0] = make_float4(b.lo.x, b.lo.y, b.lo.z, 0.0f), o[1] = make_float4(b.hi.x, b.hi.y, b.hi.z, 0.0f);
}
__device__ __forceinline__ Box3 seg_load(const float4* __restrict__ seg, const SegLevels& sl, int j, uint32_t i) {
  const float4* o = seg + 2 * ((size_t)sl.off[j] + i);
  const float4 lo = o[0], hi = o[1];
  Box3 r;
  r.lo = f3(lo.x, lo.y, lo.z), r.hi = f3(hi.x, hi.y, hi.z);
  return r;
}
// Levels j0+1 .. j0+8 from level j0.  j0 == 0: the entries are the leaf boxes (gathered through the sorted order and also
// written to boxes[ni + k], where the collapse reads them).
__global__ void __launch_bounds__(256) k_seg_reduce(const float* __restrict__ ref_box, const uint32_t* __restrict__ sorted_ids,
                                                    float* __restrict__ boxes, float4* __restrict__ seg, SegLevels sl, int j0) {
  __shared__ float s_box[8][6];
  const uint32_t k = blockIdx.x * 256u + threadIdx.x;
  const int lane = threadIdx.x & 31;
  Box3 b = box_empty();
  if (k < sl.cnt[j0]) {
    if (j0 == 0) {
      b = ref_box_load(ref_box, sorted_ids[k]);
      float2* bo = (float2*)(boxes + 6 * ((size_t)sl.cnt[0] - 1 + k));
      bo[0] = make_float2(b.lo.x, b.lo.y), bo[1] = make_float2(b.lo.z, b.hi.x), bo[2] = make_float2(b.hi.y, b.hi.z);
    } else {
      b = seg_load(seg, sl, j0, k);
    }
  }
#pragma unroll
  for (int j = 1; j <= 5; j++) {
    b = box_union(b, box_shfl_down(b, 1 << (j - 1)));
    if ((lane & ((1 << j) - 1)) == 0) seg_store(seg, sl, j0 + j, k >> j, b);
  }
  if (lane == 0) {
    float* o = s_box[threadIdx.x >> 5];
    o[0] = b.lo.x, o[1] = b.lo.y, o[2] = b.lo.z, o[3] = b.hi.x, o[4] = b.hi.y, o[5] = b.hi.z;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    const float* o = s_box[lane & 7];
    b.lo = f3(o[0], o[1], o[2]), b.hi = f3(o[3], o[4], o[5]);
#pragma unroll
    for (int j = 1; j <= 3; j++) {
      b = box_union(b, box_shfl_down(b, 1 << (j - 1)));
      if (lane < 8 && (lane & ((1 << j) - 1)) == 0) seg_store(seg, sl, j0 + 5 + j, ((k >> 5) + (uint32_t)lane) >> j, b);
    }
  }
}
__device__ __forceinline__ Box3 seg_get(const float* __restrict__ leaf_boxes, const float4* __restrict__ seg, const SegLevels& sl,
                                        int j, uint32_t i) {
  if (j == 0) {
    const float2* b = (const float2*)(leaf_boxes + 6 * (size_t)i);
    const float2 p = b[0], q = b[1], r = b[2];
    Box3 x;
    x.lo = f3(p.x, p.y, q.x), x.hi = f3(q.y, r.x, r.y);
    return x;
  }
  return seg_load(seg, sl, j, i);
}
__device__ __forceinline__ Box3 seg_query(const float* __restrict__ leaf_boxes, const float4* __restrict__ seg, const SegLevels& sl,
                                          uint32_t lo, uint32_t hi) {
  Box3 acc = box_empty();
  uint32_t l = lo, r = hi + 1u;
  for (int j = 0; l < r; j++, l >>= 1, r >>= 1) {
    if (l & 1u) acc = box_union(acc, seg_get(leaf_boxes, seg, sl, j, l++));
    if (r & 1u) acc = box_union(acc, seg_get(leaf_boxes, seg, sl, j, --r));
  }
  return acc;
}
__global__ void k_node_boxes(const uint2* __restrict__ range, int ni, float* __restrict__ boxes, const float4* __restrict__ seg,
                             SegLevels sl) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ni) return;
  const uint2 rg = range[i];
  const Box3 b = seg_query(boxes + 6 * (size_t)ni, seg, sl, rg.x, rg.y);
  float2* bo = (float2*)(boxes + 6 * (size_t)i);
  bo[0] = make_float2(b.lo.x, b.lo.y), bo[1] = make_float2(b.lo.z, b.hi.x), bo[2] = make_float2(b.hi.y, b.hi.z);
}

// ---- 7. collapse to 8-wide: ONE persistent cooperative kernel for all breadth-first levels ----------------------------
// The node layout is breadth first (a node's internal children are consecutive in the next level, its triangles
// consecutive in the index array), so a level needs the exclusive prefix sums of its nodes' child and triangle counts.
// Each WARP takes 32-node tiles of the current level in ticket order: greedy expansion + slot assignment per node
// (registers only), a warp scan, a decoupled look-back over the tile descriptors of the level (Merrill & Garland), then
// the nodes are encoded and the next level's roots written — one pass per level, no intermediate arrays.  Levels are
// separated by a grid barrier; the level sizes never leave the device (the previous version ran four launches per level
// and read two counters back to the host after each: 2.5 ms of a 4.8 ms build were that).
struct CollapseArgs {
  const int2* children;
  const uint2* range;
  const float* boxes;
  int n;
};
__device__ __forceinline__ Box3 load_box(const float* __restrict__ boxes, int ni, int c) {
  const float* b = boxes + 6 * (size_t)(c >= 0 ? c : ni + ~c);
  Box3 r;
  r.lo = f3(b[0], b[1], b[2]), r.hi = f3(b[3], b[4], b[5]);
  return r;
}
constexpr int kEmptySlot = 0x7fffffff;
constexpr int kCollapseBlock = 128;
constexpr int kCollapseLevels = kMaxStackDepth + 2;
struct CollapseLevel {
  uint32_t n, start, tri_cursor, desc_base;
};
__device__ __forceinline__ CollapseLevel level_load(const CollapseLevel* p) {  // written by another block: volatile reads
  const volatile uint32_t* v = (const volatile uint32_t*)p;
  CollapseLevel l;
  l.n = v[0], l.start = v[1], l.tri_cursor = v[2], l.desc_base = v[3];
  return l;
}
__device__ __forceinline__ void level_store(CollapseLevel* p, const CollapseLevel& l) {
  volatile uint32_t* v = (volatile uint32_t*)p;
  v[0] = l.n, v[1] = l.start, v[2] = l.tri_cursor, v[3] = l.desc_base;
}
struct CollapseState {  // device memory, zeroed before the launch except level[0]
  CollapseLevel level[kCollapseLevels];
  uint32_t ticket[kCollapseLevels];
  uint32_t depth, n_nodes, n_tris, error;
};

// Greedy expansion by largest half-area + greedy octant slot assignment of binary subtree R: the binary child of every
// slot (kEmptySlot: none), the number of internal children and of leaf triangles.
// Written as straight-line register code: every array index below is a compile-time constant after unrolling (a dynamic
// index would put ch / dv / the slot tables into local memory — the first version did, and the kernel sat at 15 % issue
// utilisation waiting for its own stack), each child's half-area and centre offset are computed once when the child
// enters the set, and the (child, slot) search runs as eight independent per-child chains.  The decisions — first
// maximum in index order, strict comparisons — are those of the loop form, which the CPU oracle keeps.
__device__ __forceinline__ void collapse_plan(const CollapseArgs& a, int R, int (&slot_child)[8], uint32_t& ci, uint32_t& ct) {
  const int ni = a.n - 1;
  const Box3 nb = load_box(a.boxes, ni, R);
  const float3 center = (nb.lo + nb.hi) * 0.5f;
  int ch[8];
  float ar[8];  // half-area of an internal child; -inf: leaf or unused (never chosen for expansion)
  float3 dv[8];
#pragma unroll
  for (int i = 0; i < 8; i++) ch[i] = kEmptySlot, ar[i] = -INFINITY, dv[i] = f3(0, 0, 0);
  auto enter = [&](int c, float& area, float3& d) {
    const Box3 cb = load_box(a.boxes, ni, c);
    area = c >= 0 ? half_area(cb) : -INFINITY;
    d = (cb.lo + cb.hi) * 0.5f - center;
  };
  int nc;
  if (R < 0) {
    ch[0] = R, nc = 1;
    enter(R, ar[0], dv[0]);
  } else {
    const int2 c2 = a.children[R];
    ch[0] = c2.x, ch[1] = c2.y, nc = 2;
    enter(c2.x, ar[0], dv[0]);
    enter(c2.y, ar[1], dv[1]);
  }
  while (nc < 8) {
    int best = -1;
    float best_area = -INFINITY;
#pragma unroll
    for (int i = 0; i < 8; i++)
      if (ar[i] > best_area) best_area = ar[i], best = i;
    if (best < 0) break;
    int cb = 0;
#pragma unroll
    for (int i = 0; i < 8; i++)
      if (i == best) cb = ch[i];
    const int2 c2 = a.children[cb];
    float ax, ay;
    float3 dx, dy;
    enter(c2.x, ax, dx);
    enter(c2.y, ay, dy);
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (i == best) ch[i] = c2.x, ar[i] = ax, dv[i] = dx;
      if (i == nc) ch[i] = c2.y, ar[i] = ay, dv[i] = dy;
    }
    nc++;
  }
  // greedy slot assignment: nc times, the (child, slot) pair of largest cost among the unassigned children and free
  // slots, first pair in (child, slot) order on ties.  cost(i, s) = ((+-dx) + (+-dy)) + (+-dz) with the signs of slot s;
  // the complementary slot 7 - s flips all three signs and IEEE addition is symmetric under negation, so
  // cost(i, 7 - s) = -cost(i, s) exactly: four sums per child, computed once, serve all eight slots of every round.
  float cst[8][4];
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int k = 0; k < 4; k++) cst[i][k] = ((-dv[i].x) + ((k & 2) ? dv[i].y : -dv[i].y)) + ((k & 1) ? dv[i].z : -dv[i].z);
  uint32_t open_i = (1u << nc) - 1u, free_s = 0xffu;
  int child_in_slot[8];
#pragma unroll
  for (int s = 0; s < 8; s++) child_in_slot[s] = -1;
  for (int it = 0; it < nc; it++) {
    float bc[8];
    int bsl[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {  // best free slot of child i (independent chains)
      float best = 0.0f;
      int bs = -1;
#pragma unroll
      for (int s = 0; s < 8; s++) {
        const float cost = s < 4 ? cst[i][s] : -cst[i][7 - s];
        if (((free_s >> s) & 1u) && (bs < 0 || cost > best)) best = cost, bs = s;
      }
      bc[i] = best, bsl[i] = bs;
    }
    float best = 0.0f;
    int bi = -1, bs = -1;
#pragma unroll
    for (int i = 0; i < 8; i++)
      if (((open_i >> i) & 1u) && (bi < 0 || bc[i] > best)) best = bc[i], bi = i, bs = bsl[i];
    open_i &= ~(1u << bi), free_s &= ~(1u << bs);
#pragma unroll
    for (int s = 0; s < 8; s++)
      if (s == bs) child_in_slot[s] = bi;
  }
  ci = 0, ct = 0;
#pragma unroll
  for (int s = 0; s < 8; s++) {
    const int i = child_in_slot[s];
    int c = kEmptySlot;
#pragma unroll
    for (int k = 0; k < 8; k++)
      if (k == i) c = ch[k];
    if (i >= 0) {
      const uint32_t cnt = c >= 0 ? (a.range[c].y - a.range[c].x + 1u) : 1u;
      if (c >= 0 && cnt > 3u)
        ci++;
      else
        ct += cnt;
    }
    slot_child[s] = c;
  }
}

// Quantisation frame and the eight child boxes of a wide node from the boxes of its binary subtree roots (shared by the
// collapse and the in-place refit, so both encode a given topology to the same bytes).
__device__ __forceinline__ void encode_frame(const float* __restrict__ boxes, int ni, int R, const int (&slot_child)[8], Node80& node) {
  const Box3 nb = load_box(boxes, ni, R);
  node.p[0] = nb.lo.x, node.p[1] = nb.lo.y, node.p[2] = nb.lo.z;
  const float3 ext = nb.hi - nb.lo;
  const uint32_t e[3] = {quant_exponent(ext.x), quant_exponent(ext.y), quant_exponent(ext.z)};
  node.e[0] = (uint8_t)e[0], node.e[1] = (uint8_t)e[1], node.e[2] = (uint8_t)e[2];
  for (int s = 0; s < 8; s++) {
    const int c = slot_child[s];
    node.qlo[0][s] = node.qlo[1][s] = node.qlo[2][s] = 0;
    node.qhi[0][s] = node.qhi[1][s] = node.qhi[2][s] = 0;
    if (c == kEmptySlot) continue;
    const Box3 cb = load_box(boxes, ni, c);
    quantize_interval(cb.lo.x, cb.hi.x, node.p[0], e[0], node.qlo[0][s], node.qhi[0][s]);
    quantize_interval(cb.lo.y, cb.hi.y, node.p[1], e[1], node.qlo[1][s], node.qhi[1][s]);
    quantize_interval(cb.lo.z, cb.hi.z, node.p[2], e[2], node.qlo[2][s], node.qhi[2][s]);
  }
}
// Encode node R (80 B), write its internal children into the next level and, per triangle slot, the SORTED POSITION of
// the reference it holds (k_tri_records turns those into 48 B records with coalesced stores).  wide_src keeps the
// binary roots {R, slot children} of the node for the in-place refit.
__device__ __forceinline__ void collapse_emit(const CollapseArgs& a, int R, const int (&slot_child)[8], uint32_t child_base,
                                              uint32_t tri_base, int* __restrict__ next_level_at, Node80* __restrict__ out,
                                              uint32_t* __restrict__ slot_pos, int* __restrict__ wide_src) {
  const int ni = a.n - 1;
  Node80 node;
  encode_frame(a.boxes, ni, R, slot_child, node);
  node.imask = 0;
  node.child_base = child_base;
  node.tri_base = tri_base;
  uint32_t tri_off = 0, int_off = 0, valid = 0, cnt16 = 0;
  wide_src[0] = R;
  for (int s = 0; s < 8; s++) {
    const int c = slot_child[s];
    wide_src[1 + s] = c;
    if (c == kEmptySlot) continue;
    const uint32_t cnt = c >= 0 ? (a.range[c].y - a.range[c].x + 1u) : 1u;
    if (c >= 0 && cnt > 3u) {
      node.imask |= (uint8_t)(1u << s);
      next_level_at[int_off++] = c;
    } else {
      valid |= (cnt == 1u ? 1u : (cnt == 2u ? 3u : 7u)) << (3 * s);
      cnt16 |= cnt << (2 * s);
      const uint32_t first = c >= 0 ? a.range[c].x : (uint32_t)~c;
      for (uint32_t j = 0; j < cnt; j++) slot_pos[tri_base + tri_off + j] = first + j;
      tri_off += cnt;
    }
  }
  node_meta_store(node, valid, cnt16);
  *out = node;
}
// In-place refit of the 8-wide nodes: the topology (slots, child and triangle bases, meta) of the last build stays, only
// the quantisation frames and child boxes are re-encoded from the new binary-node boxes.  One thread per node, no levels.
constexpr int kWideSrcInts = 9;
__global__ void k_wide_refit(const float* __restrict__ boxes, int ni, const int* __restrict__ wide_src, uint32_t n_nodes,
                             Node80* __restrict__ nodes) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  const int* ws = wide_src + kWideSrcInts * (size_t)i;
  int slot_child[8];
#pragma unroll
  for (int s = 0; s < 8; s++) slot_child[s] = ws[1 + s];
  Node80 node = nodes[i];
  encode_frame(boxes, ni, ws[0], slot_child, node);
  nodes[i] = node;
}

// Tile descriptor of the look-back: bits 62..63 = state (1 aggregate, 2 inclusive prefix), bits 32..61 = internal
// children, bits 0..31 = triangles (references <= 2^31, nodes < 2^30: the fields cannot carry into each other).
constexpr int kCollapseTile = 32;  // nodes per tile: one WARP owns a tile, so a level runs without block barriers
__global__ void __launch_bounds__(kCollapseBlock)
    k_collapse(CollapseArgs a, CollapseState* cs, unsigned long long* desc_all, int* lv_a, int* lv_b, Node80* __restrict__ nodes,
               uint32_t* __restrict__ slot_pos, int* __restrict__ wide_src, uint32_t max_nodes, uint32_t max_level, unsigned* gbar) {
  const int lane = threadIdx.x & 31;
  int L = 0;
  for (;; L++) {
    if (L >= kCollapseLevels - 1) {
      if (blockIdx.x == 0 && threadIdx.x == 0) cs->error = 2u;  // deeper than the traversal stack can hold
      break;
    }
    const CollapseLevel lv = level_load(&cs->level[L]);
    if (lv.n == 0) break;
    if ((size_t)lv.start + lv.n > max_nodes || lv.n > max_level) {
      if (blockIdx.x == 0 && threadIdx.x == 0) cs->error = 1u;  // node budget exceeded (same decision in every block)
      break;
    }
    const int* level = (L & 1) ? lv_b : lv_a;
    int* next = (L & 1) ? lv_a : lv_b;
    volatile unsigned long long* desc = desc_all + lv.desc_base;
    const uint32_t tiles = (lv.n + kCollapseTile - 1) / kCollapseTile;
    for (;;) {
      uint32_t tile = 0;
      if (lane == 0) tile = atomicAdd(&cs->ticket[L], 1u);
      tile = __shfl_sync(0xffffffffu, tile, 0);
      if (tile >= tiles) break;
      const uint32_t li = tile * kCollapseTile + (uint32_t)lane;
      const bool active = li < lv.n;
      int slot_child[8];
      uint32_t ci = 0, ct = 0;
      int R = 0;
      if (active) {
        R = level[li];
        collapse_plan(a, R, slot_child, ci, ct);
      }
      __syncwarp();
      const unsigned long long mine = ((unsigned long long)ci << 32) | ct;
      unsigned long long incl = mine;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += t;
      }
      const unsigned long long total = __shfl_sync(0xffffffffu, incl, 31);
      unsigned long long prefix = 0;
      if (tile == 0) {
        if (lane == 0) desc[0] = (2ull << 62) | total;
      } else {
        if (lane == 0) desc[tile] = (1ull << 62) | total;
        int look = (int)tile - 1;
        for (;;) {
          const int idx = look - lane;
          unsigned long long d = idx >= 0 ? desc[idx] : (2ull << 62);
          while (__any_sync(0xffffffffu, (d >> 62) == 0ull)) d = idx >= 0 ? desc[idx] : (2ull << 62);
          const unsigned done = __ballot_sync(0xffffffffu, (d >> 62) == 2ull);
          const int first_done = done ? __ffs((int)done) - 1 : 32;
          unsigned long long contrib = lane <= first_done ? (d & 0x3fffffffffffffffull) : 0ull;
#pragma unroll
          for (int off = 16; off >= 1; off >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, off);
          prefix += contrib;
          if (done) break;
          look -= 32;
        }
        if (lane == 0) desc[tile] = (2ull << 62) | (prefix + total);
      }
      if (lane == 0 && tile == tiles - 1) {  // the level's totals: the next level's size and cursors
        const unsigned long long all = prefix + total;
        CollapseLevel nx;
        nx.n = (uint32_t)(all >> 32), nx.start = lv.start + lv.n, nx.tri_cursor = lv.tri_cursor + (uint32_t)all;
        nx.desc_base = lv.desc_base + tiles;
        level_store(&cs->level[L + 1], nx);
      }
      if (active) {
        const unsigned long long excl = prefix + (incl - mine);
        const uint32_t child_off = (uint32_t)(excl >> 32), tri_off = (uint32_t)excl;
        collapse_emit(a, R, slot_child, lv.start + lv.n + child_off, lv.tri_cursor + tri_off, next + child_off,
                      nodes + lv.start + li, slot_pos, wide_src + kWideSrcInts * (size_t)(lv.start + li));
      }
      __syncwarp();
    }
    // grid barrier between levels (all blocks are resident: cooperative launch): arrival counter
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned target = (unsigned)(L + 1) * gridDim.x;
      atomicAdd(gbar, 1u);
      while (*(volatile unsigned*)gbar < target) {
      }
      __threadfence();
    }
    __syncthreads();
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const CollapseLevel lv = level_load(&cs->level[L]);
    cs->depth = (uint32_t)L, cs->n_nodes = lv.start, cs->n_tris = lv.tri_cursor;
  }
}

// 48 B triangle records in leaf-slot order from the sorted positions the collapse left per slot: one thread per slot, three
// coalesced 16 B stores (the collapse itself only writes 4 B per slot).
__global__ void k_tri_records(uint32_t n, const uint32_t* __restrict__ slot_pos, const uint32_t* __restrict__ sorted_ids,
                              const uint32_t* __restrict__ ref_tri, const float* __restrict__ pos9,
                              const uint32_t* __restrict__ tri_inst, const float* __restrict__ inst_alpha, Tri48* __restrict__ tris) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t prim = ref_tri[sorted_ids[slot_pos[i]]];
  const float* p = pos9 + 9 * (size_t)prim;
  Tri48 T;
  T.v0[0] = p[0], T.v0[1] = p[1], T.v0[2] = p[2];
  T.v1[0] = p[3], T.v1[1] = p[4], T.v1[2] = p[5];
  T.v2[0] = p[6], T.v2[1] = p[7], T.v2[2] = p[8];
  T.prim = prim;
  T.inst = tri_inst[prim];
  T.flags = (inst_alpha && inst_alpha[T.inst] < 1.0f) ? (uint32_t)TRI_FLAG_ALPHA : 0u;
  tris[i] = T;
}

static void mark(rtf_ctx* ctx, int k) {
  if (!ctx->build_ev[k]) cudaEventCreate(&ctx->build_ev[k]);
  cudaEventRecord(ctx->build_ev[k], ctx->stream);
}

int flatten_scene(rtf_ctx* ctx) {
  const uint32_t n = ctx->n_tris;
  if (n == 0) return RTF_OK;
  mark(ctx, 0);
  LAUNCH(ctx, k_flatten, div_up(n, 256), 256, 0, ctx->scene.insts, ctx->scene.n_insts, n, ctx->d_pos9.as<float>(),
         ctx->d_nrm9.as<float>(), ctx->d_tri_inst.as<uint32_t>());
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}

static int collapse_encode(rtf_ctx* ctx, uint32_t n);

// Leaf boxes in sorted order + the box of every radix-tree node (segment tree over the sorted leaves, see section 6).
static int node_boxes(rtf_ctx* ctx, uint32_t n) {
  BuildArrays& ba = ctx->ba;
  SegLevels sl;
  memset(&sl, 0, sizeof(sl));
  sl.cnt[0] = n;
  uint32_t total = 0;
  int j = 0;
  while (sl.cnt[j] > 1u) {
    sl.cnt[j + 1] = (sl.cnt[j] + 1u) / 2u;
    sl.off[j + 1] = total;
    total += sl.cnt[j + 1];
    j++;
  }
  sl.levels = j;
  int rc;
  if ((rc = ba.seg.reserve(32 * ((size_t)total + 1)))) return rc;
  float4* seg = ba.seg.as<float4>();
  for (int j0 = 0; j0 == 0 || j0 < sl.levels; j0 += 8)
    LAUNCH(ctx, k_seg_reduce, div_up(sl.cnt[j0], 256), 256, 0, ba.ref_box.as<float>(), ba.sorted_ids.as<uint32_t>(),
           ba.boxes.as<float>(), seg, sl, j0);
  if (n > 1)
    LAUNCH(ctx, k_node_boxes, div_up(n - 1, 256), 256, 0, ba.range.as<uint2>(), (int)n - 1, ba.boxes.as<float>(), seg, sl);
  return RTF_OK;
}

// Fast path for moved geometry with unchanged topology (option "refit"): the sorted reference order, the radix tree AND
// the 8-wide topology of the last full build are reused; reference boxes, radix-node boxes (segment tree), the quantised
// frames of the wide nodes (in place) and the triangle records are recomputed: four streaming kernels, no levels, no host
// round trip besides the final event.  The tree keeps the shape it had for the old positions, so its quality decays as the
// scene deforms; hits do not depend on it.  Unmoved geometry refits to the bytes of the full build.
int refit_bvh(rtf_ctx* ctx) {
  const uint32_t nt = ctx->n_tris, n = ctx->n_refs;
  BuildArrays& ba = ctx->ba;
  if (nt == 0 || n == 0 || !ctx->refit_ok || ctx->n_nodes == 0) return RTF_ERR_NOT_READY;
  int rc;
  mark(ctx, 1);
  LAUNCH(ctx, k_split_emit, div_up(nt, 256), 256, 0, ctx->d_pos9.as<float>(), nt, ba.ref_first.as<uint32_t>(), (uint32_t*)nullptr,
         ba.ref_box.as<float>());
  mark(ctx, 2);
  mark(ctx, 3);
  if ((rc = node_boxes(ctx, n))) return rc;
  mark(ctx, 4);
  LAUNCH(ctx, k_wide_refit, div_up(ctx->n_nodes, 128), 128, 0, ba.boxes.as<float>(), (int)n - 1, ba.wide_src.as<int>(),
         ctx->n_nodes, ctx->d_nodes.as<Node80>());
  LAUNCH(ctx, k_tri_records, div_up(n, 256), 256, 0, n, ba.slot_pos.as<uint32_t>(), ba.sorted_ids.as<uint32_t>(),
         ba.ref_tri.as<uint32_t>(), ctx->d_pos9.as<float>(), ctx->d_tri_inst.as<uint32_t>(), ctx->d_inst_alpha.as<float>(),
         ctx->d_tris48.as<Tri48>());
  mark(ctx, 5);
  RTF_CUDA(cudaEventSynchronize(ctx->build_ev[5]));
  for (int k = 0; k < 5; k++) cudaEventElapsedTime(&ctx->build_ms[k], ctx->build_ev[k], ctx->build_ev[k + 1]);
  cudaEventElapsedTime(&ctx->build_ms[5], ctx->build_ev[0], ctx->build_ev[5]);
  return RTF_OK;
}

int build_bvh(rtf_ctx* ctx, int bits) {
  const uint32_t nt = ctx->n_tris;
  BuildArrays& ba = ctx->ba;
  cudaStream_t st = ctx->stream;
  ctx->n_nodes = 0, ctx->depth = 0, ctx->n_refs = 0;
  if (nt == 0) return RTF_OK;
  int rc;
  if ((rc = ba.scene_bounds.reserve(64))) return rc;
  int* ob = ba.scene_bounds.as<int>();
  float* fb = ba.scene_bounds.as<float>() + 8;
  mark(ctx, 1);
  LAUNCH(ctx, k_bounds_init, 1, 32, 0, ob);
  LAUNCH(ctx, k_bounds, min(div_up(nt, 256), (unsigned)ctx->sm_count * 8u), 256, 0, ctx->d_pos9.as<float>(), nt, ob);
  LAUNCH(ctx, k_bounds_finish, 1, 32, 0, ob, fb);
  // triangle references: count -> scan -> emit (one reference per triangle when presplit <= 1)
  if ((rc = ba.ref_cnt.reserve(4 * (size_t)nt))) return rc;
  if ((rc = ba.ref_first.reserve(4 * ((size_t)nt + 1)))) return rc;
  LAUNCH(ctx, k_split_count, div_up(nt, 256), 256, 0, ctx->d_pos9.as<float>(), nt, ctx->presplit, ba.ref_cnt.as<uint32_t>());
  if ((rc = exclusive_scan_u32(ctx, ba.ref_cnt.as<uint32_t>(), ba.ref_first.as<uint32_t>(), nt, ba.scan_state))) return rc;
  uint32_t n = 0;
  RTF_CUDA(cudaMemcpyAsync(&n, ba.ref_first.as<uint32_t>() + nt, 4, cudaMemcpyDeviceToHost, st));
  RTF_CUDA(cudaStreamSynchronize(st));
  if (n < nt || (uint64_t)n > 0x7fffffffull) {
    set_error("triangle reference count out of range");
    return RTF_ERR_LIMIT;
  }
  ctx->n_refs = n;
  const int ni = (int)n - 1;
  if ((rc = ba.ref_tri.reserve(4 * (size_t)n))) return rc;
  if ((rc = ba.ref_box.reserve(24 * (size_t)n))) return rc;
  if ((rc = ba.morton.reserve(8 * (size_t)n))) return rc;
  if ((rc = ba.sorted_keys.reserve(8 * (size_t)n))) return rc;
  if ((rc = ba.sorted_ids.reserve(4 * (size_t)n))) return rc;
  if ((rc = ba.ids_alt.reserve(4 * (size_t)n))) return rc;
  if ((rc = ba.children.reserve(8 * (size_t)n))) return rc;
  if ((rc = ba.range.reserve(8 * (size_t)n))) return rc;
  if ((rc = ba.boxes.reserve(24 * (size_t)(2 * (size_t)n)))) return rc;
  LAUNCH(ctx, k_split_emit, div_up(nt, 256), 256, 0, ctx->d_pos9.as<float>(), nt, ba.ref_first.as<uint32_t>(),
         ba.ref_tri.as<uint32_t>(), ba.ref_box.as<float>());
  LAUNCH(ctx, k_morton, div_up(n, 256), 256, 0, ba.ref_box.as<float>(), n, fb, bits, ba.morton.as<uint64_t>());
  mark(ctx, 2);
  // 4. stable LSD radix sort of (Morton code, reference id) pairs, 8-bit digits (rtf_sort.cuh); ids start as 0..n-1
  if ((rc = ba.keys_alt.reserve(8 * (size_t)n))) return rc;
  if ((rc = radix_sort_pairs<uint64_t>(ctx, ba.morton.as<uint64_t>(), ba.sorted_keys.as<uint64_t>(), ba.keys_alt.as<uint64_t>(),
                                       nullptr, ba.sorted_ids.as<uint32_t>(), ba.ids_alt.as<uint32_t>(), n, 0,
                                       bits == 30 ? 30 : 63, ba.sort_hist, ba.sort_status)))
    return rc;
  mark(ctx, 3);
  if (ni > 0)
    LAUNCH(ctx, k_karras, div_up(ni, 256), 256, 0, ba.sorted_keys.as<uint64_t>(), (int)n, ba.children.as<int2>(),
           ba.range.as<uint2>());
  if ((rc = node_boxes(ctx, n))) return rc;
  ctx->refit_ok = false;
  if ((rc = collapse_encode(ctx, n))) return rc;
  ctx->refit_ok = true;
  return RTF_OK;
}

// Radix tree + node boxes -> 8-wide compressed nodes, leaf-slot triangle indices and the per-triangle records (shared by
// build and refit).  One cooperative launch + two streaming kernels; the only host synchronisation is the one at the end
// that reads {depth, nodes, triangles, error} back.
static int collapse_encode(rtf_ctx* ctx, uint32_t n) {
  BuildArrays& ba = ctx->ba;
  cudaStream_t st = ctx->stream;
  int rc;
  mark(ctx, 4);
  const size_t max_nodes = (size_t)n / 2 + 64;
  const size_t max_level = (size_t)n / 4 + 64;
  const size_t max_tiles = max_nodes / kCollapseTile + kCollapseLevels + 2;
  if ((rc = ctx->d_nodes.reserve(80 * max_nodes))) return rc;
  if ((rc = ba.slot_pos.reserve(4 * (size_t)n))) return rc;
  if ((rc = ba.wide_src.reserve(4 * kWideSrcInts * max_nodes))) return rc;
  if ((rc = ctx->d_tris48.reserve(48 * (size_t)n))) return rc;
  if ((rc = ba.level_a.reserve(4 * max_level))) return rc;
  if ((rc = ba.level_b.reserve(4 * max_level))) return rc;
  if ((rc = ba.collapse_state.reserve(sizeof(CollapseState) + 64))) return rc;
  if ((rc = ba.scan_state.reserve(8 * max_tiles))) return rc;
  CollapseState* cs = ba.collapse_state.as<CollapseState>();
  unsigned* gbar = (unsigned*)((char*)ba.collapse_state.p + sizeof(CollapseState));
  RTF_CUDA(cudaMemsetAsync(ba.collapse_state.p, 0, sizeof(CollapseState) + 64, st));
  RTF_CUDA(cudaMemsetAsync(ba.scan_state.p, 0, 8 * max_tiles, st));
  const CollapseLevel l0{1u, 0u, 0u, 0u};
  const int root = n == 1 ? ~0 : 0;
  RTF_CUDA(cudaMemcpyAsync(&cs->level[0], &l0, sizeof(l0), cudaMemcpyHostToDevice, st));
  RTF_CUDA(cudaMemcpyAsync(ba.level_a.p, &root, 4, cudaMemcpyHostToDevice, st));
  if (ctx->collapse_blocks == 0) {
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_collapse, kCollapseBlock, 0) != cudaSuccess || nb < 1) nb = 1;
    ctx->collapse_blocks = nb;
  }
  CollapseArgs ca{ba.children.as<int2>(), ba.range.as<uint2>(), ba.boxes.as<float>(), (int)n};
  unsigned long long* desc = ba.scan_state.as<unsigned long long>();
  int* lva = ba.level_a.as<int>();
  int* lvb = ba.level_b.as<int>();
  Node80* nodes = ctx->d_nodes.as<Node80>();
  uint32_t* slot_pos = ba.slot_pos.as<uint32_t>();
  int* wide_src = ba.wide_src.as<int>();
  uint32_t mn = (uint32_t)max_nodes, ml = (uint32_t)max_level;
  const unsigned want = div_up(max_level, kCollapseBlock);  // more blocks than the widest level has tiles would only spin
  const unsigned full = (unsigned)ctx->sm_count * (unsigned)ctx->collapse_blocks;
  void* args[] = {&ca, &cs, &desc, &lva, &lvb, &nodes, &slot_pos, &wide_src, &mn, &ml, &gbar};
  RTF_CUDA(cudaLaunchCooperativeKernel((const void*)k_collapse, dim3(want < full ? want : full), dim3(kCollapseBlock), args, 0, st));
  ctx->launches++;
  LAUNCH(ctx, k_tri_records, div_up(n, 256), 256, 0, n, slot_pos, ba.sorted_ids.as<uint32_t>(), ba.ref_tri.as<uint32_t>(),
         ctx->d_pos9.as<float>(), ctx->d_tri_inst.as<uint32_t>(), ctx->d_inst_alpha.as<float>(), ctx->d_tris48.as<Tri48>());
  mark(ctx, 5);
  uint32_t res[4] = {0, 0, 0, 0};  // depth, nodes, triangle slots, error
  RTF_CUDA(cudaMemcpyAsync(res, &cs->depth, 16, cudaMemcpyDeviceToHost, st));
  RTF_CUDA(cudaEventSynchronize(ctx->build_ev[5]));
  RTF_CUDA(cudaStreamSynchronize(st));
  for (int k = 0; k < 5; k++) cudaEventElapsedTime(&ctx->build_ms[k], ctx->build_ev[k], ctx->build_ev[k + 1]);
  cudaEventElapsedTime(&ctx->build_ms[5], ctx->build_ev[0], ctx->build_ev[5]);
  if (res[3] == 1u) {
    set_error("BVH collapse exceeded its node budget");
    return RTF_ERR_LIMIT;
  }
  if (res[3] == 2u || (int)res[0] > kMaxStackDepth) {
    set_error("BVH deeper than the traversal stack");
    return RTF_ERR_LIMIT;
  }
  if (res[2] != n) {
    set_error("BVH collapse lost triangles");
    return RTF_ERR_LIMIT;
  }
  ctx->n_nodes = res[1];
  ctx->depth = (int)res[0];
  return RTF_OK;
}

}  // namespace rtf
