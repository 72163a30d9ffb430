// prebake.cu — the kernels either side of the trace/shade path that the reference runs as small CUDA kernels or
// serial host loops:
//   * geometry pre-kernels: CopyVerticesInstancedKernel (RayTracer.cu:1148-1175) + the host index replication
//     (RayTracer.cu:1477-1487), CopySkinnedVerticesKernel (RayTracer.cu:1194-1247);
//   * light-probe construction: TriangleIlluminationEstimator::PrepareLightProbeGroup
//     (TriangleIlluminationEstimator.cpp:94-139);
//   * flux integration: TriangleIlluminationEstimator::SampleLightProbeGroup (:84-92), IlluminationSampler::GetArea
//     (RayTracer.hpp:204-210).
// All of them are streaming, HBM-bound kernels (one thread per vertex / triangle, 16 B vector stores where the
// record allows); arithmetic is written in glm's operation order so that the CPU oracle reproduces every bit.
#include <string.h>

#include <mutex>

#include "rtf_internal.h"

namespace rtf {

#define LAUNCH(ctx, kernel, grid, block, smem, ...)                  \
  do {                                                               \
    kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__); \
    (ctx)->launches++;                                               \
  } while (0)

static inline unsigned div_up(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

// N = normalize(M*(n,0)); T = normalize(M*(t,0)); T = normalize(T - dot(T,N)*N)      (RayTracer.cu:1157-1161)
__device__ __forceinline__ void bake_frame(const float* m, const rtf_vertex& v, float3& P, float3& N, float3& T) {
  P = mat4_mul(m, f3(v.position[0], v.position[1], v.position[2]), 1.0f);
  N = normalize3(mat4_mul(m, f3(v.normal[0], v.normal[1], v.normal[2]), 0.0f));
  T = normalize3(mat4_mul(m, f3(v.tangent[0], v.tangent[1], v.tangent[2]), 0.0f));
  T = normalize3(T - dot3(T, N) * N);
}
// An 80 B vertex as five 16 B stores.
__device__ __forceinline__ void store_vertex(rtf_vertex* out, float3 P, float pp, float3 N, float np, float3 T, float tp,
                                             const float* color, const float* tc, float tcp0, float tcp1) {
  float4* o = reinterpret_cast<float4*>(out);
  o[0] = make_float4(P.x, P.y, P.z, pp);
  o[1] = make_float4(N.x, N.y, N.z, np);
  o[2] = make_float4(T.x, T.y, T.z, tp);
  o[3] = make_float4(color[0], color[1], color[2], color[3]);
  o[4] = make_float4(tc[0], tc[1], tcp0, tcp1);
}

__global__ void k_bake_instanced(const rtf_vertex* __restrict__ verts, uint32_t n_verts, const float* __restrict__ matrices,
                                 const float* __restrict__ colors, uint32_t n_inst, rtf_vertex* __restrict__ out) {
  const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (uint64_t)n_verts * n_inst) return;
  const uint32_t inst = (uint32_t)(idx / n_verts), k = (uint32_t)(idx % n_verts);
  float m[16];
#pragma unroll
  for (int j = 0; j < 16; j++) m[j] = matrices[16 * (size_t)inst + j];
  const rtf_vertex v = verts[k];
  float3 P, N, T;
  bake_frame(m, v, P, N, T);
  float4 c = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
  if (colors) c = *reinterpret_cast<const float4*>(colors + 4 * (size_t)inst);
  store_vertex(out + idx, P, c.x, N, c.y, T, c.z, v.color, v.tex_coord, c.w, (float)inst);
}

__global__ void k_replicate_tris(const uint32_t* __restrict__ tris, uint32_t n_idx, uint32_t n_verts, uint32_t n_inst,
                                 uint32_t* __restrict__ out) {
  const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (uint64_t)n_idx * n_inst) return;
  const uint32_t inst = (uint32_t)(idx / n_idx);
  out[idx] = tris[idx % n_idx] + inst * n_verts;
}

__global__ void k_bake_skinned(const unsigned char* __restrict__ sverts, size_t stride, uint32_t n_verts,
                               const float* __restrict__ bones, uint32_t n_bones, rtf_vertex* __restrict__ out,
                               int* __restrict__ bad) {
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_verts) return;
  const rtf_skinned_vertex& sv = *reinterpret_cast<const rtf_skinned_vertex*>(sverts + (size_t)idx * stride);
  float m[16];
#pragma unroll
  for (int j = 0; j < 16; j++) m[j] = 0.0f;
  bool first = true;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const int id = k < 4 ? sv.bone_id[k] : sv.bone_id2[k - 4];
    const float w = k < 4 ? sv.weight[k] : sv.weight2[k - 4];
    if (k > 0 && id == -1) continue;  // the first influence is always applied (RayTracer.cu:1201-1202)
    if (id < 0 || (uint32_t)id >= n_bones) {
      *bad = 1;
      continue;
    }
    const float* b = bones + 16 * (size_t)id;
    if (first) {
#pragma unroll
      for (int j = 0; j < 16; j++) m[j] = b[j] * w;
      first = false;
    } else {
#pragma unroll
      for (int j = 0; j < 16; j++) m[j] = m[j] + b[j] * w;
    }
  }
  float3 P, N, T;
  bake_frame(m, sv.v, P, N, T);
  store_vertex(out + idx, P, sv.v.position_padding, N, sv.v.normal_padding, T, sv.v.tangent_padding, sv.v.color,
               sv.v.tex_coord, sv.v.tex_coord_padding[0], sv.v.tex_coord_padding[1]);
}

// ---- light probes of committed triangles --------------------------------------------------------------
__global__ void k_make_probes(DevScene sc, uint64_t tri_begin, uint64_t n, int front, int back,
                              rtf_illum_sampler* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t t = (uint32_t)(tri_begin + i);
  const InstDev& in = sc.insts[sc.tri_inst[t]];
  const uint32_t lt = t - in.first_prim;
  rtf_illum_sampler pr;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    rtf_vertex v = in.verts[in.tris[3 * lt + k]];
    const float* p = sc.pos9 + 9 * (size_t)t + 3 * k;  // M*(p,1), M*(n,0): k_flatten's values
    const float* nn = sc.nrm9 + 9 * (size_t)t + 3 * k;
    v.position[0] = p[0], v.position[1] = p[1], v.position[2] = p[2];
    v.normal[0] = nn[0], v.normal[1] = nn[1], v.normal[2] = nn[2];
    (k == 0 ? pr.a : (k == 1 ? pr.b : pr.c)) = v;
  }
  pr.direction[0] = pr.direction[1] = pr.direction[2] = 0.0f;
  pr.energy[0] = pr.energy[1] = pr.energy[2] = 0.0f;
  pr.front_face = front ? 1 : 0, pr.back_face = back ? 1 : 0, pr.pad_[0] = pr.pad_[1] = 0;
  // 268 B = 67 words: word-wise copy keeps the stores of a warp dense
  const uint32_t* src = reinterpret_cast<const uint32_t*>(&pr);
  uint32_t* dst = reinterpret_cast<uint32_t*>(out + i);
#pragma unroll
  for (int w = 0; w < 67; w++) dst[w] = src[w];
}

// ---- flux integration: fixed-shape double tree ----------------------------------------------------------
// glm::distance + Heron (RayTracer.hpp:204-210)
__device__ __forceinline__ float heron_area(float3 A, float3 B, float3 C) {
  const float a = length3(A - B), b = length3(B - C), c = length3(C - A);
  const float p = (a + b + c) * 0.5f;
  return sqrtf(p * (p - a) * (p - b) * (p - c));
}
struct Flux4 {
  double x, y, z, w;
};
__device__ __forceinline__ Flux4 flux_add(Flux4 a, Flux4 b) { return Flux4{a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }

struct LeafProbes {
  const rtf_illum_sampler* probes;
  __device__ Flux4 operator()(uint64_t i) const {
    const rtf_illum_sampler& p = probes[i];
    const float area = heron_area(f3(p.a.position[0], p.a.position[1], p.a.position[2]),
                                  f3(p.b.position[0], p.b.position[1], p.b.position[2]),
                                  f3(p.c.position[0], p.c.position[1], p.c.position[2]));
    const int faces = (p.front_face ? 1 : 0) + (p.back_face ? 1 : 0);
    return Flux4{(double)(p.energy[0] * area), (double)(p.energy[1] * area), (double)(p.energy[2] * area),
                 (double)((float)faces * area)};
  }
};
struct LeafMesh {
  const float* pos9;
  const float* energy6;
  uint64_t tri_begin;
  int faces;
  __device__ Flux4 operator()(uint64_t i) const {
    const uint64_t t = tri_begin + i;
    const float* p = pos9 + 9 * t;
    const float* e = energy6 + 6 * t;
    const float area = heron_area(f3(p[0], p[1], p[2]), f3(p[3], p[4], p[5]), f3(p[6], p[7], p[8]));
    return Flux4{(double)(e[0] * area), (double)(e[1] * area), (double)(e[2] * area), (double)((float)faces * area)};
  }
};
struct LeafPartials {
  const Flux4* v;
  __device__ Flux4 operator()(uint64_t i) const { return v[i]; }
};

// Tile b sums terms [1024 b, 1024 (b+1)): thread t adds terms t, t+256, t+512, t+768 in that order (absent = +0),
// lanes combine by an xor butterfly (16, 8, 4, 2, 1; a+b == b+a, so every lane holds the same bits), thread 0 adds the
// eight warp sums in warp order.
template <class Leaf>
__global__ void __launch_bounds__(256) k_flux_tile(Leaf leaf, uint64_t n, Flux4* __restrict__ out) {
  __shared__ Flux4 ws[8];
  const uint64_t base = (uint64_t)blockIdx.x * 1024u;
  Flux4 s{0.0, 0.0, 0.0, 0.0};
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const uint64_t i = base + threadIdx.x + 256u * j;
    Flux4 v{0.0, 0.0, 0.0, 0.0};
    if (i < n) v = leaf(i);
    s = flux_add(s, v);
  }
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    Flux4 o;
    o.x = __shfl_xor_sync(0xffffffffu, s.x, off), o.y = __shfl_xor_sync(0xffffffffu, s.y, off);
    o.z = __shfl_xor_sync(0xffffffffu, s.z, off), o.w = __shfl_xor_sync(0xffffffffu, s.w, off);
    s = flux_add(s, o);
  }
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    Flux4 r = ws[0];
#pragma unroll
    for (int w = 1; w < 8; w++) r = flux_add(r, ws[w]);
    out[blockIdx.x] = r;
  }
}
__global__ void k_flux_finish(const Flux4* __restrict__ total, float* __restrict__ out8) {
  const Flux4 t = *total;
  const float fx = (float)t.x, fy = (float)t.y, fz = (float)t.z, ar = (float)t.w;
  out8[0] = fx, out8[1] = fy, out8[2] = fz, out8[3] = ar;
  out8[4] = fx / ar, out8[5] = fy / ar, out8[6] = fz / ar, out8[7] = 0.0f;  // m_averageFlux = m_totalFlux / m_totalArea
}

template <class Leaf>
static int flux_reduce(rtf_ctx* ctx, Leaf leaf, uint64_t n, float* out8_host) {
  int rc;
  uint64_t tiles = (n + 1023) / 1024;
  if (tiles == 0) tiles = 1;
  uint64_t total = tiles + (tiles + 1023) / 1024 + 8;
  DevBuf& scratch = ctx->ws.sort_tmp;
  if ((rc = scratch.reserve(total * sizeof(Flux4) + 64))) return rc;
  Flux4* a = scratch.as<Flux4>();
  Flux4* b = a + tiles;
  LAUNCH(ctx, k_flux_tile<Leaf>, (unsigned)tiles, 256, 0, leaf, n, a);
  uint64_t m = tiles;
  while (m > 1) {
    const uint64_t t2 = (m + 1023) / 1024;
    LAUNCH(ctx, k_flux_tile<LeafPartials>, (unsigned)t2, 256, 0, LeafPartials{a}, m, b);
    Flux4* tmp = a;
    a = b, b = tmp;  // level buffers ping-pong (b's region is large enough: levels shrink 1024x)
    m = t2;
  }
  float* d_out = reinterpret_cast<float*>(scratch.as<char>() + total * sizeof(Flux4));
  LAUNCH(ctx, k_flux_finish, 1, 1, 0, a, d_out);
  RTF_CUDA(cudaGetLastError());
  RTF_CUDA(cudaMemcpyAsync(out8_host, d_out, 32, cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTF_OK;
}

int flux_probes_device(rtf_ctx* ctx, const rtf_illum_sampler* d_probes, uint64_t n, float* out8_host) {
  return flux_reduce(ctx, LeafProbes{d_probes}, n, out8_host);
}
int flux_mesh_device(rtf_ctx* ctx, const float* d_energy6, uint64_t tri_begin, uint64_t tri_end, int faces, float* out8_host) {
  return flux_reduce(ctx, LeafMesh{ctx->d_pos9.as<float>(), d_energy6, tri_begin, faces}, tri_end - tri_begin, out8_host);
}
int make_probes_device(rtf_ctx* ctx, uint64_t tri_begin, uint64_t tri_end, int front, int back, rtf_illum_sampler* d_out) {
  const uint64_t n = tri_end - tri_begin;
  if (n == 0) return RTF_OK;
  LAUNCH(ctx, k_make_probes, div_up(n, 128), 128, 0, ctx->scene, tri_begin, n, front, back, d_out);
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}

// Bakes into g.d_verts / g.d_tris and marks the geometry device-resident.
int bake_instanced(rtf_ctx* ctx, Geometry& g, const rtf_vertex* verts, size_t nv, const uint32_t* tris, size_t nt,
                   const float* matrices, const float* colors, size_t n_inst) {
  int rc;
  cudaStream_t st = ctx->stream;
  DevBuf in_v, in_t, in_m, in_c;
  const size_t out_v = nv * n_inst, out_i = 3 * nt * n_inst;
  auto fail = [&](int code) {
    cudaStreamSynchronize(st);
    in_v.release(), in_t.release(), in_m.release(), in_c.release();
    return code;
  };
  if ((rc = in_v.reserve(nv * sizeof(rtf_vertex))) || (rc = in_t.reserve(12 * nt)) || (rc = in_m.reserve(64 * n_inst)) ||
      (colors && (rc = in_c.reserve(16 * n_inst))) || (rc = g.d_verts.reserve(out_v * sizeof(rtf_vertex))) ||
      (rc = g.d_tris.reserve(4 * out_i)))
    return fail(rc);
  cudaMemcpyAsync(in_v.p, verts, nv * sizeof(rtf_vertex), cudaMemcpyHostToDevice, st);
  cudaMemcpyAsync(in_t.p, tris, 12 * nt, cudaMemcpyHostToDevice, st);
  cudaMemcpyAsync(in_m.p, matrices, 64 * n_inst, cudaMemcpyHostToDevice, st);
  if (colors) cudaMemcpyAsync(in_c.p, colors, 16 * n_inst, cudaMemcpyHostToDevice, st);
  if (out_v)
    LAUNCH(ctx, k_bake_instanced, div_up(out_v, 128), 128, 0, in_v.as<rtf_vertex>(), (uint32_t)nv, in_m.as<float>(),
           colors ? in_c.as<float>() : nullptr, (uint32_t)n_inst, g.d_verts.as<rtf_vertex>());
  if (out_i)
    LAUNCH(ctx, k_replicate_tris, div_up(out_i, 256), 256, 0, in_t.as<uint32_t>(), (uint32_t)(3 * nt), (uint32_t)nv,
           (uint32_t)n_inst, g.d_tris.as<uint32_t>());
  cudaError_t e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return fail(cuda_fail(e, "instanced geometry bake", __FILE__, __LINE__));
  fail(0);
  g.n_verts = out_v, g.n_tris = nt * n_inst, g.uploaded = true, g.baked = true;
  return RTF_OK;
}

int bake_skinned(rtf_ctx* ctx, Geometry& g, const void* sverts, size_t nv, size_t stride, const uint32_t* tris, size_t nt,
                 const float* bones, size_t n_bones) {
  int rc;
  cudaStream_t st = ctx->stream;
  DevBuf in_v, in_b, flag;
  auto fail = [&](int code) {
    cudaStreamSynchronize(st);
    in_v.release(), in_b.release(), flag.release();
    return code;
  };
  if ((rc = in_v.reserve(nv * stride)) || (rc = in_b.reserve(64 * n_bones)) || (rc = flag.reserve(4)) ||
      (rc = g.d_verts.reserve(nv * sizeof(rtf_vertex))) || (rc = g.d_tris.reserve(12 * nt)))
    return fail(rc);
  cudaMemcpyAsync(in_v.p, sverts, nv * stride, cudaMemcpyHostToDevice, st);
  cudaMemcpyAsync(in_b.p, bones, 64 * n_bones, cudaMemcpyHostToDevice, st);
  cudaMemcpyAsync(g.d_tris.p, tris, 12 * nt, cudaMemcpyHostToDevice, st);
  cudaMemsetAsync(flag.p, 0, 4, st);
  if (nv)
    LAUNCH(ctx, k_bake_skinned, div_up(nv, 128), 128, 0, in_v.as<unsigned char>(), stride, (uint32_t)nv, in_b.as<float>(),
           (uint32_t)n_bones, g.d_verts.as<rtf_vertex>(), flag.as<int>());
  int bad = 0;
  cudaMemcpyAsync(&bad, flag.p, 4, cudaMemcpyDeviceToHost, st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return fail(cuda_fail(e, "skinned geometry bake", __FILE__, __LINE__));
  fail(0);
  if (bad) {
    set_error("rtf_set_geometry_skinned: bone index out of range");
    return RTF_ERR_INVALID;
  }
  g.n_verts = nv, g.n_tris = nt, g.uploaded = true, g.baked = true;
  return RTF_OK;
}


// ---- flux integration without any allocation --------------------------------------------------------------------------
// Some hosts of this library cannot let it call cudaMalloc: an engine that took the whole device into a memory pool of
// its own at start-up (the driver allocator then has nothing left to give, and the grow-only workspace of flux_reduce
// above fails with out-of-memory), and a frame loop that cannot afford the device-wide synchronisation cudaMalloc and
// cudaFree imply.  For them the partial sums of the tree live in the array below, which is part of the library image:
// every GPU that loads librtf_b200.so holds these bytes from module load until the process exits, used or not.
//
// Capacity: kFluxStaticMaxTris triangles per call = 2^14 first-level tiles of 1024 terms, then 16 second-level sums, then
// the total.  Level 1 writes slots [0, 2^14), level 2 writes [2^14, 2^14 + 16), level 3 writes slot 0 again (level 1 is
// dead by then) — the same ping-pong as flux_reduce, so the bits of the result are the same.
constexpr uint64_t kFluxStaticMaxTris = 1ull << 24;
constexpr uint64_t kFluxStaticSlots = (kFluxStaticMaxTris / 1024) + 32;  // 16 416 x 32 B = 525 312 B per GPU
static __device__ __align__(256) Flux4 g_flux_static[kFluxStaticSlots];

// In host code `g_flux_static` is only a handle the runtime uses to find the symbol; it is not a device address and must
// never be handed to a kernel.  The address is asked for once per device, with that device current, and kept here.
constexpr int kMaxDevices = 64;
static Flux4* g_flux_static_ptr[kMaxDevices] = {};
static std::mutex g_flux_static_mutex;

static int flux_static_address(int device, Flux4** out) {
  if (device < 0 || device >= kMaxDevices) {
    set_error("rtf_illumination_flux_mesh_device_static: device ordinal beyond the 64 the address table holds");
    return RTF_ERR_LIMIT;
  }
  std::lock_guard<std::mutex> lock(g_flux_static_mutex);
  if (!g_flux_static_ptr[device]) {
    Flux4* p = nullptr;
    RTF_CUDA(cudaGetSymbolAddress((void**)&p, g_flux_static));  // the caller made `device` current (CHECK_CTX)
    if (!p) {
      set_error("cudaGetSymbolAddress returned a null address for the static flux scratch");
      return RTF_ERR_CUDA;
    }
    g_flux_static_ptr[device] = p;
  }
  *out = g_flux_static_ptr[device];
  return RTF_OK;
}

}  // namespace rtf

using namespace rtf;

extern "C" int rtf_illumination_flux_mesh_device_static(rtf_ctx* ctx, const float* dev_energy6, uint64_t tri_begin,
                                                        uint64_t tri_end, int front, int back, float* dev_out8) {
  if (!ctx) {
    set_error("null context");
    return RTF_ERR_INVALID;
  }
  RTF_CUDA(cudaSetDevice(ctx->device));
  if (!dev_energy6 || !dev_out8) {
    set_error("rtf_illumination_flux_mesh_device_static: null energy or output pointer");
    return RTF_ERR_INVALID;
  }
  if (!ctx->ready) {
    set_error("no acceleration structure (call rtf_commit)");
    return RTF_ERR_NOT_READY;
  }
  if (tri_begin > tri_end) {
    set_error("triangle range out of bounds");
    return RTF_ERR_INVALID;
  }
  const uint64_t n = tri_end - tri_begin;
  if (n > kFluxStaticMaxTris) {  // said first: the answer to "too many" is to split the range, whatever the scene holds
    set_error("rtf_illumination_flux_mesh_device_static: " + std::to_string(n) + " triangles in the range, the static scratch "
              "holds the partial sums of at most 16777216 (2^24); split the range or use rtf_illumination_flux_mesh_device");
    return RTF_ERR_LIMIT;
  }
  if (tri_end > ctx->n_tris) {
    set_error("triangle range out of bounds");
    return RTF_ERR_INVALID;
  }
  int rc;
  Flux4* a = nullptr;
  if ((rc = flux_static_address(ctx->device, &a))) return rc;
  Flux4* b = a + kFluxStaticMaxTris / 1024;
  // Nothing here has to start at zero: every slot a level reads was written by the level below in this same call, and a
  // tile's absent terms are added as +0 inside the kernel.  What the previous call left behind is never read.
  uint64_t m = (n + 1023) / 1024;
  if (m == 0) m = 1;
  const LeafMesh leaf{ctx->d_pos9.as<float>(), dev_energy6, tri_begin, (front ? 1 : 0) + (back ? 1 : 0)};
  LAUNCH(ctx, k_flux_tile<LeafMesh>, (unsigned)m, 256, 0, leaf, n, a);
  while (m > 1) {
    const uint64_t t2 = (m + 1023) / 1024;
    LAUNCH(ctx, k_flux_tile<LeafPartials>, (unsigned)t2, 256, 0, LeafPartials{a}, m, b);
    Flux4* tmp = a;
    a = b, b = tmp;
    m = t2;
  }
  LAUNCH(ctx, k_flux_finish, 1, 1, 0, a, dev_out8);
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}
