// rtf_internal.h — context, device scene and workspace shared by the translation units of librtf_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/rtf/rtf.h"
#include "rtf_btf.cuh"
#include "rtf_bvh.cuh"

namespace rtf {

constexpr int kMaxStackDepth = 64;  // wide-BVH depth the traversal stack can hold (checked at commit)

void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define RTF_CUDA(expr)                                                         \
  do {                                                                         \
    cudaError_t e_ = (expr);                                                   \
    if (e_ != cudaSuccess) return ::rtf::cuda_fail(e_, #expr, __FILE__, __LINE__); \
  } while (0)

// Grow-only device buffer (replaces the reference's CudaBuffer, CUDABuffer.cpp:13-38, which frees and
// cudaMallocs on every size change and on every entry-point call, CUDAModule.cpp:43-49).
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return RTF_OK;
    if (p) cudaFree(p);
    p = nullptr, cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc", __FILE__, __LINE__);
    cap = bytes;
    return RTF_OK;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr, cap = 0;
  }
  template <class T>
  T* as() const {
    return (T*)p;
  }
};

struct InstDev {
  float xform[16];
  uint64_t handle;  // private component handle (SBT::m_handle, RayTracer.cu:1859)
  uint32_t mat;
  uint32_t first_prim;
  uint32_t n_prims;
  uint32_t pad_;
  const rtf_vertex* verts;
  const uint32_t* tris;
};

// A material as the kernels see it: the MaterialProperties fields + the texture / BTF tables it points at
// (RayTracedMaterial, RT.hpp:247-269 -> SurfaceMaterial / SurfaceCompressedBTF, RDD.hpp:240-354).
enum { MAT_TEX_ALBEDO = 0, MAT_TEX_NORMAL = 1, MAT_TEX_METALLIC = 2, MAT_TEX_ROUGHNESS = 3 };
struct DevMaterial {
  rtf_material m;
  DevTexture tex[4];  // texels == nullptr: constant from m
  int btf;            // index into DevScene::btfs (material_type == RTF_MAT_COMPRESSED_BTF), else -1
  int rich;           // shading needs interpolated vertex attributes (texture coordinate, tangent, vertex colour)
};

// Everything a kernel needs to trace and shade, passed by value.
struct DevScene {
  const uint4* nodes = nullptr;
  const float4* tris = nullptr;
  const float* pos9 = nullptr;
  const float* nrm9 = nullptr;
  const uint32_t* tri_inst = nullptr;
  const InstDev* insts = nullptr;
  const DevMaterial* mats = nullptr;
  const DevBTF* btfs = nullptr;
  const DevTexture* env_faces = nullptr;  // 6 cube-map faces, or nullptr
  const float* inst_alpha = nullptr;      // per instance: constant alpha, or < 0 = alpha comes from the albedo texture
  uint32_t n_tris = 0, n_nodes = 0, n_insts = 0;
  int has_alpha = 0;
  int has_sss = 0;  // some material has subsurfaceFactor > 0 and subsurfaceRadius.x > 0 (BSSRDF, RF.cuh:97-133)
  int has_btf = 0;  // some material is a compressed BTF (RF.cuh:172-258)
};

struct Geometry {
  std::vector<rtf_vertex> verts;
  std::vector<uint32_t> tris;
  uint64_t version = 0;
  DevBuf d_verts, d_tris;
  bool uploaded = false;
  bool baked = false;            // instanced / skinned: d_verts and d_tris were produced on the device (prebake.cu)
  size_t n_verts = 0, n_tris = 0;  // counts of the device arrays
};
struct Texture {
  int w = 0, h = 0;
  DevBuf texels;
};
struct MaterialLinks {  // handles; 0 = none
  uint64_t tex[4] = {0, 0, 0, 0};
  uint64_t btf = 0;
};
struct BTFTables {
  rtf_btf_desc desc;  // scalars only (pointers are dead after rtf_set_btf)
  DevBuf pdf6_slices, pdf6_scales, pdf1d_basis, colors, index_ab, pdf2d_colors, pdf2d_slices, pdf2d_scales, pdf2d_indices,
      pdf3d_scales, pdf3d_slices, pdf4d_scales, pdf4d_slices, beta_angles;
  DevBTF dev() const;
  void release();
};
struct Instance {
  uint64_t geom, mat;
  float xform[16];
  uint64_t component;
};

// Intermediate arrays of the builder kept for parity read-back (rtf_debug_read).
struct BuildArrays {
  DevBuf morton, sorted_keys, sorted_ids, children, range, boxes, scene_bounds;
  DevBuf seg, wide_src;  // segment tree over the sorted leaf boxes (32 B entries); binary roots {R, 8 slot children} per wide node
  DevBuf keys_alt, ids_alt, level_a, level_b;
  DevBuf ref_cnt, ref_first, ref_tri, ref_box, scan_state;  // triangle references (pre-splitting)
  DevBuf sort_hist, sort_status;                            // radix sort scratch (rtf_sort.cuh)
  DevBuf collapse_state, slot_pos;                          // k_collapse: level table + tickets + grid barrier; per leaf slot its sorted reference
};

// Per-category device time of the wavefront kernels (option "profile"): an event pair around every launch, summed when
// read (rtf_profile_read).  Categories follow the phases the bench reports.
enum { PROF_GEN = 0, PROF_TRACE, PROF_SHADE, PROF_REDUCE, PROF_SORT, PROF_OTHER, PROF_COUNT };
struct Profiler {
  bool on = false;
  std::vector<cudaEvent_t> pool;
  size_t used = 0;
  struct Rec {
    int cat;
    cudaEvent_t a, b;
  };
  std::vector<Rec> recs;
  cudaEvent_t open_a = nullptr;
  int open_cat = 0;
  cudaEvent_t take() {
    if (used == pool.size()) {
      cudaEvent_t e = nullptr;
      cudaEventCreate(&e);
      pool.push_back(e);
    }
    return pool[used++];
  }
  void begin(int cat, cudaStream_t st) {
    if (!on) return;
    open_a = take(), open_cat = cat;
    cudaEventRecord(open_a, st);
  }
  void end(cudaStream_t st) {
    if (!on) return;
    cudaEvent_t b = take();
    cudaEventRecord(b, st);
    recs.push_back(Rec{open_cat, open_a, b});
  }
};

struct Workspace {  // wavefront queues + path state, sized for `capacity` paths
  size_t capacity = 0;
  DevBuf ray_a[2], ray_b[2], hits, counters;
  DevBuf thr, acc, dir0w, first_n, first_a, first_p;  // per-path state (SoA float4)
  DevBuf pend_o, pend_d, pend_t;                      // BSSRDF: bounce ray waiting for the subsurface path
  DevBuf sort_keys, sort_keys_alt, sort_vals, sort_vals_alt, sort_tmp, sort_kin;
  DevBuf io;  // staging for AoS entry points
  DevBuf lcg;  // option "rng" = reference: one LCG state per path of a batch
};

// Double-buffered host<->device pipeline of the host-buffer entry points (upload chunk i+1 and download chunk i-1
// while chunk i is traced).  The reference uploads, launches, synchronises and downloads serially
// (CUDAModule.cpp:43-49, 57-63).
struct HostPipe {
  bool init = false;
  cudaStream_t in = nullptr, out = nullptr;
  // kDepth chunks in flight: with two, a buffer is busy for upload + kernel + download, so neither PCIe direction stays
  // saturated (measured: 25 GB/s each way on the 112 B LiDAR records); four keep both directions and the SMs busy.
  static constexpr int kDepth = 4;
  cudaEvent_t ev_in[kDepth] = {}, ev_done[kDepth] = {}, ev_out[kDepth] = {};
  DevBuf rays[kDepth], hits[kDepth];
  DevBuf aos[kDepth], aux[kDepth];  // AoS entry points: record chunk (updated in place) + per-record side array
};

}  // namespace rtf

struct rtf_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  uint64_t launches = 0;
  int sm_count = 148;
  int trace_blocks[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // resident blocks per SM of each trace-kernel variant (occupancy query
                                                   // at its first launch, render.cu trace_grid)
  std::map<uint64_t, rtf::Geometry> geoms;
  std::map<uint64_t, rtf_material> mats;
  std::map<uint64_t, rtf::Instance> insts;
  std::map<uint64_t, rtf::Texture> textures;
  std::map<uint64_t, rtf::MaterialLinks> mat_links;
  std::map<uint64_t, rtf::BTFTables> btfs;
  uint64_t env_map[6] = {0, 0, 0, 0, 0, 0};
  bool has_env_map = false;
  // scene-sync diffing (the reference's m_updateFlag / m_version bookkeeping, RayTracerLayer.cpp:47-92, RT.hpp:305-327):
  // geom_dirty = a geometry or instance changed since the last commit (BVH rebuild); mat_dirty = only material-side
  // tables changed (materials, textures, BTFs, cube map: re-resolved without touching the BVH)
  bool geom_dirty = true, mat_dirty = true;
  bool refit_ok = false;  // the builder arrays (sorted order, radix tree, wide_src, slot_pos) describe the committed tree
  bool topo_dirty = true;  // a geometry / instance appeared, vanished or changed its triangle count, or "presplit" changed
                           // (geom_dirty && !topo_dirty = vertices or transforms moved: candidates for the refit path)
  int refit = 0;           // option "refit"
  uint64_t commits_refit = 0;
  std::vector<uint8_t> alpha_class;  // per committed instance: may ignore hits (decides TRI_FLAG_ALPHA in the BVH)
  uint64_t commits_full = 0, commits_materials = 0, commits_skipped = 0, geometry_uploads_skipped = 0;
  // committed state
  bool ready = false;
  bool bvh_ok = false;  // the committed BVH and flattened arrays are intact (ready may be false only because a texture / BTF
                        // the material tables point at was replaced: the next commit re-resolves the tables and keeps the BVH)
  uint64_t epoch = 0;   // bumped by every commit that changed the committed scene (camera accumulation restarts on a change)
  int morton_bits = 63;
  int depth = 0;
  uint32_t n_tris = 0, n_nodes = 0;
  uint32_t n_refs = 0;  // triangle references in the BVH (>= n_tris when long triangles are pre-split)
  int presplit = 8;     // option "presplit": most references one triangle may be cut into (1 = off)
  int collapse_blocks = 0;  // resident k_collapse blocks per SM (occupancy query at the first build)
  rtf::DevBuf d_pos9, d_nrm9, d_tri_inst, d_nodes, d_tris48, d_insts, d_mats, d_inst_alpha, d_btfs, d_env_faces;
  rtf::BuildArrays ba;
  rtf::Workspace ws;
  rtf::HostPipe pipe;
  size_t pipe_chunk_rays = (size_t)1 << 22;  // option "pipe_chunk_rays"
  size_t pipe_chunk_probes = (size_t)1 << 17;   // option "pipe_chunk_probes": 268 B IlluminationSampler records per chunk
  size_t pipe_chunk_samples = (size_t)1 << 19;  // option "pipe_chunk_samples": 112 B PointCloudSample records per chunk
  bool stats_hold = false;  // a chunked job is running: the ray / path counters keep adding up across its calls
  rtf::DevScene scene;
  rtf::Profiler prof;
  uint64_t last_rays = 0, last_paths = 0;
  bool stats_cached = false;  // last_rays already holds the job's total (multi-band camera frames sum it on the host)
  rtf::DevBuf d_stats;  // {rays u64, paths u64}
  // options (rtf_set_option)
  size_t batch_paths = (size_t)1 << 24;  // 2^24 paths: ~3 GB of queues + path state; +12..20 % over 2^22 (fewer, fuller waves)
  int sort_rays = 0;  // option "sort_rays": sort bounce rays by (octant, origin cell) before tracing (measured: DESIGN.md)
  int brdf = RTF_BRDF_REFERENCE;  // option "brdf"
  int rng = RTF_RNG_PHILOX;       // option "rng" (rtf_estimate_illumination, single-context camera calls)
  // RenderToCamera's change detection (m_cameraRenderingLaunchParams.m_rayTracerProperties, RayTracer.cu:253-269): the
  // environment / ray properties and scene epoch the last frame was rendered with
  rtf_env cam_last_env;
  rtf_ray_props cam_last_rp;
  uint64_t cam_last_epoch = 0;
  bool cam_last_valid = false;
  // multi-GPU (multi.cu): rank of this context in its communicator; comm is an ncclComm_t
  void* comm = nullptr;
  bool own_comm = false;
  int rank = 0, nranks = 1;
  rtf::DevBuf d_shard_out, d_small;
  cudaEvent_t phase_ev[4] = {nullptr, nullptr, nullptr, nullptr};  // compute | collective | finalize | end of the last sharded call
  bool phase_valid = false;
  // build-stage events: flatten | bounds+morton | sort | karras+refit | collapse | end
  cudaEvent_t build_ev[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  float build_ms[6] = {0, 0, 0, 0, 0, 0};
};

struct rtf_camera {
  rtf_ctx* ctx = nullptr;
  rtf_camera_props props;
  rtf::DevBuf color, normal, albedo;
};

namespace rtf {
// prebake.cu
int bake_instanced(rtf_ctx* ctx, Geometry& g, const rtf_vertex* verts, size_t nv, const uint32_t* tris, size_t nt,
                   const float* matrices, const float* colors, size_t n_inst);
int bake_skinned(rtf_ctx* ctx, Geometry& g, const void* sverts, size_t nv, size_t stride, const uint32_t* tris, size_t nt,
                 const float* bones, size_t n_bones);
int make_probes_device(rtf_ctx* ctx, uint64_t tri_begin, uint64_t tri_end, int front, int back, rtf_illum_sampler* d_out);
int flux_probes_device(rtf_ctx* ctx, const rtf_illum_sampler* d_probes, uint64_t n, float* out8_host);
int flux_mesh_device(rtf_ctx* ctx, const float* d_energy6, uint64_t tri_begin, uint64_t tri_end, int faces, float* out8_host);
// multi.cu
void shard_range(uint64_t n, int nranks, int rank, uint64_t* b, uint64_t* e);
int illum_sharded_device(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push, int front,
                         int back, int mode, float* d_out6, int pre_rc);
int render_camera_sharded(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, int outputs_mask);
// build.cu
int build_bvh(rtf_ctx* ctx, int morton_bits);
int refit_bvh(rtf_ctx* ctx);
// materials.cu
int resolve_materials(rtf_ctx* ctx, std::map<uint64_t, uint32_t>& mat_index, std::vector<DevMaterial>& mats,
                      std::vector<DevBTF>& btfs);
int resolve_environment_map(rtf_ctx* ctx, const DevTexture** dev_faces);
void release_material_tables(rtf_ctx* ctx);
// render.cu
int trace_rays_device(rtf_ctx* ctx, const float4* rays, size_t n, float4* hits, uint32_t* node_visits, uint32_t* tri_visits);
int illum_mesh_device(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push, int front,
                      int back, uint64_t tri_begin, uint64_t tri_end, int s0, int s1, float* d_accum6,
                      const rtf_illum_sampler* d_probes, uint64_t probe_base);
int illum_finalize_device(rtf_ctx* ctx, float* d_accum6, uint64_t n, int faces, int samples, rtf_illum_sampler* d_probes);
int illum_probes_lcg_device(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push,
                            const rtf_illum_sampler* d_probes, uint64_t n, uint64_t first_probe, float* d_accum6);
int scan_grid_device(rtf_ctx* ctx, const rtf_scan_grid* grid, uint64_t ray_begin, uint64_t ray_end, float* d_t,
                     uint32_t* d_prim, uint32_t* d_inst);
void scan_interleave_layout(uint64_t total, int nranks, int rank, uint32_t* shift, uint64_t* n_blocks, uint64_t* n_local);
int scan_grid_interleaved_device(rtf_ctx* ctx, const rtf_scan_grid* grid, int nranks, int rank, bool global_out, float* d_t,
                                 uint32_t* d_prim, uint32_t* d_inst);
int scan_samples_device(rtf_ctx* ctx, rtf_pc_sample* d_samples, size_t n, uint32_t* d_prim);
int render_camera_device(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, uint64_t px_begin,
                         uint64_t px_end);
int render_camera_bands(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, const int* rows, int n_bands);
int render_camera_interleaved_device(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, int band_rows,
                                     int nranks, int rank);
int render_camera_interleaved(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, int band_rows, int nranks,
                              int rank);
int sample_texture_device(rtf_ctx* ctx, const DevTexture& tex, size_t n, const float* d_uv, float* d_rgba);
int sample_environment_device(rtf_ctx* ctx, const rtf_env* env, size_t n, const float* d_pos_dir6, float* d_rgb);
int btf_eval_device(rtf_ctx* ctx, const DevBTF& btf, size_t n, const float* d_in14, float* d_rgb);
int ss_probe_device(rtf_ctx* ctx, const float4* d_rays, const unsigned long long* d_handles, size_t n, uint32_t* d_out17);
int microbench_fp32(rtf_ctx* ctx, double* tflops);
int microbench_l2(rtf_ctx* ctx, double bytes, double* gbs);
}  // namespace rtf
