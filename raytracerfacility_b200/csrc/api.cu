// api.cu — extern "C" boundary of librtf_b200.so (declared in include/rtf/rtf.h).
//
// Mirrors the module façade of the reference (layer L3): CudaModule::{Init, Terminate,
// EstimateIlluminationRayTracing, SamplePointCloud} (/root/reference/src/RayTracerFacility/CUDAModule.cpp:23-64),
// RayTracer::{RenderToCamera, BuildIAS} and the scene maps (RayTracer.hpp:358-394), CameraProperties::{Set, Resize}
// (RayTracer.cu:28-92).  Status codes instead of exceptions / exit(2) (Optix7.hpp:10-56).
#include <cmath>
#include <cstdio>
#include <cstring>

#include "rtf_internal.h"

namespace rtf {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  char buf[512];
  snprintf(buf, sizeof(buf), "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
  g_error = buf;
  return RTF_ERR_CUDA;
}
int flatten_scene(rtf_ctx* ctx);  // build.cu

}  // namespace rtf

using namespace rtf;

#define CHECK_CTX(ctx)                 \
  if (!(ctx)) {                        \
    set_error("null context");         \
    return RTF_ERR_INVALID;            \
  }                                    \
  RTF_CUDA(cudaSetDevice((ctx)->device))

extern "C" {

const char* rtf_last_error(void) { return g_error.c_str(); }
const char* rtf_version(void) { return "rtf-b200 0.1 (sm_100a)"; }

int rtf_create(int device, rtf_ctx** out) {
  if (!out) return RTF_ERR_INVALID;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    set_error("no CUDA device: librtf_b200 has no CPU fallback");
    return RTF_ERR_CUDA;
  }
  if (device < 0 || device >= count) {
    set_error("bad device index");
    return RTF_ERR_INVALID;
  }
  RTF_CUDA(cudaSetDevice(device));
  rtf_ctx* c = new rtf_ctx();
  c->device = device;
  RTF_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  c->own_stream = true;
  cudaDeviceProp prop;
  RTF_CUDA(cudaGetDeviceProperties(&prop, device));
  c->sm_count = prop.multiProcessorCount;
  *out = c;
  return RTF_OK;
}

int rtf_destroy(rtf_ctx* ctx) {
  if (!ctx) return RTF_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (auto& g : ctx->geoms) g.second.d_verts.release(), g.second.d_tris.release();
  DevBuf* bufs[] = {&ctx->d_pos9, &ctx->d_nrm9, &ctx->d_tri_inst, &ctx->d_nodes, &ctx->d_tris48, &ctx->d_insts, &ctx->d_mats,
                    &ctx->d_inst_alpha, &ctx->d_stats, &ctx->ba.morton, &ctx->ba.sorted_keys, &ctx->ba.sorted_ids,
                    &ctx->ba.children, &ctx->ba.range, &ctx->ba.boxes, &ctx->ba.seg, &ctx->ba.wide_src,
                    &ctx->ba.scene_bounds, &ctx->ba.keys_alt, &ctx->ba.ids_alt, &ctx->ba.ref_cnt, &ctx->ba.ref_first,
                    &ctx->ba.ref_tri, &ctx->ba.ref_box, &ctx->ba.scan_state, &ctx->ba.sort_hist, &ctx->ba.sort_status, &ctx->ba.level_a,
                    &ctx->ba.level_b, &ctx->ba.collapse_state, &ctx->ba.slot_pos, &ctx->ws.ray_a[0], &ctx->ws.ray_a[1], &ctx->ws.ray_b[0], &ctx->ws.ray_b[1],
                    &ctx->ws.hits, &ctx->ws.counters, &ctx->ws.thr,
                    &ctx->ws.acc, &ctx->ws.dir0w, &ctx->ws.first_n, &ctx->ws.first_a, &ctx->ws.first_p,
                    &ctx->ws.sort_keys, &ctx->ws.sort_keys_alt, &ctx->ws.sort_vals, &ctx->ws.sort_vals_alt,
                    &ctx->ws.sort_tmp, &ctx->ws.io, &ctx->ws.pend_o, &ctx->ws.pend_d, &ctx->ws.pend_t, &ctx->ws.sort_kin};
  for (DevBuf* b : bufs) b->release();
  ctx->d_btfs.release(), ctx->d_env_faces.release();
  release_material_tables(ctx);
  for (cudaEvent_t e : ctx->build_ev)
    if (e) cudaEventDestroy(e);
  if (ctx->pipe.init) {
    for (int k = 0; k < HostPipe::kDepth; k++) {
      cudaEventDestroy(ctx->pipe.ev_in[k]), cudaEventDestroy(ctx->pipe.ev_done[k]), cudaEventDestroy(ctx->pipe.ev_out[k]);
      ctx->pipe.rays[k].release(), ctx->pipe.hits[k].release(), ctx->pipe.aos[k].release(), ctx->pipe.aux[k].release();
    }
    cudaStreamDestroy(ctx->pipe.in), cudaStreamDestroy(ctx->pipe.out);
  }
  for (cudaEvent_t e : ctx->prof.pool) cudaEventDestroy(e);
  for (cudaEvent_t e : ctx->phase_ev)
    if (e) cudaEventDestroy(e);
  ctx->d_shard_out.release(), ctx->d_small.release();
  if (ctx->comm) rtf_comm_destroy(ctx);
  if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return RTF_OK;
}

int rtf_set_stream(rtf_ctx* ctx, void* cuda_stream) {
  CHECK_CTX(ctx);
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
  if (cuda_stream) {
    ctx->stream = (cudaStream_t)cuda_stream;
    ctx->own_stream = false;
  } else {
    RTF_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    ctx->own_stream = true;
  }
  return RTF_OK;
}
uint64_t rtf_launch_count(rtf_ctx* ctx) { return ctx ? ctx->launches : 0; }

// ---- scene sync --------------------------------------------------------------------------------------
int rtf_set_geometry(rtf_ctx* ctx, uint64_t handle, const void* vertices, size_t n_vertices, size_t stride,
                     const uint32_t* triangles, size_t n_triangles, uint64_t version) {
  CHECK_CTX(ctx);
  if ((!vertices && n_vertices) || (!triangles && n_triangles) || stride < sizeof(rtf_vertex)) {
    set_error("rtf_set_geometry: bad arguments (vertex stride must be >= 80)");
    return RTF_ERR_INVALID;
  }
  for (size_t i = 0; i < 3 * n_triangles; i++)
    if (triangles[i] >= n_vertices) {
      set_error("rtf_set_geometry: triangle index out of range");
      return RTF_ERR_INVALID;
    }
  {  // unchanged version: nothing to do (the caller-side check of RayTracerLayer.cpp:60-75, done here)
    auto it = ctx->geoms.find(handle);
    if (it != ctx->geoms.end() && version != 0 && it->second.version == version && !it->second.baked &&
        it->second.n_verts == n_vertices && it->second.n_tris == n_triangles) {
      ctx->geometry_uploads_skipped++;
      return RTF_OK;
    }
  }
  ctx->geom_dirty = true;
  {
    auto it = ctx->geoms.find(handle);
    if (it == ctx->geoms.end() || it->second.baked || it->second.n_tris != n_triangles) ctx->topo_dirty = true;
  }
  Geometry& g = ctx->geoms[handle];
  g.verts.resize(n_vertices);
  for (size_t i = 0; i < n_vertices; i++) memcpy(&g.verts[i], (const char*)vertices + i * stride, sizeof(rtf_vertex));
  g.tris.assign(triangles, triangles + 3 * n_triangles);
  g.version = version;
  g.uploaded = false, g.baked = false;
  g.n_verts = n_vertices, g.n_tris = n_triangles;
  return RTF_OK;
}
int rtf_set_geometry_instanced(rtf_ctx* ctx, uint64_t handle, const void* vertices, size_t n_vertices, size_t stride,
                               const uint32_t* triangles, size_t n_triangles, const float* matrices, const float* colors,
                               size_t n_instances, uint64_t version) {
  CHECK_CTX(ctx);
  if ((!vertices && n_vertices) || (!triangles && n_triangles) || stride < sizeof(rtf_vertex) || (!matrices && n_instances)) {
    set_error("rtf_set_geometry_instanced: bad arguments (vertex stride must be >= 80)");
    return RTF_ERR_INVALID;
  }
  for (size_t i = 0; i < 3 * n_triangles; i++)
    if (triangles[i] >= n_vertices) {
      set_error("rtf_set_geometry_instanced: triangle index out of range");
      return RTF_ERR_INVALID;
    }
  if ((uint64_t)n_vertices * n_instances >= 0xFFFFFFF0ull || (uint64_t)n_triangles * n_instances * 3 >= 0xFFFFFFF0ull) {
    set_error("rtf_set_geometry_instanced: more than 2^32 baked vertices or indices");
    return RTF_ERR_LIMIT;
  }
  std::vector<rtf_vertex> packed(n_vertices);
  for (size_t i = 0; i < n_vertices; i++) memcpy(&packed[i], (const char*)vertices + i * stride, sizeof(rtf_vertex));
  ctx->geom_dirty = true, ctx->topo_dirty = true;
  if (ctx->geoms.count(handle)) ctx->ready = false, ctx->bvh_ok = false;  // the committed scene may point at the arrays re-baked below
  Geometry& g = ctx->geoms[handle];
  g.verts.clear(), g.tris.clear();
  g.version = version;
  g.uploaded = false, g.baked = false, g.n_verts = g.n_tris = 0;
  return bake_instanced(ctx, g, packed.data(), n_vertices, triangles, n_triangles, matrices, colors, n_instances);
}
int rtf_set_geometry_skinned(rtf_ctx* ctx, uint64_t handle, const void* skinned_vertices, size_t n_vertices, size_t stride,
                             const uint32_t* triangles, size_t n_triangles, const float* bone_matrices, size_t n_bones,
                             uint64_t version) {
  CHECK_CTX(ctx);
  if ((!skinned_vertices && n_vertices) || (!triangles && n_triangles) || stride < sizeof(rtf_skinned_vertex) ||
      (stride & 3) || (!bone_matrices && n_bones)) {
    set_error("rtf_set_geometry_skinned: bad arguments (vertex stride must be a multiple of 4 and >= 144)");
    return RTF_ERR_INVALID;
  }
  for (size_t i = 0; i < 3 * n_triangles; i++)
    if (triangles[i] >= n_vertices) {
      set_error("rtf_set_geometry_skinned: triangle index out of range");
      return RTF_ERR_INVALID;
    }
  ctx->geom_dirty = true, ctx->topo_dirty = true;
  if (ctx->geoms.count(handle)) ctx->ready = false, ctx->bvh_ok = false;  // the committed scene may point at the arrays re-baked below
  Geometry& g = ctx->geoms[handle];
  g.verts.clear(), g.tris.clear();
  g.version = version;
  g.uploaded = false, g.baked = false, g.n_verts = g.n_tris = 0;
  return bake_skinned(ctx, g, skinned_vertices, n_vertices, stride, triangles, n_triangles, bone_matrices, n_bones);
}
int rtf_geometry_read_baked(rtf_ctx* ctx, uint64_t handle, rtf_vertex* verts_out, size_t n_vertices, uint32_t* tris_out,
                            size_t n_triangles) {
  CHECK_CTX(ctx);
  auto it = ctx->geoms.find(handle);
  if (it == ctx->geoms.end()) return RTF_ERR_NOT_FOUND;
  const Geometry& g = it->second;
  if (!g.baked || n_vertices != g.n_verts || n_triangles != g.n_tris) {
    set_error("rtf_geometry_read_baked: not a baked geometry, or counts differ");
    return RTF_ERR_INVALID;
  }
  if (verts_out && n_vertices)
    RTF_CUDA(cudaMemcpyAsync(verts_out, g.d_verts.p, n_vertices * sizeof(rtf_vertex), cudaMemcpyDeviceToHost, ctx->stream));
  if (tris_out && n_triangles)
    RTF_CUDA(cudaMemcpyAsync(tris_out, g.d_tris.p, 12 * n_triangles, cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTF_OK;
}
int rtf_set_material(rtf_ctx* ctx, uint64_t handle, const rtf_material* material) {
  CHECK_CTX(ctx);
  if (!material) return RTF_ERR_INVALID;
  if (material->material_type < RTF_MAT_DEFAULT || material->material_type > RTF_MAT_COMPRESSED_BTF) {
    set_error("rtf_set_material: unknown material_type");
    return RTF_ERR_INVALID;
  }
  auto mit = ctx->mats.find(handle);
  if (mit != ctx->mats.end() && !memcmp(&mit->second, material, sizeof(rtf_material))) return RTF_OK;  // unchanged
  ctx->mats[handle] = *material;
  ctx->mat_dirty = true;
  return RTF_OK;
}
int rtf_set_instance(rtf_ctx* ctx, uint64_t handle, uint64_t geom, uint64_t mat, const float xform[16], uint64_t comp) {
  CHECK_CTX(ctx);
  if (!xform) return RTF_ERR_INVALID;
  auto iit = ctx->insts.find(handle);
  if (iit != ctx->insts.end() && iit->second.geom == geom && iit->second.mat == mat && iit->second.component == comp &&
      !memcmp(iit->second.xform, xform, 64))
    return RTF_OK;  // unchanged (RayTracedInstance::m_version check, RayTracerLayer.cpp:76-92)
  if (iit == ctx->insts.end() || iit->second.geom != geom || iit->second.mat != mat) ctx->topo_dirty = true;
  Instance& in = ctx->insts[handle];
  in.geom = geom, in.mat = mat, in.component = comp;
  memcpy(in.xform, xform, 64);
  ctx->geom_dirty = true;
  return RTF_OK;
}
int rtf_remove_geometry(rtf_ctx* ctx, uint64_t handle) {
  CHECK_CTX(ctx);
  auto it = ctx->geoms.find(handle);
  if (it == ctx->geoms.end()) return RTF_ERR_NOT_FOUND;
  cudaStreamSynchronize(ctx->stream);
  it->second.d_verts.release(), it->second.d_tris.release();
  ctx->geoms.erase(it);
  ctx->geom_dirty = true, ctx->topo_dirty = true;
  ctx->ready = false, ctx->bvh_ok = false;  // the committed scene may point at the released arrays
  return RTF_OK;
}
int rtf_remove_material(rtf_ctx* ctx, uint64_t handle) {
  CHECK_CTX(ctx);
  ctx->mat_links.erase(handle);
  ctx->mat_dirty = true;
  return ctx->mats.erase(handle) ? RTF_OK : RTF_ERR_NOT_FOUND;
}
int rtf_remove_instance(rtf_ctx* ctx, uint64_t handle) {
  CHECK_CTX(ctx);
  ctx->geom_dirty = true, ctx->topo_dirty = true;
  return ctx->insts.erase(handle) ? RTF_OK : RTF_ERR_NOT_FOUND;
}

int rtf_commit(rtf_ctx* ctx, int morton_bits) {
  CHECK_CTX(ctx);
  if (morton_bits == 0) morton_bits = 63;
  if (morton_bits != 30 && morton_bits != 63) {
    set_error("morton_bits must be 30 or 63");
    return RTF_ERR_INVALID;
  }
  // Scene-sync diffing: nothing changed -> nothing to do (the reference only calls BuildIAS when an update flag was
  // raised, RayTracerLayer.cpp:394-400); only material-side tables changed -> they are re-resolved below and the BVH
  // stays, unless an instance's "may ignore hits" class changed (it is baked into the triangle records).
  const bool was_ready = (ctx->ready || ctx->bvh_ok) && ctx->morton_bits == morton_bits && ctx->n_tris > 0;
  if (ctx->ready && was_ready && !ctx->geom_dirty && !ctx->mat_dirty) {
    ctx->commits_skipped++;
    return RTF_OK;
  }
  const bool keep_bvh_if_possible = was_ready && !ctx->geom_dirty;
  const bool refit_if_possible = was_ready && ctx->geom_dirty && !ctx->topo_dirty && ctx->refit && ctx->n_refs > 0 && ctx->refit_ok;
  const uint32_t old_tris = ctx->n_tris;
  ctx->ready = false, ctx->bvh_ok = false;
  int rc;
  cudaStream_t st = ctx->stream;
  // materials -> dense table
  std::map<uint64_t, uint32_t> mat_index;
  std::vector<DevMaterial> mats;
  std::vector<DevBTF> btfs;
  if ((rc = resolve_materials(ctx, mat_index, mats, btfs))) return rc;
  // instances in ascending handle order = global primitive numbering
  std::vector<InstDev> insts;
  std::vector<float> alpha;
  uint64_t first = 0;
  for (auto& kv : ctx->insts) {
    const Instance& in = kv.second;
    auto g = ctx->geoms.find(in.geom);
    auto m = mat_index.find(in.mat);
    if (g == ctx->geoms.end() || m == mat_index.end()) {
      set_error("instance references an unknown geometry or material");
      return RTF_ERR_NOT_FOUND;
    }
    Geometry& geo = g->second;
    if (!geo.uploaded) {
      if ((rc = geo.d_verts.reserve(geo.verts.size() * sizeof(rtf_vertex)))) return rc;
      if ((rc = geo.d_tris.reserve(geo.tris.size() * 4))) return rc;
      RTF_CUDA(cudaMemcpyAsync(geo.d_verts.p, geo.verts.data(), geo.verts.size() * sizeof(rtf_vertex), cudaMemcpyHostToDevice, st));
      RTF_CUDA(cudaMemcpyAsync(geo.d_tris.p, geo.tris.data(), geo.tris.size() * 4, cudaMemcpyHostToDevice, st));
      geo.uploaded = true;
    }
    InstDev d;
    memcpy(d.xform, in.xform, 64);
    d.handle = in.component;
    d.mat = m->second;
    d.first_prim = (uint32_t)first;
    d.n_prims = (uint32_t)geo.n_tris;
    d.pad_ = 0;
    d.verts = geo.d_verts.as<rtf_vertex>();
    d.tris = geo.d_tris.as<uint32_t>();
    if (d.n_prims == 0) continue;  // nothing to trace; keeps first_prim strictly increasing
    first += d.n_prims;
    insts.push_back(d);
    const DevMaterial& mm = mats[d.mat];
    // any-hit alpha (RF.cuh:13-19, RDD.hpp:248-254): Default material only; the albedo texture's w when there is one
    // (flagged as -1: looked up per candidate hit), else 1 - transmission
    float al = 1.0f;
    if (mm.m.material_type == RTF_MAT_DEFAULT) al = mm.tex[MAT_TEX_ALBEDO].texels ? -1.0f : 1.0f - mm.m.transmission;
    alpha.push_back(al);
  }
  if (first >= 0xFFFFFFF0ull) {
    set_error("more than 2^32 triangles");
    return RTF_ERR_LIMIT;
  }
  std::vector<uint8_t> alpha_class(alpha.size());
  for (size_t k = 0; k < alpha.size(); k++) alpha_class[k] = alpha[k] < 1.0f ? 1 : 0;
  const bool keep_bvh = keep_bvh_if_possible && first == ctx->n_tris && alpha_class == ctx->alpha_class;
  ctx->n_tris = (uint32_t)first;
  ctx->scene = DevScene();
  if (first == 0) {  // reference: no acceleration structure -> entry points report "not ready"
    ctx->geom_dirty = true;
    return RTF_OK;
  }
  if ((rc = ctx->d_insts.reserve(insts.size() * sizeof(InstDev)))) return rc;
  if ((rc = ctx->d_mats.reserve(mats.size() * sizeof(DevMaterial)))) return rc;
  if ((rc = ctx->d_btfs.reserve((btfs.size() + 1) * sizeof(DevBTF)))) return rc;
  if (!btfs.empty())
    RTF_CUDA(cudaMemcpyAsync(ctx->d_btfs.p, btfs.data(), btfs.size() * sizeof(DevBTF), cudaMemcpyHostToDevice, st));
  const DevTexture* env_faces = nullptr;
  if ((rc = resolve_environment_map(ctx, &env_faces))) return rc;
  if ((rc = ctx->d_inst_alpha.reserve(alpha.size() * 4))) return rc;
  RTF_CUDA(cudaMemcpyAsync(ctx->d_insts.p, insts.data(), insts.size() * sizeof(InstDev), cudaMemcpyHostToDevice, st));
  RTF_CUDA(cudaMemcpyAsync(ctx->d_mats.p, mats.data(), mats.size() * sizeof(DevMaterial), cudaMemcpyHostToDevice, st));
  RTF_CUDA(cudaMemcpyAsync(ctx->d_inst_alpha.p, alpha.data(), alpha.size() * 4, cudaMemcpyHostToDevice, st));
  if ((rc = ctx->d_pos9.reserve(36 * (size_t)first))) return rc;
  if ((rc = ctx->d_nrm9.reserve(36 * (size_t)first))) return rc;
  if ((rc = ctx->d_tri_inst.reserve(4 * (size_t)first))) return rc;
  ctx->scene.n_insts = (uint32_t)insts.size();
  ctx->scene.insts = ctx->d_insts.as<InstDev>();
  if (keep_bvh) {
    RTF_CUDA(cudaStreamSynchronize(st));  // host vectors above die at return
    ctx->commits_materials++;
  } else {
  if ((rc = flatten_scene(ctx))) return rc;
  RTF_CUDA(cudaStreamSynchronize(st));  // host vectors above die at return
  if (refit_if_possible && first == old_tris && alpha_class == ctx->alpha_class) {
    if ((rc = refit_bvh(ctx))) return rc;
    ctx->commits_refit++;
  } else {
    if ((rc = build_bvh(ctx, morton_bits))) return rc;
    ctx->commits_full++;
  }
  }
  ctx->alpha_class = alpha_class;
  ctx->geom_dirty = false, ctx->mat_dirty = false, ctx->topo_dirty = false;
  ctx->morton_bits = morton_bits;
  DevScene& s = ctx->scene;
  s.nodes = ctx->d_nodes.as<uint4>(), s.tris = ctx->d_tris48.as<float4>();
  s.pos9 = ctx->d_pos9.as<float>(), s.nrm9 = ctx->d_nrm9.as<float>(), s.tri_inst = ctx->d_tri_inst.as<uint32_t>();
  s.mats = ctx->d_mats.as<DevMaterial>(), s.inst_alpha = ctx->d_inst_alpha.as<float>();
  s.btfs = ctx->d_btfs.as<DevBTF>(), s.env_faces = env_faces;
  s.n_tris = ctx->n_tris, s.n_nodes = ctx->n_nodes;
  s.has_alpha = 0;
  for (float a : alpha)
    if (a < 1.0f) s.has_alpha = 1;  // includes the "textured alpha" marker (-1)
  s.has_btf = 0;
  for (const DevMaterial& dm : mats)
    if (dm.m.material_type == RTF_MAT_COMPRESSED_BTF) s.has_btf = 1;
  s.has_sss = 0;
  for (const DevMaterial& dm : mats)
    if (dm.m.material_type != RTF_MAT_COMPRESSED_BTF && dm.m.subsurface_factor > 0.0f && dm.m.subsurface_radius[0] > 0.0f)
      s.has_sss = 1;
  ctx->ready = true, ctx->bvh_ok = true;
  ctx->epoch++;
  return RTF_OK;
}
int rtf_commit_stats(rtf_ctx* ctx, uint64_t* out5) {
  if (!ctx || !out5) return RTF_ERR_INVALID;
  out5[0] = ctx->commits_full, out5[1] = ctx->commits_materials, out5[2] = ctx->commits_skipped;
  out5[3] = ctx->geometry_uploads_skipped, out5[4] = ctx->commits_refit;
  return RTF_OK;
}
int rtf_build_timings(rtf_ctx* ctx, float* ms_out, int n) {
  if (!ctx || !ms_out || n < 6) return RTF_ERR_INVALID;
  for (int k = 0; k < 6; k++) ms_out[k] = ctx->build_ms[k];
  return RTF_OK;
}
int rtf_set_option(rtf_ctx* ctx, const char* name, int64_t value) {
  if (!ctx || !name) return RTF_ERR_INVALID;
  if (!strcmp(name, "batch_paths")) {
    if (value < 1024 || value > ((int64_t)1 << 28)) {
      set_error("batch_paths must be in [2^10, 2^28]");
      return RTF_ERR_INVALID;
    }
    ctx->batch_paths = (size_t)value;
  } else if (!strcmp(name, "pipe_chunk_rays")) {
    if (value < 1024 || value > ((int64_t)1 << 28)) {
      set_error("pipe_chunk_rays must be in [2^10, 2^28]");
      return RTF_ERR_INVALID;
    }
    ctx->pipe_chunk_rays = (size_t)value;
  } else if (!strcmp(name, "presplit")) {
    if (value < 1 || value > 64) {
      set_error("presplit must be in [1, 64]");
      return RTF_ERR_INVALID;
    }
    if (ctx->presplit != (int)value) ctx->geom_dirty = true, ctx->topo_dirty = true;
    ctx->presplit = (int)value;  // takes effect at the next rtf_commit
  } else if (!strcmp(name, "brdf")) {
    if (value != RTF_BRDF_REFERENCE && value != RTF_BRDF_LAMBERT_GGX) {
      set_error("brdf must be RTF_BRDF_REFERENCE (0) or RTF_BRDF_LAMBERT_GGX (1)");
      return RTF_ERR_INVALID;
    }
    ctx->brdf = (int)value;
  } else if (!strcmp(name, "pipe_chunk_probes") || !strcmp(name, "pipe_chunk_samples")) {
    if (value < 1 || value > ((int64_t)1 << 28)) {
      set_error("pipe chunk sizes must be in [1, 2^28]");
      return RTF_ERR_INVALID;
    }
    (name[11] == 'p' ? ctx->pipe_chunk_probes : ctx->pipe_chunk_samples) = (size_t)value;
  } else if (!strcmp(name, "profile")) {  // per-category kernel time (rtf_profile_read); switching resets the sums
    cudaStreamSynchronize(ctx->stream);
    ctx->prof.on = value != 0;
    ctx->prof.recs.clear(), ctx->prof.used = 0;
  } else if (!strcmp(name, "refit")) {
    ctx->refit = value != 0;
  } else if (!strcmp(name, "sort_rays")) {
    ctx->sort_rays = value != 0;
  } else if (!strcmp(name, "rng")) {
    if (value != RTF_RNG_PHILOX && value != RTF_RNG_REFERENCE_LCG) {
      set_error("rng must be RTF_RNG_PHILOX or RTF_RNG_REFERENCE_LCG");
      return RTF_ERR_INVALID;
    }
    ctx->rng = (int)value;
  } else {
    set_error(std::string("unknown option ") + name);
    return RTF_ERR_INVALID;
  }
  return RTF_OK;
}
// Device milliseconds per kernel category since the last read (option "profile" = 1): {ray generation, traversal,
// shading, reduction / resolve, ray sort, other}; the records are consumed.
int rtf_profile_read(rtf_ctx* ctx, float* ms6) {
  CHECK_CTX(ctx);
  if (!ms6) return RTF_ERR_INVALID;
  for (int k = 0; k < PROF_COUNT; k++) ms6[k] = 0.0f;
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  for (const Profiler::Rec& r : ctx->prof.recs) {
    float ms = 0.0f;
    if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess && r.cat >= 0 && r.cat < PROF_COUNT) ms6[r.cat] += ms;
  }
  ctx->prof.recs.clear(), ctx->prof.used = 0;
  return RTF_OK;
}
int rtf_triangle_count(rtf_ctx* ctx, uint64_t* n) {
  if (!ctx || !n) return RTF_ERR_INVALID;
  *n = ctx->n_tris;
  return RTF_OK;
}

static int fetch_stats(rtf_ctx* ctx) {
  unsigned long long s[2] = {0, 0};
  if (ctx->d_stats.p) RTF_CUDA(cudaMemcpy(s, ctx->d_stats.p, 16, cudaMemcpyDeviceToHost));
  ctx->last_rays = s[0];
  return RTF_OK;
}
static int pipe_init(rtf_ctx* ctx);
static int not_ready() {
  set_error("no acceleration structure (call rtf_commit)");
  return RTF_ERR_NOT_READY;
}

// ---- illumination ----------------------------------------------------------------------------------------
// CudaModule::EstimateIlluminationRayTracing (CUDAModule.cpp:39-51, RayTracer.cu:606-653): AoS probes in host memory,
// updated in place.  The reference uploads all probes, launches, synchronises and downloads them, serially.  Here probe
// chunks run as a three-stage pipeline — H2D of chunk i+1 (copy-in stream), path tracing of chunk i (context stream),
// D2H of chunk i-1 (copy-out stream) — so both PCIe directions and the SMs are busy at once.  A probe's RNG streams are
// keyed by its index in the caller's array, so chunking does not change a single bit of the result.
int rtf_estimate_illumination(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, rtf_illum_sampler* probes, size_t n,
                              uint32_t seed, float push) {
  CHECK_CTX(ctx);
  if (!env || !rp || (!probes && n)) return RTF_ERR_INVALID;
  if (!ctx->ready) return not_ready();  // reference: silent return (RayTracer.cu:610-611)
  if (n == 0) return RTF_OK;            // reference: message + return (RayTracer.cu:612-615)
  int rc;
  cudaStream_t st = ctx->stream;
  if (ctx->rng == RTF_RNG_REFERENCE_LCG) {  // verification mode: the reference's stream, samples in sequence (render.cu)
    if ((rc = ctx->ws.io.reserve(n * (sizeof(rtf_illum_sampler) + 24)))) return rc;
    if ((rc = ctx->d_stats.reserve(64))) return rc;
    rtf_illum_sampler* d_probes = ctx->ws.io.as<rtf_illum_sampler>();
    float* d_acc = (float*)(d_probes + n);
    RTF_CUDA(cudaMemcpyAsync(d_probes, probes, n * sizeof(rtf_illum_sampler), cudaMemcpyHostToDevice, st));
    RTF_CUDA(cudaMemsetAsync(d_acc, 0, n * 24, st));
    if ((rc = illum_probes_lcg_device(ctx, env, rp, seed, push, d_probes, n, 0, d_acc))) return rc;
    if ((rc = illum_finalize_device(ctx, d_acc, n, 2, rp->samples, d_probes))) return rc;
    RTF_CUDA(cudaMemcpyAsync(probes, d_probes, n * sizeof(rtf_illum_sampler), cudaMemcpyDeviceToHost, st));
    RTF_CUDA(cudaStreamSynchronize(st));
    return fetch_stats(ctx);
  }
  if ((rc = pipe_init(ctx))) return rc;
  HostPipe& p = ctx->pipe;
  const size_t chunk = ctx->pipe_chunk_probes ? ctx->pipe_chunk_probes : 1;
  const size_t cm = n < chunk ? n : chunk;
  for (int k = 0; k < HostPipe::kDepth; k++)
    if ((rc = p.aos[k].reserve(cm * sizeof(rtf_illum_sampler))) || (rc = p.aux[k].reserve(cm * 24))) return rc;
  RTF_CUDA(cudaStreamSynchronize(st));  // earlier work of the caller on the context stream
  if ((rc = ctx->d_stats.reserve(64))) return rc;  // first job of a context: the wavefront workspace does not exist yet
  RTF_CUDA(cudaMemsetAsync(ctx->d_stats.p, 0, 16, st));
  ctx->stats_hold = true, ctx->last_paths = 0;
  size_t i = 0;
  for (size_t off = 0; off < n && !rc; off += chunk, i++) {
    const size_t m = (n - off) < chunk ? (n - off) : chunk;
    const int b = (int)(i % HostPipe::kDepth);
    rtf_illum_sampler* d_probes = p.aos[b].as<rtf_illum_sampler>();
    float* d_acc = p.aux[b].as<float>();
    if (i >= (size_t)HostPipe::kDepth) cudaStreamWaitEvent(p.in, p.ev_out[b], 0);  // the chunk before has left aos[b]
    cudaMemcpyAsync(d_probes, probes + off, m * sizeof(rtf_illum_sampler), cudaMemcpyHostToDevice, p.in);
    cudaEventRecord(p.ev_in[b], p.in);
    cudaStreamWaitEvent(st, p.ev_in[b], 0);
    cudaMemsetAsync(d_acc, 0, m * 24, st);
    if ((rc = illum_mesh_device(ctx, env, rp, seed, push, 1, 1, off, off + m, 0, rp->samples, d_acc, d_probes, off))) break;
    if ((rc = illum_finalize_device(ctx, d_acc, m, 2, rp->samples, d_probes))) break;
    cudaEventRecord(p.ev_done[b], st);
    cudaStreamWaitEvent(p.out, p.ev_done[b], 0);
    cudaMemcpyAsync(probes + off, d_probes, m * sizeof(rtf_illum_sampler), cudaMemcpyDeviceToHost, p.out);
    cudaEventRecord(p.ev_out[b], p.out);
  }
  ctx->stats_hold = false;
  cudaStreamSynchronize(p.in);
  cudaStreamSynchronize(p.out);
  RTF_CUDA(cudaStreamSynchronize(st));
  if (rc) return rc;
  RTF_CUDA(cudaGetLastError());
  return fetch_stats(ctx);
}

int rtf_estimate_illumination_mesh_device(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push,
                                          int front, int back, uint64_t tri_begin, uint64_t tri_end, int s0, int s1,
                                          float* dev_accum6) {
  CHECK_CTX(ctx);
  if (!env || !rp || !dev_accum6) return RTF_ERR_INVALID;
  if (!ctx->ready) return not_ready();
  if (tri_end > ctx->n_tris || tri_begin > tri_end || s0 < 0 || s1 > rp->samples || s0 > s1) {
    set_error("triangle or sample range out of bounds");
    return RTF_ERR_INVALID;
  }
  return illum_mesh_device(ctx, env, rp, seed, push, front, back, tri_begin, tri_end, s0, s1, dev_accum6 + 6 * tri_begin,
                           nullptr, 0);
}
int rtf_illumination_finalize_device(rtf_ctx* ctx, float* dev_accum6, uint64_t n, int faces, int samples) {
  CHECK_CTX(ctx);
  return illum_finalize_device(ctx, dev_accum6, n, faces, samples, nullptr);
}
int rtf_estimate_illumination_mesh(rtf_ctx* ctx, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push,
                                   int front, int back, uint64_t tri_begin, uint64_t tri_end, int s0, int s1, int normalise,
                                   float* out6) {
  CHECK_CTX(ctx);
  if (!env || !rp || !out6) return RTF_ERR_INVALID;
  if (!ctx->ready) return not_ready();
  if (tri_end > ctx->n_tris || tri_begin > tri_end || s0 < 0 || s1 > rp->samples || s0 > s1) {
    set_error("triangle or sample range out of bounds");
    return RTF_ERR_INVALID;
  }
  const uint64_t n = tri_end - tri_begin;
  if (n == 0) return RTF_OK;
  int rc;
  if ((rc = ctx->ws.io.reserve(n * 24))) return rc;
  float* d_acc = ctx->ws.io.as<float>();
  RTF_CUDA(cudaMemsetAsync(d_acc, 0, n * 24, ctx->stream));
  if ((rc = illum_mesh_device(ctx, env, rp, seed, push, front, back, tri_begin, tri_end, s0, s1, d_acc, nullptr, 0))) return rc;
  if (normalise)
    if ((rc = illum_finalize_device(ctx, d_acc, n, (front ? 1 : 0) + (back ? 1 : 0), s1 - s0, nullptr))) return rc;
  RTF_CUDA(cudaMemcpyAsync(out6, d_acc, n * 24, cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  return fetch_stats(ctx);
}
int rtf_prepare_light_probes(rtf_ctx* ctx, uint64_t tri_begin, uint64_t tri_end, int front, int back, rtf_illum_sampler* out) {
  CHECK_CTX(ctx);
  if (!ctx->ready) return not_ready();
  if (tri_end > ctx->n_tris || tri_begin > tri_end || (!out && tri_end > tri_begin)) {
    set_error("triangle range out of bounds");
    return RTF_ERR_INVALID;
  }
  const uint64_t n = tri_end - tri_begin;
  if (n == 0) return RTF_OK;
  int rc;
  if ((rc = ctx->ws.io.reserve(n * sizeof(rtf_illum_sampler)))) return rc;
  if ((rc = make_probes_device(ctx, tri_begin, tri_end, front, back, ctx->ws.io.as<rtf_illum_sampler>()))) return rc;
  RTF_CUDA(cudaMemcpyAsync(out, ctx->ws.io.p, n * sizeof(rtf_illum_sampler), cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTF_OK;
}
int rtf_illumination_flux(rtf_ctx* ctx, const rtf_illum_sampler* probes, size_t n, float* out8) {
  CHECK_CTX(ctx);
  if (!out8 || (!probes && n)) return RTF_ERR_INVALID;
  int rc;
  if ((rc = ctx->ws.io.reserve(n * sizeof(rtf_illum_sampler) + 16))) return rc;
  if (n) RTF_CUDA(cudaMemcpyAsync(ctx->ws.io.p, probes, n * sizeof(rtf_illum_sampler), cudaMemcpyHostToDevice, ctx->stream));
  return flux_probes_device(ctx, ctx->ws.io.as<rtf_illum_sampler>(), n, out8);
}
int rtf_illumination_flux_mesh_device(rtf_ctx* ctx, const float* dev_energy6, uint64_t tri_begin, uint64_t tri_end, int front,
                                      int back, float* out8) {
  CHECK_CTX(ctx);
  if (!out8 || !dev_energy6) return RTF_ERR_INVALID;
  if (!ctx->ready) return not_ready();
  if (tri_end > ctx->n_tris || tri_begin > tri_end) {
    set_error("triangle range out of bounds");
    return RTF_ERR_INVALID;
  }
  return flux_mesh_device(ctx, dev_energy6, tri_begin, tri_end, (front ? 1 : 0) + (back ? 1 : 0), out8);
}
int rtf_last_job_stats(rtf_ctx* ctx, uint64_t* rays, uint64_t* paths) {
  CHECK_CTX(ctx);
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  int rc = ctx->stats_cached ? RTF_OK : fetch_stats(ctx);
  if (rc) return rc;
  if (rays) *rays = ctx->last_rays;
  if (paths) *paths = ctx->last_paths;
  return RTF_OK;
}

// ---- LiDAR --------------------------------------------------------------------------------------------------
// CudaModule::SamplePointCloud (CUDAModule.cpp:53-64, RayTracer.cu:758-795): AoS samples in host memory, updated in
// place.  224 B cross PCIe per ray (112 up, 112 down) against ~0.3 ns of traversal, so the call is a copy pipeline:
// chunk i+1 goes up while chunk i is traced and chunk i-1 comes down (full duplex), instead of the reference's serial
// upload -> launch -> synchronise -> download.
int rtf_scan_point_cloud(rtf_ctx* ctx, const rtf_env* env, rtf_pc_sample* samples, size_t n, uint32_t* prim_ids_out) {
  CHECK_CTX(ctx);
  (void)env;  // the reference passes it but PCS never reads it (RayTracer.cu:758-795)
  if (!samples && n) return RTF_ERR_INVALID;
  if (!ctx->ready) return not_ready();
  if (n == 0) return RTF_OK;
  if (n >= 0xFFFFFFFFull) {
    set_error("use rtf_scan_grid_trace for > 2^32 rays");
    return RTF_ERR_LIMIT;
  }
  int rc;
  cudaStream_t st = ctx->stream;
  if ((rc = pipe_init(ctx))) return rc;
  HostPipe& p = ctx->pipe;
  const size_t chunk = ctx->pipe_chunk_samples ? ctx->pipe_chunk_samples : 1;
  const size_t cm = n < chunk ? n : chunk;
  for (int k = 0; k < HostPipe::kDepth; k++)
    if ((rc = p.aos[k].reserve(cm * sizeof(rtf_pc_sample))) || (rc = p.aux[k].reserve(cm * 4))) return rc;
  RTF_CUDA(cudaStreamSynchronize(st));
  size_t i = 0;
  for (size_t off = 0; off < n && !rc; off += chunk, i++) {
    const size_t m = (n - off) < chunk ? (n - off) : chunk;
    const int b = (int)(i % HostPipe::kDepth);
    rtf_pc_sample* d = p.aos[b].as<rtf_pc_sample>();
    uint32_t* d_prim = p.aux[b].as<uint32_t>();
    if (i >= (size_t)HostPipe::kDepth) cudaStreamWaitEvent(p.in, p.ev_out[b], 0);
    cudaMemcpyAsync(d, samples + off, m * sizeof(rtf_pc_sample), cudaMemcpyHostToDevice, p.in);
    cudaEventRecord(p.ev_in[b], p.in);
    cudaStreamWaitEvent(st, p.ev_in[b], 0);
    if ((rc = scan_samples_device(ctx, d, m, d_prim))) break;
    cudaEventRecord(p.ev_done[b], st);
    cudaStreamWaitEvent(p.out, p.ev_done[b], 0);
    cudaMemcpyAsync(samples + off, d, m * sizeof(rtf_pc_sample), cudaMemcpyDeviceToHost, p.out);
    if (prim_ids_out) cudaMemcpyAsync(prim_ids_out + off, d_prim, m * 4, cudaMemcpyDeviceToHost, p.out);
    cudaEventRecord(p.ev_out[b], p.out);
  }
  cudaStreamSynchronize(p.in);
  cudaStreamSynchronize(p.out);
  RTF_CUDA(cudaStreamSynchronize(st));
  if (rc) return rc;
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}
// column = uint(size.x / distance.x), row = uint(size.y / distance.y) (PointCloudScanner.cpp:71-75); the float -> unsigned
// conversion is only defined for finite quotients in [0, 2^32)
int rtf_scan_grid_dims(const rtf_scan_grid* g, uint32_t* column, uint32_t* row) {
  if (!g || !column || !row) return RTF_ERR_INVALID;
  for (int k = 0; k < 2; k++) {
    const float q = g->size[k] / g->distance[k];
    if (!(g->distance[k] > 0.0f) || !(g->size[k] >= 0.0f) || !(q < 4294967296.0f)) {
      set_error("scan grid: distance must be > 0, size >= 0, both finite, and size / distance < 2^32");
      return RTF_ERR_INVALID;
    }
  }
  *column = (uint32_t)(g->size[0] / g->distance[0]);
  *row = (uint32_t)(g->size[1] / g->distance[1]);
  return RTF_OK;
}
int rtf_scan_grid_ray_count(const rtf_scan_grid* g, uint64_t* n) {
  if (!g || !n) return RTF_ERR_INVALID;
  uint32_t column = 0, row = 0;
  int rc = rtf_scan_grid_dims(g, &column, &row);
  if (rc) return rc;
  *n = (uint64_t)column * (uint64_t)row;
  return RTF_OK;
}
int rtf_scan_grid_trace_device(rtf_ctx* ctx, const rtf_scan_grid* grid, uint64_t b, uint64_t e, float* dev_t, uint32_t* dev_prim) {
  CHECK_CTX(ctx);
  if (!grid || (!dev_t && !dev_prim)) {
    set_error("rtf_scan_grid_trace_device: null grid, or no output array");
    return RTF_ERR_INVALID;
  }
  return scan_grid_device(ctx, grid, b, e, dev_t, dev_prim, nullptr);
}
int rtf_scan_grid_trace(rtf_ctx* ctx, const rtf_scan_grid* grid, uint64_t b, uint64_t e, float* t_out, uint32_t* prim_out,
                        uint32_t* inst_out) {
  CHECK_CTX(ctx);
  if (!grid) return RTF_ERR_INVALID;
  if (!ctx->ready) return not_ready();
  if (e <= b) return RTF_OK;
  const uint64_t n = e - b;
  int rc;
  if ((rc = ctx->ws.io.reserve(n * 12))) return rc;
  float* dt = ctx->ws.io.as<float>();
  uint32_t* dp = (uint32_t*)(dt + n);
  uint32_t* di = dp + n;
  if ((rc = scan_grid_device(ctx, grid, b, e, dt, dp, di))) return rc;
  if (t_out) RTF_CUDA(cudaMemcpyAsync(t_out, dt, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (prim_out) RTF_CUDA(cudaMemcpyAsync(prim_out, dp, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (inst_out) RTF_CUDA(cudaMemcpyAsync(inst_out, di, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTF_OK;
}

// ---- raw closest hit ------------------------------------------------------------------------------------------
int rtf_trace_closest_device(rtf_ctx* ctx, const float* dev_rays, size_t n, void* dev_hits) {
  CHECK_CTX(ctx);
  if ((!dev_rays || !dev_hits) && n) return RTF_ERR_INVALID;
  return trace_rays_device(ctx, (const float4*)dev_rays, n, (float4*)dev_hits, nullptr, nullptr);
}
int rtf_trace_closest_counts(rtf_ctx* ctx, const float* rays, size_t n, void* hits, uint32_t* node_visits, uint32_t* tri_visits) {
  CHECK_CTX(ctx);
  if ((!rays || !hits) && n) return RTF_ERR_INVALID;
  if (!ctx->ready) return not_ready();
  if (n == 0) return RTF_OK;
  int rc;
  const size_t rb = n * 32, hb = n * 16, cb = n * 4;
  if ((rc = ctx->ws.io.reserve(rb + hb + 2 * cb))) return rc;
  float4* dr = ctx->ws.io.as<float4>();
  float4* dh = (float4*)((char*)ctx->ws.io.p + rb);
  uint32_t* dn = (uint32_t*)((char*)ctx->ws.io.p + rb + hb);
  uint32_t* dtv = dn + n;
  const bool counts = node_visits || tri_visits;
  RTF_CUDA(cudaMemcpyAsync(dr, rays, rb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = trace_rays_device(ctx, dr, n, dh, counts ? dn : nullptr, counts ? dtv : nullptr))) return rc;
  RTF_CUDA(cudaMemcpyAsync(hits, dh, hb, cudaMemcpyDeviceToHost, ctx->stream));
  if (node_visits) RTF_CUDA(cudaMemcpyAsync(node_visits, dn, cb, cudaMemcpyDeviceToHost, ctx->stream));
  if (tri_visits) RTF_CUDA(cudaMemcpyAsync(tri_visits, dtv, cb, cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTF_OK;
}
static int pipe_init(rtf_ctx* ctx) {
  HostPipe& p = ctx->pipe;
  if (p.init) return RTF_OK;
  RTF_CUDA(cudaStreamCreateWithFlags(&p.in, cudaStreamNonBlocking));
  RTF_CUDA(cudaStreamCreateWithFlags(&p.out, cudaStreamNonBlocking));
  for (int k = 0; k < HostPipe::kDepth; k++) {
    RTF_CUDA(cudaEventCreateWithFlags(&p.ev_in[k], cudaEventDisableTiming));
    RTF_CUDA(cudaEventCreateWithFlags(&p.ev_done[k], cudaEventDisableTiming));
    RTF_CUDA(cudaEventCreateWithFlags(&p.ev_out[k], cudaEventDisableTiming));
  }
  p.init = true;
  return RTF_OK;
}
// Host rays in, host hits out.  Large batches run as a three-stage pipeline over chunks: H2D of chunk i+1 (copy-in
// stream), traversal of chunk i (context stream) and D2H of chunk i-1 (copy-out stream) overlap; PCIe is full duplex,
// so the job is bounded by the larger of the two copy directions instead of their sum plus the kernel.  Pinned host
// buffers give true overlap; pageable ones still work (the runtime stages them).
int rtf_trace_closest(rtf_ctx* ctx, const float* rays, size_t n, void* hits) {
  CHECK_CTX(ctx);
  if ((!rays || !hits) && n) return RTF_ERR_INVALID;
  if (!ctx->ready) return not_ready();
  if (n == 0) return RTF_OK;
  const size_t chunk = ctx->pipe_chunk_rays;
  if (n <= chunk) return rtf_trace_closest_counts(ctx, rays, n, hits, nullptr, nullptr);
  int rc;
  if ((rc = pipe_init(ctx))) return rc;
  HostPipe& p = ctx->pipe;
  for (int k = 0; k < HostPipe::kDepth; k++)
    if ((rc = p.rays[k].reserve(chunk * 32)) || (rc = p.hits[k].reserve(chunk * 16))) return rc;
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));  // earlier work of the caller on the context stream
  size_t i = 0;
  for (size_t off = 0; off < n; off += chunk, i++) {
    const size_t m = (n - off) < chunk ? (n - off) : chunk;
    const int b = (int)(i % HostPipe::kDepth);
    if (i >= (size_t)HostPipe::kDepth) RTF_CUDA(cudaStreamWaitEvent(p.in, p.ev_done[b], 0));  // the chunk before no longer reads rays[b]
    RTF_CUDA(cudaMemcpyAsync(p.rays[b].p, rays + off * 8, m * 32, cudaMemcpyHostToDevice, p.in));
    RTF_CUDA(cudaEventRecord(p.ev_in[b], p.in));
    RTF_CUDA(cudaStreamWaitEvent(ctx->stream, p.ev_in[b], 0));
    if (i >= (size_t)HostPipe::kDepth) RTF_CUDA(cudaStreamWaitEvent(ctx->stream, p.ev_out[b], 0));  // hits[b] of the chunk before has left the device
    if ((rc = trace_rays_device(ctx, p.rays[b].as<float4>(), m, p.hits[b].as<float4>(), nullptr, nullptr))) return rc;
    RTF_CUDA(cudaEventRecord(p.ev_done[b], ctx->stream));
    RTF_CUDA(cudaStreamWaitEvent(p.out, p.ev_done[b], 0));
    RTF_CUDA(cudaMemcpyAsync((char*)hits + off * 16, p.hits[b].p, m * 16, cudaMemcpyDeviceToHost, p.out));
    RTF_CUDA(cudaEventRecord(p.ev_out[b], p.out));
  }
  RTF_CUDA(cudaStreamSynchronize(p.out));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTF_OK;
}

// ---- camera -------------------------------------------------------------------------------------------------------
int rtf_camera_create(rtf_ctx* ctx, rtf_camera** out) {
  CHECK_CTX(ctx);
  if (!out) return RTF_ERR_INVALID;
  rtf_camera* c = new rtf_camera();
  c->ctx = ctx;
  memset(&c->props, 0, sizeof(c->props));
  c->props.accumulate = 1, c->props.fov = 120.0f, c->props.max_distance = 50.0f, c->props.gamma = 2.2f;  // RayTracer.hpp:48-61
  c->props.focal_length = 1.0f;
  *out = c;
  return RTF_OK;
}
int rtf_camera_destroy(rtf_camera* cam) {
  if (!cam) return RTF_OK;
  cudaSetDevice(cam->ctx->device);
  cudaStreamSynchronize(cam->ctx->stream);
  cam->color.release(), cam->normal.release(), cam->albedo.release();
  delete cam;
  return RTF_OK;
}
int rtf_camera_get_props(rtf_camera* cam, rtf_camera_props** props) {
  if (!cam || !props) return RTF_ERR_INVALID;
  *props = &cam->props;
  return RTF_OK;
}
int rtf_camera_resize(rtf_camera* cam, int width, int height) {  // CameraProperties::Resize, RayTracer.cu:48-92
  if (!cam || width < 0 || height < 0) return RTF_ERR_INVALID;
  if (cam->props.width == width && cam->props.height == height) return RTF_OK;
  RTF_CUDA(cudaSetDevice(cam->ctx->device));
  RTF_CUDA(cudaStreamSynchronize(cam->ctx->stream));
  cam->props.width = width, cam->props.height = height, cam->props.modified = 1;
  const size_t bytes = (size_t)width * height * 16;
  int rc;
  if ((rc = cam->color.reserve(bytes)) || (rc = cam->normal.reserve(bytes)) || (rc = cam->albedo.reserve(bytes))) return rc;
  return RTF_OK;
}

static void mat4_mul_d(const double* a, const double* b, double* c) {  // column-major c = a*b
  for (int col = 0; col < 4; col++)
    for (int row = 0; row < 4; row++) {
      double s = 0;
      for (int k = 0; k < 4; k++) s += a[k * 4 + row] * b[col * 4 + k];
      c[col * 4 + row] = s;
    }
}
static bool mat4_inverse_d(const double* m, double* inv) {
  double a[4][8];
  for (int r = 0; r < 4; r++)
    for (int c = 0; c < 4; c++) a[r][c] = m[c * 4 + r], a[r][4 + c] = r == c ? 1.0 : 0.0;
  for (int i = 0; i < 4; i++) {
    int piv = i;
    for (int r = i + 1; r < 4; r++)
      if (fabs(a[r][i]) > fabs(a[piv][i])) piv = r;
    if (a[piv][i] == 0.0) return false;
    if (piv != i)
      for (int c = 0; c < 8; c++) std::swap(a[i][c], a[piv][c]);
    const double d = a[i][i];
    for (int c = 0; c < 8; c++) a[i][c] /= d;
    for (int r = 0; r < 4; r++)
      if (r != i) {
        const double f = a[r][i];
        for (int c = 0; c < 8; c++) a[r][c] -= f * a[i][c];
      }
  }
  for (int r = 0; r < 4; r++)
    for (int c = 0; c < 4; c++) inv[c * 4 + r] = a[r][4 + c];
  return true;
}
// CameraProperties::Set, RayTracer.cu:28-46.  Host-side setup math is done in double and rounded once.
int rtf_camera_set(rtf_camera* cam, const float position[3], const float q[4]) {
  if (!cam || !position || !q) return RTF_ERR_INVALID;
  rtf_camera_props& p = cam->props;
  if (p.width <= 0 || p.height <= 0) {
    set_error("resize the camera before Set");
    return RTF_ERR_INVALID;
  }
  const double w = q[0], x = q[1], y = q[2], z = q[3];
  auto rot = [&](const double v[3], double o[3]) {  // glm quat * vec3
    const double uv[3] = {y * v[2] - z * v[1], z * v[0] - x * v[2], x * v[1] - y * v[0]};
    const double uuv[3] = {y * uv[2] - z * uv[1], z * uv[0] - x * uv[2], x * uv[1] - y * uv[0]};
    for (int k = 0; k < 3; k++) o[k] = v[k] + (uv[k] * w + uuv[k]) * 2.0;
  };
  auto norm = [](double v[3]) {
    const double l = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    v[0] /= l, v[1] /= l, v[2] /= l;
  };
  const double f0[3] = {0, 0, -1}, u0[3] = {0, 1, 0};
  double front[3], up[3];
  rot(f0, front), rot(u0, up);
  norm(front), norm(up);
  const double aspect = (double)p.width / (double)p.height;
  const double half_fov = (double)(p.fov * 0.5f) * (M_PI / 180.0);  // radians(m_fov * 0.5f): the halved fov quirk
  const double th = tan(half_fov / 2.0), zn = 0.1, zf = 100.0;
  double proj[16] = {0};
  proj[0] = 1.0 / (aspect * th), proj[5] = 1.0 / th, proj[10] = -(zf + zn) / (zf - zn), proj[11] = -1.0;
  proj[14] = -(2.0 * zf * zn) / (zf - zn);
  const double eye[3] = {position[0], position[1], position[2]};
  double s[3] = {front[1] * up[2] - front[2] * up[1], front[2] * up[0] - front[0] * up[2], front[0] * up[1] - front[1] * up[0]};
  norm(s);
  const double u[3] = {s[1] * front[2] - s[2] * front[1], s[2] * front[0] - s[0] * front[2], s[0] * front[1] - s[1] * front[0]};
  double view[16] = {0};
  view[0] = s[0], view[4] = s[1], view[8] = s[2];
  view[1] = u[0], view[5] = u[1], view[9] = u[2];
  view[2] = -front[0], view[6] = -front[1], view[10] = -front[2];
  view[12] = -(s[0] * eye[0] + s[1] * eye[1] + s[2] * eye[2]);
  view[13] = -(u[0] * eye[0] + u[1] * eye[1] + u[2] * eye[2]);
  view[14] = front[0] * eye[0] + front[1] * eye[1] + front[2] * eye[2];
  view[15] = 1.0;
  double pv[16], inv[16];
  mat4_mul_d(proj, view, pv);
  if (!mat4_inverse_d(pv, inv)) {
    set_error("singular projection*view");
    return RTF_ERR_INVALID;
  }
  bool changed = false;
  for (int k = 0; k < 16; k++) {
    const float v = (float)inv[k];
    if (v != p.inverse_projection_view[k]) changed = true;
    p.inverse_projection_view[k] = v;
  }
  if (changed) p.modified = 1;
  for (int k = 0; k < 3; k++) p.from[k] = position[k];
  double hv[3] = {front[1] * up[2] - front[2] * up[1], front[2] * up[0] - front[0] * up[2], front[0] * up[1] - front[1] * up[0]};
  norm(hv);
  for (int k = 0; k < 3; k++) {
    p.horizontal[k] = (float)(half_fov * aspect * hv[k]);  // "cosFovY" is the angle in radians, RayTracer.cu:40-45
    p.vertical[k] = (float)(half_fov * up[k]);
  }
  return RTF_OK;
}

}  // extern "C"
namespace rtf {
// RenderToCamera (RayTracer.cu:244-603) for a list of row bands [rows[2k], rows[2k+1]) of ONE frame: the change
// detection runs once, every band sees the same frame id, the id advances once.  Returns 1 / 0 like the reference's bool
// (0 = no acceleration structure), < 0 on error.  Blocks until the bands are rendered.
// RenderToCamera's bookkeeping around the launch (RayTracer.cu:244-269, 349-352): 1 = render, 0 / 1-without-work as the
// reference returns them.
static int camera_job_begin(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, bool* skip) {
  rtf_camera_props& p = cam->props;
  *skip = true;
  if (p.width == 0 || p.height == 0) return 1;  // RayTracer.cu:246-247
  if (!ctx->ready) return 0;                    // RayTracer.cu:248-249
  *skip = false;
  bool changed = p.modified != 0;
  p.modified = 0;
  if (!ctx->cam_last_valid || memcmp(&ctx->cam_last_env, env, sizeof(rtf_env)) != 0) ctx->cam_last_env = *env, changed = true;
  if (!ctx->cam_last_valid || ctx->cam_last_rp.bounces != rp->bounces || ctx->cam_last_rp.samples != rp->samples)
    ctx->cam_last_rp = *rp, changed = true;
  // m_requireUpdate: BuildIAS / an SBT update restarts the accumulation (RayTracer.cu:254, RayTracerLayer.cpp:382-389)
  if (ctx->cam_last_epoch != ctx->epoch) ctx->cam_last_epoch = ctx->epoch, changed = true;
  ctx->cam_last_valid = true;
  if (!p.accumulate || changed) p.frame_id = 0;  // RayTracer.cu:266-269
  return 1;
}
static int camera_job_end(rtf_ctx* ctx, rtf_camera* cam, uint64_t paths) {
  cam->props.frame_id++;  // the launch saw the frame id before the increment (RayTracer.cu:349-352)
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  unsigned long long st[2] = {0, 0};
  if (ctx->d_stats.p) RTF_CUDA(cudaMemcpy(st, ctx->d_stats.p, 16, cudaMemcpyDeviceToHost));
  ctx->last_rays = st[0];
  ctx->last_paths = paths;
  ctx->stats_cached = true;
  return 1;
}
int render_camera_bands(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, const int* rows, int n_bands) {
  if (!env || !cam || !rp || (!rows && n_bands)) return RTF_ERR_INVALID;
  rtf_camera_props& p = cam->props;
  for (int k = 0; k < n_bands; k++)
    if (rows[2 * k] < 0 || (p.height > 0 && rows[2 * k + 1] > p.height) || rows[2 * k] > rows[2 * k + 1]) {
      set_error("rtf_render_camera_rows: bad row range");
      return RTF_ERR_INVALID;
    }
  bool skip;
  const int st = camera_job_begin(ctx, env, cam, rp, &skip);
  if (skip) return st;
  int rc;
  if ((rc = ctx->d_stats.reserve(64))) return rc;
  RTF_CUDA(cudaMemsetAsync(ctx->d_stats.p, 0, 16, ctx->stream));  // the job counters run over all bands
  uint64_t paths = 0;
  ctx->stats_hold = true;
  for (int k = 0; k < n_bands && !rc; k++) {
    rc = render_camera_device(ctx, env, cam, rp, (uint64_t)rows[2 * k] * p.width, (uint64_t)rows[2 * k + 1] * p.width);
    paths += ctx->last_paths;
  }
  ctx->stats_hold = false;
  if (rc) return rc;
  return camera_job_end(ctx, cam, paths);
}
int render_camera_interleaved(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, int band_rows, int nranks,
                              int rank) {
  if (!env || !cam || !rp) return RTF_ERR_INVALID;
  bool skip;
  const int st = camera_job_begin(ctx, env, cam, rp, &skip);
  if (skip) return st;
  int rc;
  if ((rc = render_camera_interleaved_device(ctx, env, cam, rp, band_rows, nranks, rank))) return rc;
  return camera_job_end(ctx, cam, ctx->last_paths);
}
}  // namespace rtf
extern "C" {
static int render_camera_rows(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, int row_begin,
                              int row_end) {
  const int rows[2] = {row_begin, row_end};
  return render_camera_bands(ctx, env, cam, rp, rows, 1);
}
int rtf_render_camera(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp) {
  CHECK_CTX(ctx);
  return render_camera_rows(ctx, env, cam, rp, 0, cam ? cam->props.height : 0);
}
int rtf_render_camera_rows(rtf_ctx* ctx, const rtf_env* env, rtf_camera* cam, const rtf_ray_props* rp, int row_begin,
                           int row_end) {
  CHECK_CTX(ctx);
  return render_camera_rows(ctx, env, cam, rp, row_begin, row_end);
}
int rtf_camera_device_buffer(rtf_camera* cam, int which, float** dev) {
  if (!cam || !dev) return RTF_ERR_INVALID;
  *dev = which == RTF_OUT_COLOR ? cam->color.as<float>() : (which == RTF_OUT_NORMAL ? cam->normal.as<float>() : cam->albedo.as<float>());
  return RTF_OK;
}
int rtf_camera_download(rtf_camera* cam, int which, float* host) {
  if (!cam || !host) return RTF_ERR_INVALID;
  RTF_CUDA(cudaSetDevice(cam->ctx->device));
  float* d = nullptr;
  rtf_camera_device_buffer(cam, which, &d);
  const size_t bytes = (size_t)cam->props.width * cam->props.height * 16;
  RTF_CUDA(cudaMemcpyAsync(host, d, bytes, cudaMemcpyDeviceToHost, cam->ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(cam->ctx->stream));
  return RTF_OK;
}

// The inverse of rtf_camera_download: restores a frame buffer (e.g. the colour buffer of an earlier session, so that
// progressive accumulation — CameraRendering.cu:112-128 blends into what the buffer holds — can go on).
int rtf_camera_upload(rtf_camera* cam, int which, const float* host) {
  if (!cam || !host) return RTF_ERR_INVALID;
  RTF_CUDA(cudaSetDevice(cam->ctx->device));
  float* d = nullptr;
  rtf_camera_device_buffer(cam, which, &d);
  const size_t bytes = (size_t)cam->props.width * cam->props.height * 16;
  if (!d || bytes == 0) return RTF_ERR_INVALID;
  RTF_CUDA(cudaMemcpyAsync(d, host, bytes, cudaMemcpyHostToDevice, cam->ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(cam->ctx->stream));
  return RTF_OK;
}

// ---- parity read-back ----------------------------------------------------------------------------------------------
static const void* dbg_ptr(rtf_ctx* c, int what, uint64_t* bytes) {
  const uint64_t nt = c->n_tris, n = c->n_refs, ni = n ? n - 1 : 0;  // n: triangle references
  switch (what) {
    case RTF_DBG_REF_TRI: *bytes = n * 4; return c->ba.ref_tri.p;
    case RTF_DBG_REF_BOX: *bytes = n * 24; return c->ba.ref_box.p;
    case RTF_DBG_MORTON: *bytes = n * 8; return c->ba.morton.p;
    case RTF_DBG_SORTED_IDS: *bytes = n * 4; return c->ba.sorted_ids.p;
    case RTF_DBG_BIN_CHILDREN: *bytes = ni * 8; return c->ba.children.p;
    case RTF_DBG_BIN_BOXES: *bytes = (n ? 2 * n - 1 : 0) * 24; return c->ba.boxes.p;
    case RTF_DBG_WIDE_NODES: *bytes = (uint64_t)c->n_nodes * 80; return c->d_nodes.p;
    case RTF_DBG_WIDE_TRIS: *bytes = n * 48; return c->d_tris48.p;
    case RTF_DBG_FLAT_POS: *bytes = nt * 36; return c->d_pos9.p;
    case RTF_DBG_FLAT_NRM: *bytes = nt * 36; return c->d_nrm9.p;
    case RTF_DBG_FLAT_INST: *bytes = nt * 4; return c->d_tri_inst.p;
    case RTF_DBG_SCENE_BOUNDS: *bytes = 24; return c->ba.scene_bounds.p ? (char*)c->ba.scene_bounds.p + 32 : nullptr;
  }
  *bytes = 0;
  return nullptr;
}
int rtf_debug_size(rtf_ctx* ctx, int what, uint64_t* bytes) {
  if (!ctx || !bytes) return RTF_ERR_INVALID;
  dbg_ptr(ctx, what, bytes);
  return RTF_OK;
}
int rtf_debug_read(rtf_ctx* ctx, int what, void* out, uint64_t bytes) {
  CHECK_CTX(ctx);
  uint64_t b = 0;
  const void* p = dbg_ptr(ctx, what, &b);
  if (bytes < b || (b && !p)) return RTF_ERR_INVALID;
  if (b) {
    RTF_CUDA(cudaStreamSynchronize(ctx->stream));
    RTF_CUDA(cudaMemcpy(out, p, b, cudaMemcpyDeviceToHost));
  }
  return RTF_OK;
}
int rtf_debug_upload_bvh(rtf_ctx* ctx, const void* nodes, uint64_t n_nodes, const void* tris, uint64_t n_tris, int depth) {
  CHECK_CTX(ctx);
  if (!nodes || !tris || depth > kMaxStackDepth) return RTF_ERR_INVALID;
  int rc;
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  if ((rc = ctx->d_nodes.reserve(n_nodes * 80)) || (rc = ctx->d_tris48.reserve(n_tris * 48))) return rc;
  RTF_CUDA(cudaMemcpy(ctx->d_nodes.p, nodes, n_nodes * 80, cudaMemcpyHostToDevice));
  RTF_CUDA(cudaMemcpy(ctx->d_tris48.p, tris, n_tris * 48, cudaMemcpyHostToDevice));
  ctx->scene.nodes = ctx->d_nodes.as<uint4>(), ctx->scene.tris = ctx->d_tris48.as<float4>();
  ctx->n_nodes = (uint32_t)n_nodes, ctx->depth = depth;
  ctx->scene.n_nodes = (uint32_t)n_nodes;
  if (!ctx->ready) ctx->n_tris = (uint32_t)n_tris, ctx->scene.n_tris = (uint32_t)n_tris;
  ctx->n_refs = (uint32_t)n_tris;
  ctx->refit_ok = false;  // the builder's arrays do not describe an uploaded tree
  ctx->ready = true;
  return RTF_OK;
}
int rtf_sss_probe(rtf_ctx* ctx, const float* rays, const uint64_t* handles, size_t n, uint32_t* out17) {
  CHECK_CTX(ctx);
  if ((!rays || !handles || !out17) && n) return RTF_ERR_INVALID;
  if (!ctx->ready) return not_ready();
  if (n == 0) return RTF_OK;
  int rc;
  const size_t rb = n * 32, hb = n * 8, ob = n * 68;
  if ((rc = ctx->ws.io.reserve(rb + hb + ob + 512))) return rc;
  char* base = ctx->ws.io.as<char>();
  float4* d_rays = (float4*)base;
  unsigned long long* d_h = (unsigned long long*)(base + ((rb + 255) & ~(size_t)255));
  uint32_t* d_out = (uint32_t*)((char*)d_h + ((hb + 255) & ~(size_t)255));
  RTF_CUDA(cudaMemcpyAsync(d_rays, rays, rb, cudaMemcpyHostToDevice, ctx->stream));
  RTF_CUDA(cudaMemcpyAsync(d_h, handles, hb, cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = ss_probe_device(ctx, d_rays, d_h, n, d_out))) return rc;
  RTF_CUDA(cudaMemcpyAsync(out17, d_out, ob, cudaMemcpyDeviceToHost, ctx->stream));
  RTF_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTF_OK;
}
int rtf_microbench_fp32(rtf_ctx* ctx, double* tflops) {
  CHECK_CTX(ctx);
  return microbench_fp32(ctx, tflops);
}
int rtf_microbench_l2(rtf_ctx* ctx, double bytes, double* gbs) {
  CHECK_CTX(ctx);
  return microbench_l2(ctx, bytes, gbs);
}

}  // extern "C"
