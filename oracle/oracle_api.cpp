// oracle_api.cpp — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// extern "C" surface of the CPU oracle, loaded with ctypes by tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs.  The product (raytracerfacility_b200/) never links it.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "oracle_prebake.h"
#include "oracle_shade.h"

using namespace orc;

#define ORC_API extern "C" __attribute__((visibility("default")))

template <class F>
static void parallel_for(uint64_t n, int nthreads, F fn) {
  if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
  if (nthreads < 1) nthreads = 1;
  if ((uint64_t)nthreads > n) nthreads = n ? (int)n : 1;
  if (nthreads == 1) {
    fn(0, 0, n);
    return;
  }
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; t++) {
    uint64_t b = n * (uint64_t)t / (uint64_t)nthreads, e = n * (uint64_t)(t + 1) / (uint64_t)nthreads;
    th.emplace_back([=] { fn(t, b, e); });
  }
  for (auto& x : th) x.join();
}

ORC_API const char* orc_version() { return "rtf-oracle 0.2 (shading, helpers, RNG, BTF pinned by the reference's own code compiled on the host; traversal unpinned by the reference: OptiX is closed, pinned by brute force)"; }
ORC_API int orc_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

// ---- primitives -----------------------------------------------------------------------------
ORC_API void orc_philox(const uint32_t* ctr, const uint32_t* key, uint32_t* out) { philox4x32_10(ctr, key, out); }
ORC_API void orc_lcg(uint32_t v0, uint32_t v1, int n, float* out, uint32_t* state_after_init) {
  LCG g;
  g.init(v0, v1);
  if (state_after_init) *state_after_init = g.state;
  for (int i = 0; i < n; i++) out[i] = g.next();
}
// Test switch, see g_libm_trig in oracle_math.h.  Returns the previous value.
ORC_API int orc_set_libm_trig(int on) {
  const int was = g_libm_trig;
  g_libm_trig = on;
  return was;
}
ORC_API void orc_sincos2pi(int n, const float* u, float* s, float* c) {
  for (int i = 0; i < n; i++) sincos2pi(u[i], &s[i], &c[i]);
}
ORC_API void orc_path_rng(uint32_t seed, uint32_t stream, uint64_t unit, uint32_t bounce, uint32_t extra, float* r4) {
  path_rng(seed, stream, unit, bounce, extra, r4);
}
ORC_API void orc_sample_hemisphere(float r0, float r1, const float* n, float* out) {
  V3 d = sample_hemisphere(r0, r1, v3(n[0], n[1], n[2]));
  out[0] = d.x, out[1] = d.y, out[2] = d.z;
}
ORC_API void orc_sample_lobe(float r0, float r1, const float* n, float alpha, float* out) {
  V3 d = sample_lobe(r0, r1, v3(n[0], n[1], n[2]), alpha);
  out[0] = d.x, out[1] = d.y, out[2] = d.z;
}
ORC_API void orc_reflect(const float* i, const float* n, float* out) {
  V3 d = reflect(v3(i[0], i[1], i[2]), v3(n[0], n[1], n[2]));
  out[0] = d.x, out[1] = d.y, out[2] = d.z;
}
ORC_API void orc_tangent_space(const float* n, float* out9) {
  V3 nn = v3(n[0], n[1], n[2]), t, b;
  tangent_space(nn, t, b);
  out9[0] = t.x, out9[1] = t.y, out9[2] = t.z, out9[3] = b.x, out9[4] = b.y, out9[5] = b.z;
  out9[6] = nn.x, out9[7] = nn.y, out9[8] = nn.z;
}
ORC_API int orc_cube_map_lookup(const float* d, float* uv) { return cube_map_lookup(v3(d[0], d[1], d[2]), uv[0], uv[1]); }
ORC_API int orc_ray_sphere(const float* o, const float* d, float radius, float* t01) {
  return ray_sphere(v3(o[0], o[1], o[2]), v3(d[0], d[1], d[2]), radius, t01[0], t01[1]) ? 1 : 0;
}
ORC_API void orc_env_light(const rtf_env* env, const float* pos, const float* dir, float* out) {
  V3 c = env_light(v3(pos[0], pos[1], pos[2]), v3(dir[0], dir[1], dir[2]), *env);
  out[0] = c.x, out[1] = c.y, out[2] = c.z;
}

// ---- scene ----------------------------------------------------------------------------------
ORC_API Scene* orc_scene_create() { return new Scene(); }
ORC_API void orc_scene_destroy(Scene* s) { delete s; }
ORC_API int orc_scene_add_geometry(Scene* s, const void* verts, uint64_t nv, uint64_t stride, const uint32_t* tris,
                                   uint64_t nt) {
  Geometry g;
  g.verts.resize(nv);
  for (uint64_t i = 0; i < nv; i++) memcpy(&g.verts[i], (const char*)verts + i * stride, sizeof(rtf_vertex));
  g.tris.assign(tris, tris + 3 * nt);
  s->geoms.push_back(std::move(g));
  return (int)s->geoms.size() - 1;
}
ORC_API int orc_scene_add_material(Scene* s, const rtf_material* m) {
  s->mats.push_back(*m);
  return (int)s->mats.size() - 1;
}
// Instances must be added in ascending handle order (the global primitive numbering).
ORC_API int orc_scene_add_instance(Scene* s, int geom, int mat, const float* xform16, uint64_t handle) {
  Instance in;
  in.geom = (uint32_t)geom, in.mat = (uint32_t)mat, in.handle = handle, in.first_prim = 0;
  memcpy(in.xform, xform16, 64);
  s->insts.push_back(in);
  return (int)s->insts.size() - 1;
}
ORC_API void orc_scene_commit(Scene* s, int bits, int presplit) { commit(*s, bits == 30 ? 30 : 63, presplit < 1 ? 1 : presplit); }
ORC_API uint64_t orc_scene_nrefs(Scene* s) { return s->bvh.nref; }
ORC_API void orc_scene_set_brdf(Scene* s, int brdf) { s->brdf = brdf; }
ORC_API uint64_t orc_scene_ntris(Scene* s) { return s->tri_inst.size(); }
ORC_API uint64_t orc_scene_nnodes(Scene* s) { return s->bvh.nodes.size(); }
ORC_API int orc_scene_depth(Scene* s) { return s->bvh.depth; }

static const void* dbg_ptr(Scene* s, int what, uint64_t* bytes) {
  BVH& b = s->bvh;
  switch (what) {
    case RTF_DBG_MORTON: *bytes = b.morton.size() * 8; return b.morton.data();
    case RTF_DBG_SORTED_IDS: *bytes = b.order.size() * 4; return b.order.data();
    case RTF_DBG_BIN_CHILDREN: *bytes = b.children.size() * 4; return b.children.data();
    case RTF_DBG_BIN_PARENT: *bytes = b.parent.size() * 4; return b.parent.data();
    case RTF_DBG_BIN_BOXES: *bytes = b.boxes.size() * 24; return b.boxes.data();
    case RTF_DBG_WIDE_NODES: *bytes = b.nodes.size() * 80; return b.nodes.data();
    case RTF_DBG_WIDE_TRIS: *bytes = b.tris.size() * 48; return b.tris.data();
    case RTF_DBG_FLAT_POS: *bytes = s->pos9.size() * 4; return s->pos9.data();
    case RTF_DBG_FLAT_NRM: *bytes = s->nrm9.size() * 4; return s->nrm9.data();
    case RTF_DBG_FLAT_INST: *bytes = s->tri_inst.size() * 4; return s->tri_inst.data();
    case RTF_DBG_SCENE_BOUNDS: *bytes = 24; return &b.scene;
    case RTF_DBG_REF_TRI: *bytes = b.ref_tri.size() * 4; return b.ref_tri.data();
    case RTF_DBG_REF_BOX: *bytes = b.ref_box.size() * 24; return b.ref_box.data();
  }
  *bytes = 0;
  return nullptr;
}
ORC_API uint64_t orc_debug_size(Scene* s, int what) {
  uint64_t b = 0;
  dbg_ptr(s, what, &b);
  return b;
}
ORC_API int orc_debug_read(Scene* s, int what, void* out, uint64_t bytes) {
  uint64_t b = 0;
  const void* p = dbg_ptr(s, what, &b);
  if (!p && b) return -1;
  if (bytes < b) return -1;
  if (b) memcpy(out, p, b);
  return 0;
}

// ---- closest hit ----------------------------------------------------------------------------
// rays: n x {o.xyz, tmin, d.xyz, tmax}; hits: n x {t,u,v,prim}
ORC_API void orc_trace(Scene* s, const float* rays, uint64_t n, void* hits, uint32_t* node_visits, uint32_t* tri_visits,
                       int nthreads) {
  Hit* H = (Hit*)hits;
  parallel_for(n, nthreads, [=](int, uint64_t b, uint64_t e) {
    for (uint64_t i = b; i < e; i++) {
      const float* r = rays + 8 * i;
      Ray ray{v3(r[0], r[1], r[2]), r[3], v3(r[4], r[5], r[6]), r[7]};
      TraceStats st;
      H[i] = trace(s->bvh, ray, &st);
      if (node_visits) node_visits[i] = st.nodes;
      if (tri_visits) tri_visits[i] = st.tris;
    }
  });
}
ORC_API void orc_trace_brute(Scene* s, const float* rays, uint64_t n, void* hits, int nthreads) {
  Hit* H = (Hit*)hits;
  parallel_for(n, nthreads, [=](int, uint64_t b, uint64_t e) {
    for (uint64_t i = b; i < e; i++) {
      const float* r = rays + 8 * i;
      Ray ray{v3(r[0], r[1], r[2]), r[3], v3(r[4], r[5], r[6]), r[7]};
      H[i] = trace_brute(s->bvh, ray);
    }
  });
}

// ---- illumination ---------------------------------------------------------------------------
struct JobStats {
  uint64_t rays, nodes, tris, paths;
};
// Committed triangles [tri_begin, tri_end) with step tri_stride are the probes; out6 holds one
// record per visited triangle (dense, in visiting order).
ORC_API void orc_illum_mesh(Scene* s, const rtf_env* env, const rtf_ray_props* rp, uint32_t seed, float push, int front,
                            int back, uint64_t tri_begin, uint64_t tri_end, uint64_t tri_stride, int s0, int s1,
                            int normalise, float* out6, JobStats* stats, int nthreads) {
  if (tri_stride < 1) tri_stride = 1;
  uint64_t cnt = tri_end > tri_begin ? (tri_end - tri_begin + tri_stride - 1) / tri_stride : 0;
  int nt = nthreads <= 0 ? (int)std::thread::hardware_concurrency() : nthreads;
  std::vector<JobStats> per(nt > 0 ? nt : 1, JobStats{0, 0, 0, 0});
  JobStats* perp = per.data();
  parallel_for(cnt, nt, [=](int tid, uint64_t b, uint64_t e) {
    for (uint64_t k = b; k < e; k++) {
      uint64_t t = tri_begin + k * tri_stride;
      const float* p = &s->pos9[9 * t];
      const float* n = &s->nrm9[9 * t];
      V3 P[3] = {v3(p[0], p[1], p[2]), v3(p[3], p[4], p[5]), v3(p[6], p[7], p[8])};
      V3 N[3] = {v3(n[0], n[1], n[2]), v3(n[3], n[4], n[5]), v3(n[6], n[7], n[8])};
      TraceStats st;
      uint64_t rays = 0;
      illum_probe(*s, *env, *rp, seed, push, t, P, N, front != 0, back != 0, s0, s1, out6 + 6 * k, &rays, &st);
      if (normalise) illum_finalize(out6 + 6 * k, (front ? 1 : 0) + (back ? 1 : 0), s1 - s0);
      perp[tid].rays += rays, perp[tid].nodes += st.nodes, perp[tid].tris += st.tris;
      perp[tid].paths += (uint64_t)((front ? 1 : 0) + (back ? 1 : 0)) * (uint64_t)(s1 - s0);
    }
  });
  if (stats) {
    *stats = JobStats{0, 0, 0, 0};
    for (auto& x : per) stats->rays += x.rays, stats->nodes += x.nodes, stats->tris += x.tris, stats->paths += x.paths;
  }
}
// orc_illum_probes with the REFERENCE's random stream: one LCG per probe seeded (probe index, seed) as __raygen__IE does
// (IE.cu:47), drawn sequentially in the reference's order.  Test mode only (see t_seq_rng, oracle_math.h): compared with
// the reference's own device programs run on the host (ref_optix_driver.cpp).  BSSRDF materials: the draws are made at
// the reference's positions (oracle_shade.h, `deferred`), but the spatial probe itself is the repo's (4 nearest), which
// equals the reference's only where at most one same-handle surface lies in the probe's range.  Single thread.
ORC_API int orc_illum_probes_lcg(Scene* s, const rtf_env* env, const rtf_ray_props* rp, rtf_illum_sampler* probes, uint64_t n,
                                 uint32_t seed, float push, uint64_t* rays_out) {
  uint64_t rays = 0;
  for (uint64_t i = 0; i < n; i++) {
    rtf_illum_sampler& pr = probes[i];
    V3 P[3] = {v3(pr.a.position[0], pr.a.position[1], pr.a.position[2]), v3(pr.b.position[0], pr.b.position[1], pr.b.position[2]),
               v3(pr.c.position[0], pr.c.position[1], pr.c.position[2])};
    V3 N[3] = {v3(pr.a.normal[0], pr.a.normal[1], pr.a.normal[2]), v3(pr.b.normal[0], pr.b.normal[1], pr.b.normal[2]),
               v3(pr.c.normal[0], pr.c.normal[1], pr.c.normal[2])};
    LCG g;
    g.init((uint32_t)i, seed);
    t_seq_rng = &g;
    g_libm_trig = 1;  // the BTF evaluator's angles through libm too, as the reference computes them
    float o6[6];
    bool f = pr.front_face != 0, bk = pr.back_face != 0;
    illum_probe(*s, *env, *rp, seed, push, i, P, N, f, bk, 0, rp->samples, o6, &rays, nullptr);
    t_seq_rng = nullptr;
    g_libm_trig = 0;
    int faces = (f ? 1 : 0) + (bk ? 1 : 0);
    if (faces * rp->samples != 0) {
      illum_finalize(o6, faces, rp->samples);
      for (int k = 0; k < 3; k++) pr.energy[k] = o6[k], pr.direction[k] = o6[3 + k];
    }
  }
  if (rays_out) *rays_out = rays;
  return 0;
}
// CudaModule::EstimateIlluminationRayTracing semantics over the 268 B AoS (CUDAModule.cpp:39-50).
ORC_API void orc_illum_probes(Scene* s, const rtf_env* env, const rtf_ray_props* rp, rtf_illum_sampler* probes, uint64_t n,
                              uint32_t seed, float push, JobStats* stats, int nthreads) {
  int nt = nthreads <= 0 ? (int)std::thread::hardware_concurrency() : nthreads;
  std::vector<JobStats> per(nt > 0 ? nt : 1, JobStats{0, 0, 0, 0});
  JobStats* perp = per.data();
  parallel_for(n, nt, [=](int tid, uint64_t b, uint64_t e) {
    for (uint64_t i = b; i < e; i++) {
      rtf_illum_sampler& pr = probes[i];
      V3 P[3] = {v3(pr.a.position[0], pr.a.position[1], pr.a.position[2]),
                 v3(pr.b.position[0], pr.b.position[1], pr.b.position[2]),
                 v3(pr.c.position[0], pr.c.position[1], pr.c.position[2])};
      V3 N[3] = {v3(pr.a.normal[0], pr.a.normal[1], pr.a.normal[2]), v3(pr.b.normal[0], pr.b.normal[1], pr.b.normal[2]),
                 v3(pr.c.normal[0], pr.c.normal[1], pr.c.normal[2])};
      float o6[6];
      TraceStats st;
      uint64_t rays = 0;
      bool f = pr.front_face != 0, bk = pr.back_face != 0;
      illum_probe(*s, *env, *rp, seed, push, i, P, N, f, bk, 0, rp->samples, o6, &rays, &st);
      int faces = (f ? 1 : 0) + (bk ? 1 : 0);
      if (faces * rp->samples != 0) {  // IE.cu:125: outputs untouched when sampleSize == 0
        illum_finalize(o6, faces, rp->samples);
        for (int k = 0; k < 3; k++) pr.energy[k] = o6[k], pr.direction[k] = o6[3 + k];
      }
      perp[tid].rays += rays, perp[tid].nodes += st.nodes, perp[tid].tris += st.tris;
      perp[tid].paths += (uint64_t)faces * (uint64_t)rp->samples;
    }
  });
  if (stats) {
    *stats = JobStats{0, 0, 0, 0};
    for (auto& x : per) stats->rays += x.rays, stats->nodes += x.nodes, stats->tris += x.tris, stats->paths += x.paths;
  }
}

// ---- LiDAR ----------------------------------------------------------------------------------
ORC_API void orc_scan_point_cloud(Scene* s, rtf_pc_sample* samples, uint64_t n, uint32_t* prim_out, int nthreads) {
  parallel_for(n, nthreads, [=](int, uint64_t b, uint64_t e) {
    for (uint64_t i = b; i < e; i++) {
      rtf_pc_sample& sm = samples[i];
      V3 d = v3(sm.direction[0], sm.direction[1], sm.direction[2]);
      Ray ray{v3(sm.start[0], sm.start[1], sm.start[2]), 1e-3f, d, 1e20f};  // PCS.cu:58-62
      Hit h = trace(s->bvh, ray);
      if (prim_out) prim_out[i] = h.prim;
      if (h.prim == 0xFFFFFFFFu) {  // __miss__PCS_R, PCS.cu:36-40 (+ HitInfo() defaults, PCS.cu:55)
        sm.hit = 0, sm.handle = 0;
        memset(&sm.hit_info, 0, sizeof(sm.hit_info));
        for (int k = 0; k < 4; k++) sm.hit_info.color[k] = 1.0f;
      } else {
        sm.hit = 1;
        sm.handle = s->insts[s->tri_inst[h.prim]].handle;
        full_hit_info(*s, h, d, sm.hit_info);
      }
    }
  });
}
// PointCloudScanner::Scan ray grid, PointCloudScanner.cpp:70-92.
static inline void grid_dims(const rtf_scan_grid& g, uint32_t& column, uint32_t& row) {
  column = (uint32_t)(g.size[0] / g.distance[0]);
  row = (uint32_t)(g.size[1] / g.distance[1]);
}
static inline Ray grid_ray(const rtf_scan_grid& g, uint32_t column, uint32_t row, uint64_t i) {
  const int column_start = -(int)(column / 2);
  const int row_start = -(int)(row / 2);
  const int ci = (int)(i / row), ri = (int)(i % row);
  V3 center = v3(g.center[0], g.center[1], g.center[2]);
  V3 left = v3(g.left[0], g.left[1], g.left[2]), up = v3(g.up[0], g.up[1], g.up[2]);
  V3 pos = (center + (left * (float)(column_start + ci)) * g.distance[0]) + (up * (float)(row_start + ri)) * g.distance[1];
  V3 dir = normalize(v3(g.direction[0], g.direction[1], g.direction[2]));
  return Ray{pos, 1e-3f, dir, 1e20f};
}
ORC_API uint64_t orc_scan_grid_count(const rtf_scan_grid* g) {
  uint32_t c, r;
  grid_dims(*g, c, r);
  return (uint64_t)c * (uint64_t)r;
}
ORC_API void orc_scan_grid(Scene* s, const rtf_scan_grid* g, uint64_t ray_begin, uint64_t ray_end, uint64_t stride,
                           float* t_out, uint32_t* prim_out, uint32_t* inst_out, JobStats* stats, int nthreads) {
  uint32_t column, row;
  grid_dims(*g, column, row);
  if (stride < 1) stride = 1;
  uint64_t cnt = ray_end > ray_begin ? (ray_end - ray_begin + stride - 1) / stride : 0;
  int nt = nthreads <= 0 ? (int)std::thread::hardware_concurrency() : nthreads;
  std::vector<JobStats> per(nt > 0 ? nt : 1, JobStats{0, 0, 0, 0});
  JobStats* perp = per.data();
  parallel_for(cnt, nt, [=](int tid, uint64_t b, uint64_t e) {
    for (uint64_t k = b; k < e; k++) {
      Ray ray = grid_ray(*g, column, row, ray_begin + k * stride);
      TraceStats st;
      Hit h = trace(s->bvh, ray, &st);
      if (t_out) t_out[k] = h.t;
      if (prim_out) prim_out[k] = h.prim;
      if (inst_out) inst_out[k] = h.prim == 0xFFFFFFFFu ? 0xFFFFFFFFu : s->tri_inst[h.prim];
      perp[tid].rays++, perp[tid].nodes += st.nodes, perp[tid].tris += st.tris;
    }
  });
  if (stats) {
    *stats = JobStats{0, 0, 0, 0};
    for (auto& x : per) stats->rays += x.rays, stats->nodes += x.nodes, stats->tris += x.tris;
  }
}
ORC_API void orc_scan_grid_rays(const rtf_scan_grid* g, uint64_t ray_begin, uint64_t ray_end, float* rays8) {
  uint32_t column, row;
  grid_dims(*g, column, row);
  for (uint64_t i = ray_begin; i < ray_end; i++) {
    Ray r = grid_ray(*g, column, row, i);
    float* o = rays8 + 8 * (i - ray_begin);
    o[0] = r.o.x, o[1] = r.o.y, o[2] = r.o.z, o[3] = r.tmin, o[4] = r.d.x, o[5] = r.d.y, o[6] = r.d.z, o[7] = r.tmax;
  }
}

// ---- camera ---------------------------------------------------------------------------------
// Pixels pixel_begin + k*stride (k = 0..) of the W*H raster (index = x + W*y).  prev_color: the
// colour buffer before this frame (float4 per visited pixel, may be NULL).  Outputs float4 per visited pixel.
ORC_API void orc_camera(Scene* s, const rtf_env* env, const rtf_ray_props* rp, const rtf_camera_props* cam,
                        uint64_t pixel_begin, uint64_t pixel_end, uint64_t stride, const float* prev_color, float* color,
                        float* normal, float* albedo, JobStats* stats, int nthreads) {
  if (stride < 1) stride = 1;
  uint64_t cnt = pixel_end > pixel_begin ? (pixel_end - pixel_begin + stride - 1) / stride : 0;
  int nt = nthreads <= 0 ? (int)std::thread::hardware_concurrency() : nthreads;
  std::vector<JobStats> per(nt > 0 ? nt : 1, JobStats{0, 0, 0, 0});
  JobStats* perp = per.data();
  parallel_for(cnt, nt, [=](int tid, uint64_t b, uint64_t e) {
    for (uint64_t k = b; k < e; k++) {
      uint64_t px = pixel_begin + k * stride;
      int ix = (int)(px % (uint64_t)cam->width), iy = (int)(px / (uint64_t)cam->width);
      TraceStats st;
      uint64_t rays = 0;
      PixelOut po = camera_pixel(*s, *env, *rp, *cam, ix, iy, &rays, &st);
      const float zero[4] = {0, 0, 0, 0};
      camera_resolve(*cam, po, prev_color ? prev_color + 4 * k : zero, color + 4 * k, normal + 4 * k, albedo + 4 * k);
      perp[tid].rays += rays, perp[tid].nodes += st.nodes, perp[tid].tris += st.tris, perp[tid].paths += rp->samples;
    }
  });
  if (stats) {
    *stats = JobStats{0, 0, 0, 0};
    for (auto& x : per) stats->rays += x.rays, stats->nodes += x.nodes, stats->tris += x.tris, stats->paths += x.paths;
  }
}

// orc_camera over the whole frame with the REFERENCE's random stream: one LCG per pixel seeded (ix + width * iy, frame id)
// as __raygen__CR does (CR.cu:42-44), drawn sequentially.  Test mode only (see orc_illum_probes_lcg).  color is in-out.
ORC_API int orc_camera_lcg(Scene* s, const rtf_env* env, const rtf_ray_props* rp, const rtf_camera_props* cam, float* color,
                           float* normal, float* albedo, uint64_t* rays_out) {
  uint64_t rays = 0;
  for (int iy = 0; iy < cam->height; iy++)
    for (int ix = 0; ix < cam->width; ix++) {
      const size_t k = (size_t)ix + (size_t)cam->width * (size_t)iy;
      LCG g;
      g.init((uint32_t)((float)ix + (float)cam->width * (float)iy), (uint32_t)cam->frame_id);
      t_seq_rng = &g;
      g_libm_trig = 1;
      PixelOut po = camera_pixel(*s, *env, *rp, *cam, ix, iy, &rays, nullptr);
      t_seq_rng = nullptr;
      g_libm_trig = 0;
      float prev[4] = {color[4 * k], color[4 * k + 1], color[4 * k + 2], color[4 * k + 3]};
      camera_resolve(*cam, po, prev, color + 4 * k, normal + 4 * k, albedo + 4 * k);
    }
  if (rays_out) *rays_out = rays;
  return 0;
}

// ---- textures, cube map, compressed BTF (oracle_btf.h) ----------------------------------------------------------
ORC_API int orc_scene_add_texture(Scene* s, int w, int h, const float* rgba) {
  Texture t;
  t.w = w, t.h = h;
  t.rgba.assign(rgba, rgba + 4 * (size_t)w * h);
  s->textures.push_back(std::move(t));
  return (int)s->textures.size() - 1;
}
ORC_API void orc_scene_set_material_textures(Scene* s, int mat, int albedo, int normal, int metallic, int roughness) {
  if (s->mat_extra.size() < s->mats.size()) s->mat_extra.resize(s->mats.size());
  MatExtra& mx = s->mat_extra[mat];
  mx.tex[0] = albedo, mx.tex[1] = normal, mx.tex[2] = metallic, mx.tex[3] = roughness;
}
ORC_API void orc_scene_set_environment_map(Scene* s, const int* faces6) {
  for (int k = 0; k < 6; k++) s->env_faces[k] = faces6 ? faces6[k] : -1;
}
ORC_API int orc_scene_add_btf(Scene* s, const rtf_btf_desc* d) {
  s->btfs.emplace_back();
  btf_assign(s->btfs.back(), *d);
  return (int)s->btfs.size() - 1;
}
ORC_API void orc_scene_set_material_btf(Scene* s, int mat, int btf) {
  if (s->mat_extra.size() < s->mats.size()) s->mat_extra.resize(s->mats.size());
  s->mat_extra[mat].btf = btf;
}
ORC_API void orc_sample_texture(Scene* s, int tex, uint64_t n, const float* uv, float* rgba) {
  for (uint64_t i = 0; i < n; i++) tex2d(s->textures[tex], uv[2 * i], uv[2 * i + 1], rgba + 4 * i);
}
ORC_API void orc_sample_environment(Scene* s, const rtf_env* env, uint64_t n, const float* pd6, float* rgb) {
  for (uint64_t i = 0; i < n; i++) {
    const float* q = pd6 + 6 * i;
    V3 c = env_light(v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), *env, s);
    rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
  }
}
ORC_API void orc_btf_eval(Scene* s, int btf, uint64_t n, const float* tc, const float* view, const float* illum,
                          const float* normal, const float* tangent, float* rgb, int nthreads) {
  const BTF* b = &s->btfs[btf];
  parallel_for(n, nthreads, [=](int, uint64_t b0, uint64_t e0) {
    for (uint64_t i = b0; i < e0; i++) {
      V3 c = btf_get_value(*b, tc[2 * i], tc[2 * i + 1], v3(view[3 * i], view[3 * i + 1], view[3 * i + 2]),
                           v3(illum[3 * i], illum[3 * i + 1], illum[3 * i + 2]),
                           v3(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]),
                           v3(tangent[3 * i], tangent[3 * i + 1], tangent[3 * i + 2]));
      rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
    }
  });
}
// BTFBase::GetValueDeg layer: q6 = {u, v, illumTheta, illumPhi, viewTheta, viewPhi} (degrees)
ORC_API void orc_btf_value_deg(Scene* s, int btf, uint64_t n, const float* q6, float* rgb) {
  const BTF& b = s->btfs[btf];
  for (uint64_t i = 0; i < n; i++) {
    const float* q = q6 + 6 * i;
    V3 c = btf_value_deg(b, q[0], q[1], q[2], q[3], q[4], q[5]);
    rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
  }
}
ORC_API void orc_btf_beta_angles(Scene* s, int btf, float* out) {
  const BTF& b = s->btfs[btf];
  for (size_t i = 0; i < b.beta_angles.size(); i++) out[i] = b.beta_angles[i];
}
ORC_API void orc_inverse_trig(int n, const float* y, const float* x, float* atan2_out, float* asin_out, float* acos_out) {
  for (int i = 0; i < n; i++) atan2_out[i] = atan2p(y[i], x[i]), asin_out[i] = asinp(y[i]), acos_out[i] = acosp(y[i]);
}
ORC_API void orc_sample_lambert_ggx(int n, const float* albedo3, float metallic, float roughness, const float* normal3,
                                    const float* view3, const float* r3, float* weight3, float* dir3) {
  for (int i = 0; i < n; i++) {
    V3 l;
    V3 w = sample_lambert_ggx(v3(albedo3[0], albedo3[1], albedo3[2]), metallic, roughness, v3(normal3[0], normal3[1], normal3[2]),
                              v3(view3[0], view3[1], view3[2]), r3[3 * i], r3[3 * i + 1], r3[3 * i + 2], l);
    weight3[3 * i] = w.x, weight3[3 * i + 1] = w.y, weight3[3 * i + 2] = w.z;
    dir3[3 * i] = l.x, dir3[3 * i + 1] = l.y, dir3[3 * i + 2] = l.z;
  }
}

// ---- geometry pre-kernels, light probes, flux (oracle_prebake.h) -----------------------------------------
ORC_API void orc_bake_instanced(const void* verts, uint64_t nv, uint64_t stride, const uint32_t* tris, uint64_t nt,
                                const float* matrices, const float* colors, uint64_t n_inst, rtf_vertex* out_v, uint32_t* out_t) {
  std::vector<rtf_vertex> packed(nv);
  for (uint64_t i = 0; i < nv; i++) memcpy(&packed[i], (const char*)verts + i * stride, sizeof(rtf_vertex));
  bake_instanced(packed.data(), nv, tris, nt, matrices, colors, n_inst, out_v, out_t);
}
ORC_API int orc_bake_skinned(const void* sverts, uint64_t nv, uint64_t stride, const float* bones, uint64_t n_bones,
                             rtf_vertex* out_v) {
  return bake_skinned((const unsigned char*)sverts, stride, nv, bones, n_bones, out_v);
}
ORC_API void orc_make_probes(Scene* s, uint64_t tri_begin, uint64_t tri_end, int front, int back, rtf_illum_sampler* out) {
  make_probes(*s, tri_begin, tri_end, front, back, out);
}
ORC_API void orc_flux_probes(const rtf_illum_sampler* probes, uint64_t n, float* out8) { flux_probes(probes, n, out8); }
ORC_API void orc_flux_mesh(Scene* s, const float* energy6, uint64_t tri_begin, uint64_t tri_end, int faces, float* out8) {
  flux_mesh(*s, energy6, tri_begin, tri_end, faces, out8);
}

// ---- BSSRDF spatial probe (oracle_sss.h) -------------------------------------------------------------------
ORC_API void orc_sss_probe(Scene* s, const float* rays, const uint64_t* handles, uint64_t n, uint32_t* out17, int brute) {
  auto handle_of = [&](uint32_t inst) { return s->insts[inst].handle; };
  for (uint64_t i = 0; i < n; i++) {
    const float* r = rays + 8 * i;
    const V3 o = v3(r[0], r[1], r[2]), d = v3(r[4], r[5], r[6]);
    const SSRecords rec = brute ? ss_probe_brute(s->bvh, o, d, r[3], r[7], handles[i], handle_of)
                                : ss_probe(s->bvh, o, d, r[3], r[7], handles[i], handle_of);
    uint32_t* out = out17 + 17 * i;
    out[0] = (uint32_t)rec.n;
    for (int k = 0; k < 4; k++) out[1 + k] = rec.prim[k], out[5 + k] = f2u(rec.t[k]), out[9 + k] = f2u(rec.u[k]), out[13 + k] = f2u(rec.v[k]);
  }
}
ORC_API void orc_rotate_about(const float* v, float u_turns, const float* axis, float* out) {
  const V3 r = rotate_about(v3(v[0], v[1], v[2]), u_turns, v3(axis[0], axis[1], axis[2]));
  out[0] = r.x, out[1] = r.y, out[2] = r.z;
}
