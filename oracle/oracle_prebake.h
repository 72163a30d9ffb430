// oracle_prebake.h — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Scalar restatement of the reference's geometry pre-kernels and of the host loops either side of the
// illumination entry point:
//   bake_instanced  = CopyVerticesInstancedKernel (src/RayTracerFacility/RayTracer.cu:1148-1175) + the index
//                     replication of RayTracedGeometry::BuildGAS (RayTracer.cu:1477-1487, without the N leading
//                     degenerate triples its resize-then-push_back produces — unhittable, see include/rtf/rtf.h)
//   bake_skinned    = CopySkinnedVerticesKernel (RayTracer.cu:1194-1247)
//   make_probes     = TriangleIlluminationEstimator::PrepareLightProbeGroup (TriangleIlluminationEstimator.cpp:94-139)
//   flux_*          = TriangleIlluminationEstimator::SampleLightProbeGroup (:84-92) with
//                     IlluminationSampler::GetArea (RayTracer.hpp:204-210); the sum runs in double over the fixed
//                     tree documented in include/rtf/rtf.h (the reference adds floats serially).
// bake_instanced / bake_skinned are pinned by the reference's own two kernels compiled unmodified (oracle/_ref/libprebake_ref.so,
// tests/golden/prebake_ref.npz, byte-equal).  make_probes / flux_*: parity unpinned by the reference (it has no tests and
// TriangleIlluminationEstimator.cpp does not compile here); pinned by closed forms in tests/test_oracle_prebake.py.
#pragma once
#include <vector>

#include "oracle_shade.h"

namespace orc {

static inline void bake_frame(const float* m, const rtf_vertex& v, V3& P, V3& N, V3& T) {
  P = mat4_mul_point(m, v3(v.position[0], v.position[1], v.position[2]), 1.0f);
  N = normalize(mat4_mul_point(m, v3(v.normal[0], v.normal[1], v.normal[2]), 0.0f));
  T = normalize(mat4_mul_point(m, v3(v.tangent[0], v.tangent[1], v.tangent[2]), 0.0f));
  T = normalize(T - dot(T, N) * N);
}
static inline void put_frame(rtf_vertex& o, V3 P, V3 N, V3 T) {
  o.position[0] = P.x, o.position[1] = P.y, o.position[2] = P.z;
  o.normal[0] = N.x, o.normal[1] = N.y, o.normal[2] = N.z;
  o.tangent[0] = T.x, o.tangent[1] = T.y, o.tangent[2] = T.z;
}

static inline void bake_instanced(const rtf_vertex* verts, uint64_t nv, const uint32_t* tris, uint64_t nt,
                                  const float* matrices, const float* colors, uint64_t n_inst, rtf_vertex* out_v,
                                  uint32_t* out_t) {
  for (uint64_t idx = 0; idx < nv * n_inst; idx++) {
    const uint64_t inst = idx / nv, k = idx % nv;
    V3 P, N, T;
    bake_frame(matrices + 16 * inst, verts[k], P, N, T);
    rtf_vertex o = verts[k];  // texCoord, colour
    put_frame(o, P, N, T);
    const float one[4] = {1.0f, 1.0f, 1.0f, 1.0f};
    const float* c = colors ? colors + 4 * inst : one;
    o.position_padding = c[0], o.normal_padding = c[1], o.tangent_padding = c[2];
    o.tex_coord_padding[0] = c[3], o.tex_coord_padding[1] = (float)inst;
    out_v[idx] = o;
  }
  for (uint64_t inst = 0; inst < n_inst; inst++)
    for (uint64_t i = 0; i < 3 * nt; i++) out_t[inst * 3 * nt + i] = tris[i] + (uint32_t)(inst * nv);
}

static inline int bake_skinned(const unsigned char* sverts, uint64_t stride, uint64_t nv, const float* bones, uint64_t n_bones,
                               rtf_vertex* out_v) {
  for (uint64_t idx = 0; idx < nv; idx++) {
    rtf_skinned_vertex sv;
    memcpy(&sv, sverts + idx * stride, sizeof(sv));
    float m[16];
    for (int j = 0; j < 16; j++) m[j] = 0.0f;
    bool first = true;
    for (int k = 0; k < 8; k++) {
      const int id = k < 4 ? sv.bone_id[k] : sv.bone_id2[k - 4];
      const float w = k < 4 ? sv.weight[k] : sv.weight2[k - 4];
      if (k > 0 && id == -1) continue;
      if (id < 0 || (uint64_t)id >= n_bones) return -1;
      const float* b = bones + 16 * (uint64_t)id;
      for (int j = 0; j < 16; j++) m[j] = first ? b[j] * w : m[j] + b[j] * w;
      first = false;
    }
    V3 P, N, T;
    bake_frame(m, sv.v, P, N, T);
    rtf_vertex o = sv.v;
    put_frame(o, P, N, T);
    out_v[idx] = o;
  }
  return 0;
}

static inline void make_probes(const Scene& sc, uint64_t tri_begin, uint64_t tri_end, int front, int back,
                               rtf_illum_sampler* out) {
  for (uint64_t t = tri_begin; t < tri_end; t++) {
    const Instance& in = sc.insts[sc.tri_inst[t]];
    const Geometry& g = sc.geoms[in.geom];
    const uint64_t lt = t - in.first_prim;
    rtf_illum_sampler pr;
    memset(&pr, 0, sizeof(pr));
    for (int k = 0; k < 3; k++) {
      rtf_vertex v = g.verts[g.tris[3 * lt + k]];
      for (int c = 0; c < 3; c++) v.position[c] = sc.pos9[9 * t + 3 * k + c], v.normal[c] = sc.nrm9[9 * t + 3 * k + c];
      (k == 0 ? pr.a : (k == 1 ? pr.b : pr.c)) = v;
    }
    pr.front_face = front ? 1 : 0, pr.back_face = back ? 1 : 0;
    out[t - tri_begin] = pr;
  }
}

static inline float heron_area(V3 A, V3 B, V3 C) {
  const float a = length(A - B), b = length(B - C), c = length(C - A);
  const float p = (a + b + c) * 0.5f;
  return sqrtf(p * (p - a) * (p - b) * (p - c));
}
struct Flux4 {
  double v[4];
};
// One level of the tree: 1024-term tiles; thread t adds terms t, t+256, t+512, t+768; xor butterfly over the 32 lanes;
// the eight warp sums in order.
template <class Leaf>
static inline std::vector<Flux4> flux_level(Leaf leaf, uint64_t n) {
  uint64_t tiles = (n + 1023) / 1024;
  if (tiles == 0) tiles = 1;
  std::vector<Flux4> out(tiles);
  for (uint64_t b = 0; b < tiles; b++) {
    Flux4 th[256];
    for (int t = 0; t < 256; t++) {
      Flux4 s{{0, 0, 0, 0}};
      for (int j = 0; j < 4; j++) {
        const uint64_t i = b * 1024 + (uint64_t)t + 256u * (uint64_t)j;
        Flux4 v{{0, 0, 0, 0}};
        if (i < n) v = leaf(i);
        for (int c = 0; c < 4; c++) s.v[c] = s.v[c] + v.v[c];
      }
      th[t] = s;
    }
    for (int off = 16; off >= 1; off >>= 1) {
      Flux4 nx[256];
      for (int t = 0; t < 256; t++)
        for (int c = 0; c < 4; c++) nx[t].v[c] = th[t].v[c] + th[t ^ off].v[c];
      memcpy(th, nx, sizeof(th));
    }
    Flux4 r = th[0];
    for (int w = 1; w < 8; w++)
      for (int c = 0; c < 4; c++) r.v[c] = r.v[c] + th[32 * w].v[c];
    out[b] = r;
  }
  return out;
}
template <class Leaf>
static inline void flux_reduce(Leaf leaf, uint64_t n, float* out8) {
  std::vector<Flux4> lvl = flux_level(leaf, n);
  while (lvl.size() > 1) {
    std::vector<Flux4> cur;
    cur.swap(lvl);
    lvl = flux_level([&](uint64_t i) { return cur[i]; }, cur.size());
  }
  const float fx = (float)lvl[0].v[0], fy = (float)lvl[0].v[1], fz = (float)lvl[0].v[2], ar = (float)lvl[0].v[3];
  out8[0] = fx, out8[1] = fy, out8[2] = fz, out8[3] = ar;
  out8[4] = fx / ar, out8[5] = fy / ar, out8[6] = fz / ar, out8[7] = 0.0f;
}
static inline void flux_probes(const rtf_illum_sampler* probes, uint64_t n, float* out8) {
  flux_reduce(
      [&](uint64_t i) {
        const rtf_illum_sampler& p = probes[i];
        const float area = heron_area(v3(p.a.position[0], p.a.position[1], p.a.position[2]),
                                      v3(p.b.position[0], p.b.position[1], p.b.position[2]),
                                      v3(p.c.position[0], p.c.position[1], p.c.position[2]));
        const int faces = (p.front_face ? 1 : 0) + (p.back_face ? 1 : 0);
        return Flux4{{(double)(p.energy[0] * area), (double)(p.energy[1] * area), (double)(p.energy[2] * area),
                      (double)((float)faces * area)}};
      },
      n, out8);
}
static inline void flux_mesh(const Scene& sc, const float* energy6, uint64_t tri_begin, uint64_t tri_end, int faces, float* out8) {
  flux_reduce(
      [&](uint64_t i) {
        const uint64_t t = tri_begin + i;
        const float* p = &sc.pos9[9 * t];
        const float* e = energy6 + 6 * t;
        const float area = heron_area(v3(p[0], p[1], p[2]), v3(p[3], p[4], p[5]), v3(p[6], p[7], p[8]));
        return Flux4{{(double)(e[0] * area), (double)(e[1] * area), (double)(e[2] * area), (double)((float)faces * area)}};
      },
      tri_end - tri_begin, out8);
}

}  // namespace orc
