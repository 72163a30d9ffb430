// oracle_bvh.h — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Scalar CPU restatement of the acceleration structure that replaces the closed-source parts of
// the reference path: optixAccelBuild (RT.cu:1528-1584 GAS, RT.cu:1654-1707 IAS) and optixTrace's
// traversal + built-in triangle intersection (10 call sites).  OptiX itself is
// NVIDIA OptiX 7.4.0/8.0.0 (CMakeLists.txt:27, cmake/FindOptiX.cmake:36,42) and is absent from
// /root/reference, so the published algorithms are restated instead:
//   * Morton codes + LBVH radix tree: Karras, "Maximizing Parallelism in the Construction of BVHs,
//     Octrees, and k-d Trees", HPG 2012.
//   * 8-wide compressed nodes (80 B) and their traversal order: Ylitie, Karras, Laine,
//     "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs", HPG 2017.
//   * ray/triangle: Woop, Benthin, Wald, "Watertight Ray/Triangle Intersection", JCGT 2013.
// PARITY PINNING: the reference holds no tests, fixtures or golden vectors and cannot be
// built here, so these functions are pinned by (1) a brute-force all-triangles intersector below
// (ground truth for hits, independent of any BVH) and (2) closed-form cases in tests/.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <vector>

#include "oracle_math.h"

namespace orc {

struct Node80 {
  float p[3];
  uint8_t e[3];
  uint8_t imask;
  uint32_t child_base;
  uint32_t tri_base;
  uint8_t meta[8];
  uint8_t qlo[3][8];  // [axis][slot]
  uint8_t qhi[3][8];
};
static_assert(sizeof(Node80) == 80, "node must be 80 B");

struct Tri48 {
  float v0[3];
  uint32_t prim;
  float v1[3];
  uint32_t inst;
  float v2[3];
  uint32_t flags;
};
static_assert(sizeof(Tri48) == 48, "triangle must be 48 B");

// Node meta (8 bytes, little-endian words w0 = meta[0..3], w1 = meta[4..7]):
//   w0 bits 0..23 = VALID: bit 3s+j set iff slot s is a leaf slot holding more than j triangles (j = 0..2)
//   w1 bits 0..7  = imask (copy), bits 8..23 = CNT: 2-bit field s = triangles of leaf slot s (0 for inner / empty slots)
// The triangles of a node's leaf slots lie consecutively from tri_base in slot order, so triangle j of slot s is
// tri_base + (sum of the CNT fields below s) + j.  The traversal keeps its pending triangles as the 24-bit mask
// "bit 3s+j" — a fixed position per slot, so the node visit needs no per-child variable shift.
static inline void node_meta_store(Node80& nd, uint32_t valid, uint32_t cnt16) {
  const uint32_t w0 = valid & 0xffffffu, w1 = (uint32_t)nd.imask | ((cnt16 & 0xffffu) << 8);
  for (int k = 0; k < 4; k++) nd.meta[k] = (uint8_t)(w0 >> (8 * k)), nd.meta[4 + k] = (uint8_t)(w1 >> (8 * k));
}
static inline uint32_t node_meta_w0(const Node80& nd) {
  return (uint32_t)nd.meta[0] | ((uint32_t)nd.meta[1] << 8) | ((uint32_t)nd.meta[2] << 16) | ((uint32_t)nd.meta[3] << 24);
}
static inline uint32_t node_meta_w1(const Node80& nd) {
  return (uint32_t)nd.meta[4] | ((uint32_t)nd.meta[5] << 8) | ((uint32_t)nd.meta[6] << 16) | ((uint32_t)nd.meta[7] << 24);
}

struct Box {
  V3 lo, hi;
};

struct BVH {
  uint32_t n = 0;     // triangles
  uint32_t nref = 0;  // triangle references (>= n when long triangles are pre-split, see split_refs)
  int bits = 63;
  Box scene;
  std::vector<uint32_t> ref_first;   // per triangle: index of its first reference (n + 1 entries)
  std::vector<uint32_t> ref_tri;     // per reference: its triangle
  std::vector<Box> ref_box;          // per reference: bounds of the triangle clipped to the reference's slab
  std::vector<uint64_t> morton;      // per reference, unsorted
  std::vector<uint32_t> order;       // sorted position -> reference
  std::vector<int32_t> children;     // 2 per internal binary node; >=0 internal index, <0 = ~leaf position
  std::vector<int32_t> parent;       // per binary node (internals [0,n-1), leaves [n-1,2n-1)); root = -1
  std::vector<uint32_t> range_first; // per internal node
  std::vector<uint32_t> range_last;
  std::vector<Box> boxes;            // per binary node
  std::vector<Node80> nodes;         // wide nodes, BFS order
  std::vector<Tri48> tris;           // per triangle reference, BVH (leaf-slot) order
  int depth = 0;                     // number of wide levels
};

// ---- Morton codes ---------------------------------------------------------------------------
static inline uint32_t expand10(uint32_t v) {  // 10 bits -> every third bit
  v &= 0x3ffu;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}
static inline uint64_t expand21(uint64_t v) {  // 21 bits -> every third bit
  v &= 0x1fffffull;
  v = (v | (v << 32)) & 0x1f00000000ffffull;
  v = (v | (v << 16)) & 0x1f0000ff0000ffull;
  v = (v | (v << 8)) & 0x100f00f00f00f00full;
  v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
  v = (v | (v << 2)) & 0x1249249249249249ull;
  return v;
}
static inline uint32_t quantize_axis(float c, float lo, float inv, float scale, float maxq) {
  float nx = (c - lo) * inv;
  float q = fminf(fmaxf(nx * scale, 0.0f), maxq);
  return (uint32_t)q;
}
static inline uint64_t morton_code(V3 c, V3 lo, V3 inv, int bits) {
  if (bits == 30) {
    uint32_t x = quantize_axis(c.x, lo.x, inv.x, 1024.0f, 1023.0f);
    uint32_t y = quantize_axis(c.y, lo.y, inv.y, 1024.0f, 1023.0f);
    uint32_t z = quantize_axis(c.z, lo.z, inv.z, 1024.0f, 1023.0f);
    return (uint64_t)((expand10(x) << 2) | (expand10(y) << 1) | expand10(z));
  }
  uint64_t x = quantize_axis(c.x, lo.x, inv.x, 2097152.0f, 2097151.0f);
  uint64_t y = quantize_axis(c.y, lo.y, inv.y, 2097152.0f, 2097151.0f);
  uint64_t z = quantize_axis(c.z, lo.z, inv.z, 2097152.0f, 2097151.0f);
  return (expand21(x) << 2) | (expand21(y) << 1) | expand21(z);
}

static inline int clz32(uint32_t x) { return x ? __builtin_clz(x) : 32; }
static inline int clz64(uint64_t x) { return x ? __builtin_clzll(x) : 64; }

// Karras 2012, section 4: common-prefix length with index tie-break.
static inline int karras_delta(const uint64_t* keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  uint64_t a = keys[i], b = keys[j];
  if (a == b) return 64 + clz32((uint32_t)i ^ (uint32_t)j);
  return clz64(a ^ b);
}

static inline float half_area(const Box& b) {
  V3 d = b.hi - b.lo;
  return (d.x * d.y + d.y * d.z) + d.z * d.x;
}
static inline Box box_union(const Box& a, const Box& b) { return Box{vmin(a.lo, b.lo), vmax(a.hi, b.hi)}; }

// Exponent e (biased, as stored) with 2^(e-127) * 255 >= extent.
static inline uint32_t quant_exponent(float extent) {
  float f = extent / 255.0f;
  uint32_t bits = f2u(f);
  uint32_t e = (bits >> 23) & 0xffu;
  if (bits & 0x7fffffu) e += 1;
  if (e < 1) e = 1;
  if (u2f(e << 23) * 255.0f < extent) e += 1;
  if (e > 254) e = 254;
  return e;
}


// ---- triangle references (pre-splitting) -------------------------------------------------------
// Long thin triangles that run diagonally (sorghum leaves are 2 mm x 45 mm slivers) have boxes far larger than
// themselves; a ray then tests dozens of triangles it cannot hit.  Before the Morton sort such a triangle is cut
// into k slabs along the longest axis of its box ("early split clipping", Ernst & Greiner 2007); every slab becomes
// one REFERENCE with the bounds of the clipped triangle.  References of one triangle all point to the same 48 B
// triangle record, so the intersection result does not change (equal t resolves to the lower primitive id).
struct SplitPlan {
  int k, axis;
};
constexpr float kSplitGain = 0.5f;  // k = floor(gain * box half-area / (2 * triangle area)), >= 2 to split
static inline SplitPlan split_plan(V3 a, V3 b, V3 c, int max_split) {
  SplitPlan sp{1, 0};
  const V3 lo = vmin(vmin(a, b), c), hi = vmax(vmax(a, b), c);
  const V3 e = hi - lo;
  sp.axis = (e.x >= e.y && e.x >= e.z) ? 0 : (e.y >= e.z ? 1 : 2);
  if (max_split <= 1) return sp;
  const V3 ab = b - a, ac = c - a;
  const V3 n = v3(ab.y * ac.z - ab.z * ac.y, ab.z * ac.x - ab.x * ac.z, ab.x * ac.y - ab.y * ac.x);
  const float area2 = sqrtf((n.x * n.x + n.y * n.y) + n.z * n.z);  // twice the triangle's area
  const float ha = (e.x * e.y + e.y * e.z) + e.z * e.x;
  if (!(area2 > 0.0f) || !(ha < INFINITY)) return sp;
  const float r = (ha / area2) * kSplitGain;  // a flat box tight around the triangle has half-area == area2
  if (!(r >= 2.0f)) return sp;
  sp.k = r >= (float)max_split ? max_split : (int)r;
  return sp;
}
// Bounds of triangle (a,b,c) clipped to slab j of k along `axis`.  Planes: lo + j*(ext/k); the same float value
// closes slab j-1 and opens slab j, so the slabs tile the box without gaps.  Clipped edge points are rounded, hence
// the other two axes are padded by 2^-21 of the box magnitude and clamped to the triangle's own box.
static inline Box split_slab_box(V3 a, V3 b, V3 c, int axis, int k, int j) {
  const float A[3][3] = {{a.x, a.y, a.z}, {b.x, b.y, b.z}, {c.x, c.y, c.z}};
  float lo[3], hi[3];
  for (int d = 0; d < 3; d++) {
    lo[d] = fminf(fminf(A[0][d], A[1][d]), A[2][d]);
    hi[d] = fmaxf(fmaxf(A[0][d], A[1][d]), A[2][d]);
  }
  if (k <= 1) return Box{v3(lo[0], lo[1], lo[2]), v3(hi[0], hi[1], hi[2])};
  const float w = (hi[axis] - lo[axis]) / (float)k;
  const float x0 = j == 0 ? lo[axis] : fminf(lo[axis] + (float)j * w, hi[axis]);
  const float x1 = j == k - 1 ? hi[axis] : fminf(lo[axis] + (float)(j + 1) * w, hi[axis]);
  float blo[3] = {INFINITY, INFINITY, INFINITY}, bhi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int v = 0; v < 3; v++) {
    if (A[v][axis] >= x0 && A[v][axis] <= x1)
      for (int d = 0; d < 3; d++) blo[d] = fminf(blo[d], A[v][d]), bhi[d] = fmaxf(bhi[d], A[v][d]);
  }
  for (int ed = 0; ed < 3; ed++) {
    const float* P = A[ed];
    const float* Q = A[(ed + 1) % 3];
    if (P[axis] > Q[axis]) {
      const float* T = P;
      P = Q, Q = T;
    }
    for (int s = 0; s < 2; s++) {
      const float xp = s ? x1 : x0;
      if (P[axis] < xp && xp < Q[axis]) {
        const float t = (xp - P[axis]) / (Q[axis] - P[axis]);
        for (int d = 0; d < 3; d++) {
          const float v = fmaf(Q[d] - P[d], t, P[d]);
          blo[d] = fminf(blo[d], v), bhi[d] = fmaxf(bhi[d], v);
        }
      }
    }
  }
  for (int d = 0; d < 3; d++) {
    if (d == axis) {
      blo[d] = x0, bhi[d] = x1;
    } else {
      const float pad = fmaxf(fabsf(lo[d]), fabsf(hi[d])) * 4.76837158203125e-07f;  // 2^-21
      blo[d] = fmaxf(blo[d] - pad, lo[d]);
      bhi[d] = fminf(bhi[d] + pad, hi[d]);
      if (!(blo[d] <= bhi[d])) blo[d] = lo[d], bhi[d] = hi[d];  // empty slab (cannot happen for a valid triangle)
    }
  }
  return Box{v3(blo[0], blo[1], blo[2]), v3(bhi[0], bhi[1], bhi[2])};
}

static inline void build(BVH& bvh, uint32_t ntri, const float* pos9, const uint32_t* tri_inst, int bits,
                         int max_split = 1) {
  bvh.n = ntri;
  bvh.bits = bits;
  const float INF = INFINITY;
  Box sc{v3(INF, INF, INF), v3(-INF, -INF, -INF)};
  bvh.ref_first.assign((size_t)ntri + 1, 0);
  bvh.ref_tri.clear();
  bvh.ref_box.clear();
  for (uint32_t t = 0; t < ntri; t++) {
    const float* p = pos9 + 9 * (size_t)t;
    V3 a = v3(p[0], p[1], p[2]), b = v3(p[3], p[4], p[5]), c = v3(p[6], p[7], p[8]);
    const SplitPlan sp = split_plan(a, b, c, max_split);
    bvh.ref_first[t] = (uint32_t)bvh.ref_tri.size();
    for (int j = 0; j < sp.k; j++) {
      bvh.ref_tri.push_back(t);
      bvh.ref_box.push_back(split_slab_box(a, b, c, sp.axis, sp.k, j));
    }
    sc.lo = vmin(sc.lo, vmin(vmin(a, b), c));
    sc.hi = vmax(sc.hi, vmax(vmax(a, b), c));
  }
  bvh.ref_first[ntri] = (uint32_t)bvh.ref_tri.size();
  const uint32_t n = bvh.nref = (uint32_t)bvh.ref_tri.size();
  const std::vector<Box>& tbox = bvh.ref_box;
  bvh.scene = sc;
  V3 ext = sc.hi - sc.lo;
  V3 inv = v3(ext.x > 0 ? 1.0f / ext.x : 0.0f, ext.y > 0 ? 1.0f / ext.y : 0.0f, ext.z > 0 ? 1.0f / ext.z : 0.0f);
  bvh.morton.resize(n);
  for (uint32_t t = 0; t < n; t++) {
    V3 c = (tbox[t].lo + tbox[t].hi) * 0.5f;
    bvh.morton[t] = morton_code(c, sc.lo, inv, bits);
  }
  bvh.order.resize(n);
  for (uint32_t t = 0; t < n; t++) bvh.order[t] = t;
  std::stable_sort(bvh.order.begin(), bvh.order.end(),
                   [&](uint32_t a, uint32_t b) { return bvh.morton[a] < bvh.morton[b]; });
  std::vector<uint64_t> keys(n);
  for (uint32_t i = 0; i < n; i++) keys[i] = bvh.morton[bvh.order[i]];

  // ---- Karras radix tree ----
  const int N = (int)n;
  const int ni = N > 0 ? N - 1 : 0;
  bvh.children.assign(2 * (size_t)ni, 0);
  bvh.parent.assign((size_t)(N > 0 ? 2 * N - 1 : 0), -1);
  bvh.range_first.assign(ni, 0);
  bvh.range_last.assign(ni, 0);
  for (int i = 0; i < ni; i++) {
    int d = (karras_delta(keys.data(), N, i, i + 1) - karras_delta(keys.data(), N, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = karras_delta(keys.data(), N, i, i - d);
    int lmax = 2;
    while (karras_delta(keys.data(), N, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
      if (karras_delta(keys.data(), N, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = karras_delta(keys.data(), N, i, j);
    int s = 0;
    int t = l;
    do {
      t = (t + 1) >> 1;
      if (karras_delta(keys.data(), N, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + std::min(d, 0);
    int lo = std::min(i, j), hi = std::max(i, j);
    int left = (lo == gamma) ? ~gamma : gamma;
    int right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    bvh.children[2 * i] = left;
    bvh.children[2 * i + 1] = right;
    bvh.range_first[i] = (uint32_t)lo;
    bvh.range_last[i] = (uint32_t)hi;
    bvh.parent[left >= 0 ? left : (ni + ~left)] = i;
    bvh.parent[right >= 0 ? right : (ni + ~right)] = i;
  }
  // ---- refit (bottom-up; min/max are exact so any order gives the same boxes) ----
  bvh.boxes.resize(N > 0 ? 2 * (size_t)N - 1 : 0);
  for (int k = 0; k < N; k++) bvh.boxes[ni + k] = tbox[bvh.order[k]];
  {
    std::vector<int> pending(ni, 0);
    for (int k = 0; k < N; k++) {
      int node = bvh.parent[ni + k];
      while (node >= 0) {
        if (++pending[node] < 2) break;
        int l = bvh.children[2 * node], r = bvh.children[2 * node + 1];
        const Box& bl = bvh.boxes[l >= 0 ? l : ni + ~l];
        const Box& br = bvh.boxes[r >= 0 ? r : ni + ~r];
        bvh.boxes[node] = box_union(bl, br);
        node = bvh.parent[node];
      }
    }
  }
  // ---- greedy surface-area collapse to 8-wide, breadth first ----
  bvh.nodes.clear();
  bvh.tris.assign(n, Tri48{});
  bvh.depth = 0;
  if (N == 0) return;
  auto bin_box = [&](int c) -> const Box& { return bvh.boxes[c >= 0 ? c : ni + ~c]; };
  auto bin_count = [&](int c) -> uint32_t { return c >= 0 ? bvh.range_last[c] - bvh.range_first[c] + 1 : 1u; };
  auto bin_first = [&](int c) -> uint32_t { return c >= 0 ? bvh.range_first[c] : (uint32_t)~c; };
  std::vector<int> level;
  level.push_back(N == 1 ? ~0 : 0);
  uint32_t tri_cursor = 0;
  while (!level.empty()) {
    bvh.depth++;
    uint32_t level_start = (uint32_t)bvh.nodes.size();
    uint32_t next_start = level_start + (uint32_t)level.size();
    std::vector<int> next;
    for (size_t li = 0; li < level.size(); li++) {
      int R = level[li];
      int ch[8];
      int nc;
      if (R < 0) {  // single-triangle scene
        ch[0] = R;
        nc = 1;
      } else {
        ch[0] = bvh.children[2 * R];
        ch[1] = bvh.children[2 * R + 1];
        nc = 2;
      }
      while (nc < 8) {
        int best = -1;
        float best_area = -INF;
        for (int i = 0; i < nc; i++) {
          if (ch[i] < 0) continue;
          float a = half_area(bin_box(ch[i]));
          if (a > best_area) best_area = a, best = i;
        }
        if (best < 0) break;
        int c = ch[best];
        ch[best] = bvh.children[2 * c];
        ch[nc++] = bvh.children[2 * c + 1];
      }
      const Box& nb = bin_box(R);
      V3 center = (nb.lo + nb.hi) * 0.5f;
      float cost[8][8];
      for (int i = 0; i < nc; i++) {
        const Box& cb = bin_box(ch[i]);
        V3 d = (cb.lo + cb.hi) * 0.5f - center;
        for (int s = 0; s < 8; s++)
          cost[i][s] = (((s & 4) ? d.x : -d.x) + ((s & 2) ? d.y : -d.y)) + ((s & 1) ? d.z : -d.z);
      }
      int slot_of[8], child_in_slot[8];
      for (int i = 0; i < 8; i++) slot_of[i] = -1, child_in_slot[i] = -1;
      for (int it = 0; it < nc; it++) {
        float best = -INF;
        int bi = -1, bs = -1;
        for (int i = 0; i < nc; i++) {
          if (slot_of[i] >= 0) continue;
          for (int s = 0; s < 8; s++) {
            if (child_in_slot[s] >= 0) continue;
            if (bi < 0 || cost[i][s] > best) best = cost[i][s], bi = i, bs = s;
          }
        }
        slot_of[bi] = bs;
        child_in_slot[bs] = bi;
      }
      Node80 node;
      memset(&node, 0, sizeof(node));
      node.p[0] = nb.lo.x, node.p[1] = nb.lo.y, node.p[2] = nb.lo.z;
      V3 ext3 = nb.hi - nb.lo;
      uint32_t e[3] = {quant_exponent(ext3.x), quant_exponent(ext3.y), quant_exponent(ext3.z)};
      for (int a = 0; a < 3; a++) node.e[a] = (uint8_t)e[a];
      node.child_base = next_start + (uint32_t)next.size();
      node.tri_base = tri_cursor;
      uint32_t tri_off = 0, valid = 0, cnt16 = 0;
      for (int s = 0; s < 8; s++) {
        int i = child_in_slot[s];
        if (i < 0) continue;
        int c = ch[i];
        const Box& cb = bin_box(c);
        const float lo3[3] = {cb.lo.x, cb.lo.y, cb.lo.z}, hi3[3] = {cb.hi.x, cb.hi.y, cb.hi.z};
        for (int a = 0; a < 3; a++) {
          float inv_scale = u2f((254u - e[a]) << 23);  // 2^-(e-127), exact
          float scale = u2f(e[a] << 23);
          float ql = floorf((lo3[a] - node.p[a]) * inv_scale);
          float qh = ceilf((hi3[a] - node.p[a]) * inv_scale);
          ql = fminf(fmaxf(ql, 0.0f), 255.0f);
          qh = fminf(fmaxf(qh, 0.0f), 255.0f);
          while (ql > 0.0f && node.p[a] + ql * scale > lo3[a]) ql -= 1.0f;
          while (qh < 255.0f && node.p[a] + qh * scale < hi3[a]) qh += 1.0f;
          node.qlo[a][s] = (uint8_t)ql;
          node.qhi[a][s] = (uint8_t)qh;
        }
        uint32_t cnt = bin_count(c);
        if (c >= 0 && cnt > 3) {
          node.imask |= (uint8_t)(1u << s);
          next.push_back(c);
        } else {
          // leaf slot s: `cnt` (1..3) triangles, stored consecutively in slot order; bit 3s+j of the valid mask says
          // "slot s holds more than j triangles", the 2-bit field s of the count word is cnt (node_meta below)
          valid |= (cnt == 1 ? 1u : (cnt == 2 ? 3u : 7u)) << (3 * s);
          cnt16 |= cnt << (2 * s);
          uint32_t first = bin_first(c);
          for (uint32_t j = 0; j < cnt; j++) {
            uint32_t prim = bvh.ref_tri[bvh.order[first + j]];
            Tri48& T = bvh.tris[tri_cursor + tri_off + j];
            const float* p = pos9 + 9 * (size_t)prim;
            T.v0[0] = p[0], T.v0[1] = p[1], T.v0[2] = p[2];
            T.v1[0] = p[3], T.v1[1] = p[4], T.v1[2] = p[5];
            T.v2[0] = p[6], T.v2[1] = p[7], T.v2[2] = p[8];
            T.prim = prim;
            T.inst = tri_inst ? tri_inst[prim] : 0;
            T.flags = 0;
          }
          tri_off += cnt;
        }
      }
      tri_cursor += tri_off;
      node_meta_store(node, valid, cnt16);
      bvh.nodes.push_back(node);
    }
    level.swap(next);
  }
}

// ---- rays, hits -------------------------------------------------------------------------------
struct Ray {
  V3 o;
  float tmin;
  V3 d;
  float tmax;
};
struct Hit {
  float t, u, v;
  uint32_t prim;  // 0xFFFFFFFF = miss
};
struct RayPrep {  // per-ray shear matrix of the watertight test (axis permutation folded in)
  float mx[3], my[3], mz[3];
};
static inline RayPrep prep_ray(V3 d) {
  float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
  int kz = (ax >= ay && ax >= az) ? 0 : (ay >= az ? 1 : 2);
  int kx = (kz + 1) % 3, ky = (kz + 2) % 3;  // no winding swap: nothing is culled, the sign test is symmetric
  const float dd[3] = {d.x, d.y, d.z};
  RayPrep r;
  float Sz = 1.0f / dd[kz];
  float Sx = dd[kx] * Sz;
  float Sy = dd[ky] * Sz;
  for (int a = 0; a < 3; a++) r.mx[a] = r.my[a] = r.mz[a] = 0.0f;
  r.mx[kx] = 1.0f, r.mx[kz] = -Sx;
  r.my[ky] = 1.0f, r.my[kz] = -Sy;
  r.mz[kz] = Sz;
  return r;
}
static inline float shear_row(const float m[3], const float a[3]) { return fmaf(m[2], a[2], fmaf(m[1], a[1], m[0] * a[0])); }
// Woop/Benthin/Wald 2013, no back-face culling.  Updates `best` if this triangle is the new
// closest hit; equal t resolves to the lower primitive id.
static inline bool intersect_tri(const Ray& ray, const RayPrep& rp, const Tri48& tri, Hit& best) {
  const float o[3] = {ray.o.x, ray.o.y, ray.o.z};
  float A[3], B[3], C[3];
  for (int a = 0; a < 3; a++) A[a] = tri.v0[a] - o[a], B[a] = tri.v1[a] - o[a], C[a] = tri.v2[a] - o[a];
  const float Ax = shear_row(rp.mx, A), Ay = shear_row(rp.my, A);
  const float Bx = shear_row(rp.mx, B), By = shear_row(rp.my, B);
  const float Cx = shear_row(rp.mx, C), Cy = shear_row(rp.my, C);
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
  const float Az = shear_row(rp.mz, A), Bz = shear_row(rp.mz, B), Cz = shear_row(rp.mz, C);
  const float T = fmaf(U, Az, fmaf(V, Bz, W * Cz));
  const float rcp = 1.0f / det;
  const float t = T * rcp;
  if (!(t >= ray.tmin)) return false;
  if (t > best.t) return false;
  if (t == best.t && tri.prim >= best.prim) return false;
  best.t = t;
  best.u = V * rcp;
  best.v = W * rcp;
  best.prim = tri.prim;
  return true;
}

static inline Hit trace_brute(const BVH& bvh, const Ray& ray) {
  Hit best{ray.tmax, 0.0f, 0.0f, 0xFFFFFFFFu};
  RayPrep rp = prep_ray(ray.d);
  for (uint32_t i = 0; i < bvh.nref; i++) intersect_tri(ray, rp, bvh.tris[i], best);
  return best;
}

struct TraceStats {
  uint32_t nodes = 0, tris = 0;
};

// Optional per-candidate filter (the any-hit program, RF.cuh:5-23): return false to ignore the hit.
struct NoFilter {
  bool operator()(const Tri48&, const Hit&) const { return true; }
};

// Sign bit of an IEEE result; a NaN counts as non-negative (the GPU's arithmetic returns the canonical NaN 0x7fffffff).
static inline bool sign_neg(float d) { return d != d ? false : std::signbit(d); }
static inline float safe_dir(float d) {
  const float ooeps = u2f(0x17800000u);  // 2^-80
  return fabsf(d) < ooeps ? copysignf(ooeps, d) : d;
}

// Ylitie et al. 2017, Algorithm 1 without the warp-level postponing heuristics: the per-ray visit
// order is a pure function of (BVH, ray), which is what the GPU kernel reproduces.
template <class Filter = NoFilter>
static inline Hit trace(const BVH& bvh, const Ray& ray, TraceStats* stats = nullptr, bool any_hit = false,
                        Filter filter = Filter()) {
  Hit best{ray.tmax, 0.0f, 0.0f, 0xFFFFFFFFu};
  if (bvh.nodes.empty()) return best;
  const RayPrep rp = prep_ray(ray.d);
  const float dx = safe_dir(ray.d.x), dy = safe_dir(ray.d.y), dz = safe_dir(ray.d.z);
  const float idir[3] = {1.0f / dx, 1.0f / dy, 1.0f / dz};
  const bool neg[3] = {std::signbit(dx), std::signbit(dy), std::signbit(dz)};
  const float org[3] = {ray.o.x, ray.o.y, ray.o.z};
  const uint32_t oct_inv = (neg[0] ? 0u : 4u) | (neg[1] ? 0u : 2u) | (neg[2] ? 0u : 1u);
  const float SLACK = 1.00000095367431640625f;    // 1 + 2^-20
  const float FAR_SLACK = 2.384185791015625e-07f;  // 2^-22
  struct Group {
    uint32_t base, bits;
  };
  std::vector<Group> stack;
  stack.reserve(64);
  Group cur{0u, 0x80000000u};
  for (;;) {
    // invariant: cur holds at least one un-visited internal child hit (bits 24..31)
    Group tg{0u, 0u};
    {
      const uint32_t hits = cur.bits;
      const uint32_t off = 31u - (uint32_t)__builtin_clz(hits);
      cur.bits &= ~(1u << off);
      if (cur.bits & 0xff000000u) stack.push_back(cur);
      const uint32_t slot = (off - 24u) ^ oct_inv;
      const uint32_t rel = (uint32_t)__builtin_popcount(hits & ~(0xffffffffu << slot));
      const Node80& nd = bvh.nodes[cur.base + rel];
      if (stats) stats->nodes++;
      // Plane distance t = q*s + o with the byte q turned into the float 1 + q*2^-15 (one byte-permute on the GPU,
      // no int->float conversion): t = (1 + q*2^-15) * (s*2^15) + (o - s*2^15).  Near planes are pulled in and far planes
      // pushed out by 2^-22 of the operands that cancel, which bounds the rounding error (see rtf_trace.cuh).
      float sc[3], oo[3], oof[3];
      for (int a = 0; a < 3; a++) {
        sc[a] = (u2f((uint32_t)nd.e[a] << 23) * idir[a]) * 32768.0f;
        const float o = (nd.p[a] - org[a]) * idir[a];
        const float m = fabsf(o) + fabsf(sc[a]);
        oo[a] = fmaf(m, -FAR_SLACK, o) - sc[a];
        oof[a] = fmaf(m, FAR_SLACK, o) - sc[a];
      }
      const float best_s = best.t * SLACK;
      // A child is entered iff none of (tfar - tnear), (best_s - tnear), (tfar - tmin) is negative — the slab test
      // max(tnear, tmin) <= min(tfar, best_s) written as three sign bits (the GPU gathers them with funnel shifts).
      uint32_t hit8 = 0;
      for (int s = 0; s < 8; s++) {
        float tn[3], tf[3];
        for (int a = 0; a < 3; a++) {
          const float qn = u2f(0x3F800000u | ((uint32_t)(neg[a] ? nd.qhi[a][s] : nd.qlo[a][s]) << 8));
          const float qf = u2f(0x3F800000u | ((uint32_t)(neg[a] ? nd.qlo[a][s] : nd.qhi[a][s]) << 8));
          tn[a] = fmaf(qn, sc[a], oo[a]);
          tf[a] = fmaf(qf, sc[a], oof[a]);
        }
        const float tnear = fmaxf(fmaxf(tn[0], tn[1]), tn[2]);
        const float tfar = fminf(fminf(tf[0], tf[1]), tf[2]);
        if (!(sign_neg(tfar - tnear) || sign_neg(best_s - tnear) || sign_neg(tfar - ray.tmin))) hit8 |= 1u << s;
      }
      // inner children: hit bit of slot s moves to priority (s ^ oct_inv); leaf slots: bits 3s..3s+2 of the valid mask
      const uint32_t w0 = node_meta_w0(nd), w1 = node_meta_w1(nd);
      const uint32_t inner = hit8 & nd.imask;
      uint32_t perm = 0;
      for (int s = 0; s < 8; s++)
        if (inner & (1u << s)) perm |= 1u << (s ^ oct_inv);
      uint32_t spread = 0;
      for (int s = 0; s < 8; s++)
        if (hit8 & (1u << s)) spread |= 7u << (3 * s);
      cur.base = nd.child_base;
      cur.bits = (perm << 24) | w1;
      tg.base = nd.tri_base;
      tg.bits = spread & w0;
    }
    while (tg.bits) {
      const uint32_t pos = 31u - (uint32_t)__builtin_clz(tg.bits);
      tg.bits &= ~(1u << pos);
      const uint32_t ts = pos / 3u, tj = pos - 3u * ts;
      uint32_t before = 0;  // triangles of the leaf slots below ts (2-bit count fields of cur.bits, bits 8..23)
      for (uint32_t s = 0; s < ts; s++) before += (cur.bits >> (8 + 2 * s)) & 3u;
      const Tri48& tri = bvh.tris[tg.base + before + tj];
      if (stats) stats->tris++;
      Hit cand = best;
      if (intersect_tri(ray, rp, tri, cand) && filter(tri, cand)) {
        best = cand;
        if (any_hit) return best;
      }
    }
    if ((cur.bits & 0xff000000u) == 0) {
      if (stack.empty()) break;
      cur = stack.back();
      stack.pop_back();
    }
  }
  return best;
}

}  // namespace orc
