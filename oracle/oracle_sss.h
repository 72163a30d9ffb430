// oracle_sss.h — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Scalar restatement of the reference's BSSRDF spatial probe: SSAnyHit / BSSRDF (include/RayTracerFacility/RayTracer/
// BSSDF.cuh:18-89), called from ClosestHitFunc (RayFunctions.cuh:97-133).
// The reference's record set depends on the order in which the OptiX driver reports any-hit candidates (an accepted
// candidate shrinks the ray interval), so it is not a function of the scene; the order-independent statement used by
// the product and restated here: records = the up-to-four NEAREST same-handle intersections in [tmin, tmax], ascending
// (t, primitive id), a triangle counted once however many pre-split references it has.
// Parity unpinned by the reference (no tests, OptiX closed); pinned by brute force (ss_probe_brute) and closed forms.
#pragma once
#include "oracle_bvh.h"

namespace orc {

struct SSRecords {
  int n = 0;
  float t[4] = {0, 0, 0, 0}, u[4] = {0, 0, 0, 0}, v[4] = {0, 0, 0, 0};
  uint32_t prim[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
};
static inline void ss_insert(SSRecords& r, float t, float u, float v, uint32_t prim) {
  for (int k = 0; k < r.n; k++)
    if (r.prim[k] == prim) return;
  int pos = r.n;
  for (int k = r.n - 1; k >= 0; k--)
    if (t < r.t[k] || (t == r.t[k] && prim < r.prim[k])) pos = k;
  if (pos >= 4) return;
  for (int k = 3; k > pos; k--)
    if (k <= r.n) r.t[k] = r.t[k - 1], r.u[k] = r.u[k - 1], r.v[k] = r.v[k - 1], r.prim[k] = r.prim[k - 1];
  r.t[pos] = t, r.u[pos] = u, r.v[pos] = v, r.prim[pos] = prim;
  if (r.n < 4) r.n++;
}

// Every candidate in [tmin, tmax] reaches the filter because it never accepts one (the interval does not shrink).
template <class HandleOf>
struct SSCollector {
  SSRecords* rec;
  uint64_t handle;
  HandleOf handle_of;
  bool operator()(const Tri48& tri, const Hit& cand) const {
    if (handle_of(tri.inst) == handle) ss_insert(*rec, cand.t, cand.u, cand.v, tri.prim);
    return false;
  }
};
template <class HandleOf>
static inline SSRecords ss_probe(const BVH& bvh, V3 o, V3 d, float tmin, float tmax, uint64_t handle, HandleOf handle_of) {
  SSRecords rec;
  Ray ray{o, tmin, d, tmax};
  SSCollector<HandleOf> c{&rec, handle, handle_of};
  trace(bvh, ray, nullptr, false, c);
  return rec;
}
template <class HandleOf>
static inline SSRecords ss_probe_brute(const BVH& bvh, V3 o, V3 d, float tmin, float tmax, uint64_t handle, HandleOf handle_of) {
  SSRecords rec;
  Ray ray{o, tmin, d, tmax};
  const RayPrep rp = prep_ray(d);
  for (uint32_t i = 0; i < bvh.nref; i++) {
    Hit cand{tmax, 0.0f, 0.0f, 0xFFFFFFFFu};
    if (intersect_tri(ray, rp, bvh.tris[i], cand) && handle_of(bvh.tris[i].inst) == handle)
      ss_insert(rec, cand.t, cand.u, cand.v, cand.prim);
  }
  return rec;
}

// glm::rotate(v, angle, normal) (glm/gtx/rotate_vector.hpp -> glm/ext/matrix_transform.inl), angle = 2*pi*u.
static inline V3 rotate_about(V3 v, float u_turns, V3 axis_in) {
  float s, c;
  if (t_seq_rng) {  // test mode (oracle_math.h): the reference's own libm calls on 2 pi u (BSSDF.cuh:54)
    const float a = 2.0f * 3.14159265358979323846264338327950288f * u_turns;
    c = cosf(a), s = sinf(a);
  } else {
    sincos2pi(u_turns, &s, &c);
  }
  const V3 axis = normalize(axis_in);
  const V3 temp = (1.0f - c) * axis;
  const V3 c0 = v3(c + temp.x * axis.x, temp.x * axis.y + s * axis.z, temp.x * axis.z - s * axis.y);
  const V3 c1 = v3(temp.y * axis.x - s * axis.z, c + temp.y * axis.y, temp.y * axis.z + s * axis.x);
  const V3 c2 = v3(temp.z * axis.x + s * axis.y, temp.z * axis.y - s * axis.x, c + temp.z * axis.z);
  return v3((c0.x * v.x + c1.x * v.y) + c2.x * v.z, (c0.y * v.x + c1.y * v.y) + c2.y * v.z,
            (c0.z * v.x + c1.z * v.y) + c2.z * v.z);
}

}  // namespace orc
