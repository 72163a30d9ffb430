// oracle_btf.h — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Scalar CPU restatement of the reference's compressed-BTF (MLVQ) evaluator, its linear-memory stand-ins for GL
// textures, and the cube-map lookup.  Paths relative to /root/reference/include/RayTracerFacility/:
//   MLVQ/BTFBase.cuh:49-81            BTFBase::GetValueDeg
//   MLVQ/PDF6D.cuh:32-95              PDF6D::GetValDeg2 (texel lookup, angle re-parameterisation)
//   MLVQ/SharedCoordinates.cuh:49-165 interpolation set-up, ConvertThetaPhiToBetaAlpha
//   MLVQ/PDF4D.cuh:33-63, PDF3D.cuh:31-46, PDF2D.cuh:46-141, PDF1D.cuh:24-69 (Hermite), IndexAB.cuh:26-46,
//   MLVQ/VectorColor.cuh:22-28, MLVQ/CIELab.cuh:8-186 (YCbCr / LogLuv -> RGB)
//   RayTracer/RayDataDefinations.hpp:286-354  SurfaceCompressedBTF::ComputeAngles / GetValue
//   RayTracer/RayDataDefinations.hpp:240-284  SurfaceMaterial texture look-ups
//   RayTracer/RayTracerUtilities.cuh:54-80    SampleCubeMap
// PARITY PINNING: btf_value_deg and everything below it are pinned by the reference's own MLVQ headers compiled on the
// host (ref_btf_driver.cpp, tests/golden/btf_ref.npz); the cube-map addressing by RayTracerUtilities.cuh compiled on the
// host (ref_env_driver.cpp, tests/golden/env_ref.npz).  Texture filtering has no reference source (it is hardware):
// pinned by closed forms in tests/ (texel centres, wrap periodicity).
// Departures: sin/cos/asin/acos/atan2 use the portable polynomials of oracle_math.h (reproducible on host and device);
// texture filtering uses full float weights (tex2D keeps 8 fractional bits).
#pragma once
#include <cstdint>
#include <vector>

#include "../include/rtf/rtf.h"
#include "oracle_math.h"

namespace orc {

// ---- images (tex2D<float4>, wrap + linear + normalised coordinates, RT.cu:2246-2252) ----------------------------
struct Texture {
  int w = 0, h = 0;
  std::vector<float> rgba;  // 4 per texel, row-major
};
static inline int wrapi(int i, int n) {
  int m = i % n;
  return m < 0 ? m + n : m;
}
static inline void tex2d(const Texture& t, float u, float v, float out[4]) {
  const float x = (u - floorf(u)) * (float)t.w - 0.5f, y = (v - floorf(v)) * (float)t.h - 0.5f;
  const float fx = floorf(x), fy = floorf(y);
  const float a = x - fx, b = y - fy;
  const int x0 = wrapi((int)fx, t.w), x1 = wrapi((int)fx + 1, t.w), y0 = wrapi((int)fy, t.h), y1 = wrapi((int)fy + 1, t.h);
  const float* t00 = &t.rgba[4 * ((size_t)y0 * t.w + x0)];
  const float* t10 = &t.rgba[4 * ((size_t)y0 * t.w + x1)];
  const float* t01 = &t.rgba[4 * ((size_t)y1 * t.w + x0)];
  const float* t11 = &t.rgba[4 * ((size_t)y1 * t.w + x1)];
  const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
  for (int k = 0; k < 4; k++) out[k] = ((w00 * t00[k] + w10 * t10[k]) + w01 * t01[k]) + w11 * t11[k];
}
// SampleCubeMap, RTU.cuh:54-80.  faces: +X -X +Y -Y +Z -Z.
static inline int cube_map_lookup(V3 d, float& u, float& v) {
  const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
  float ma;
  int face;
  if (az >= ax && az >= ay) {
    face = d.z < 0.0f ? 5 : 4;
    ma = 0.5f / az;
    u = d.z < 0.0f ? -d.x : d.x, v = -d.y;
  } else if (ay >= ax) {
    face = d.y < 0.0f ? 3 : 2;
    ma = 0.5f / ay;
    u = d.x, v = d.y > 0.0f ? d.z : -d.z;
  } else {
    face = d.x < 0.0f ? 1 : 0;
    ma = 0.5f / ax;
    u = d.x < 0.0f ? d.z : -d.z, v = -d.y;
  }
  u = u * ma + 0.5f, v = v * ma + 0.5f;
  return face;
}
static inline void sample_cube_map(const Texture* const faces[6], V3 d, float out[4]) {
  float u, v;
  const int face = cube_map_lookup(d, u, v);
  tex2d(*faces[face], u, v, out);
}

// ---- compressed BTF ------------------------------------------------------------------------------------------------
struct BTF {
  rtf_btf_desc d;  // scalars only: its pointers are the caller's and may dangle; the tables live in the vectors below
  std::vector<int32_t> pdf6_slices, index_ab, pdf2d_colors, pdf2d_slices, pdf2d_indices, pdf3d_slices, pdf4d_slices;
  std::vector<float> pdf6_scales, pdf1d_basis, colors, pdf2d_scales, pdf3d_scales, pdf4d_scales, beta_angles;
};
// CompressedBTF.cpp:162-186: the beta angles (degrees) at which the 1-D slices are tabulated.
static inline void btf_beta_angles(BTF& b) {
  const int n = b.d.num_beta;
  b.beta_angles.resize(n);
  if (b.d.use_cos_beta) {
    for (int i = 0; i < n; i++) {
      float s = -1.0f + 2.0f * i / (n - 1);
      if (s > 1.0f) s = 1.0f;
      b.beta_angles[i] = degreesf(asinp(s));
    }
    b.beta_angles[0] = -90.f, b.beta_angles[(n - 1) / 2] = 0.f, b.beta_angles[n - 1] = 90.f;
  } else {
    const float step = 180.f / (n - 1);
    for (int i = 0; i < n; i++) b.beta_angles[i] = i * step - 90.f;
    b.beta_angles[(n - 1) / 2] = 0.f, b.beta_angles[n - 1] = 90.0f;
  }
}
static inline void btf_assign(BTF& b, const rtf_btf_desc& d) {
  b.d = d;
  b.pdf6_slices.assign(d.pdf6_slices, d.pdf6_slices + (size_t)d.pdf6_rows * d.pdf6_cols);
  b.pdf6_scales.assign(d.pdf6_scales, d.pdf6_scales + (size_t)d.pdf6_rows * d.pdf6_cols);
  b.pdf1d_basis.assign(d.pdf1d_basis, d.pdf1d_basis + (size_t)d.n_pdf1d * d.num_beta);
  b.colors.assign(d.colors, d.colors + (size_t)d.n_colors * d.n_channels);
  b.index_ab.assign(d.index_ab, d.index_ab + (size_t)d.n_index_slices * d.num_beta);
  b.pdf2d_colors.assign(d.pdf2d_colors, d.pdf2d_colors + (size_t)d.n_pdf2d_color * d.num_alpha);
  b.pdf2d_slices.assign(d.pdf2d_slices, d.pdf2d_slices + (size_t)d.n_pdf2d_lum * d.num_alpha);
  b.pdf2d_scales.assign(d.pdf2d_scales, d.pdf2d_scales + (size_t)d.n_pdf2d_lum * d.num_alpha);
  b.pdf2d_indices.assign(d.pdf2d_indices, d.pdf2d_indices + (size_t)d.n_pdf2d * d.pdf2d_slice_len);
  b.pdf3d_scales.assign(d.pdf3d_scales, d.pdf3d_scales + (size_t)d.n_pdf3d * d.num_theta);
  b.pdf3d_slices.assign(d.pdf3d_slices, d.pdf3d_slices + (size_t)d.n_pdf3d * d.num_theta);
  b.pdf4d_scales.assign(d.pdf4d_scales, d.pdf4d_scales + (size_t)d.n_pdf4d * d.num_phi);
  b.pdf4d_slices.assign(d.pdf4d_slices, d.pdf4d_slices + (size_t)d.n_pdf4d * d.num_phi);
  btf_beta_angles(b);
}

// SharedCoordinates (SharedCoordinates.cuh:9-123): interpolation cell + weight per angle.
struct BtfCoords {
  int i_beta, i_alpha, i_theta, i_phi;
  float w_beta, w_alpha, w_theta, w_phi;
};
static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
static inline float clampf_(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

// PDF1D::GetVal, PDF1D.cuh:24-69 (Fergusson / cubic Hermite)
static inline float btf_pdf1(const BTF& b, int slice, const BtfCoords& tc) {
  const int nb = b.d.num_beta, i = tc.i_beta;
  const float* row = &b.pdf1d_basis[(size_t)slice * nb];
  const float w = tc.w_beta, p0 = row[i], p1 = row[i + 1];
  const float m0h = i == 0 ? p1 - p0 : 0.5f * (p1 - row[i - 1]);
  // as written (PDF1D.cuh:51-55): the "standard way" tangent re-reads index i+1, i.e. 0.5*(p1 - p0)
  const float m1h = i == nb - 2 ? p1 - p0 : 0.5f * (row[i + 1] - p0);
  const float t2 = w * w, t3 = t2 * w;
  const float h01 = -2.0f * t3 + 3.0f * t2, h00 = 1.0f - h01, h11 = t3 - t2, h10 = h11 - t2 + w;
  return ((h00 * p0 + h10 * m0h) + h01 * p1) + h11 * m1h;
}
// IndexAB::GetVal, IndexAB.cuh:36-46 + VectorColor::Get, VectorColor.cuh:22-28
static inline void btf_iab(const BTF& b, int slice, const BtfCoords& tc, float ab[2]) {
  const int* row = &b.index_ab[(size_t)slice * b.d.num_beta];
  const int nc = b.d.n_channels;
  const int c0 = row[tc.i_beta], c1 = row[tc.i_beta + 1];
  for (int k = 0; k < 2; k++)
    ab[k] = (1.f - tc.w_beta) * b.colors[(size_t)c0 * nc + k] + tc.w_beta * b.colors[(size_t)c1 * nc + k];
}
// CIELab.cuh:8-186
static inline V3 btf_clamp0(V3 c) { return v3(c.x < 0.f ? 0.f : c.x, c.y < 0.f ? 0.f : c.y, c.z < 0.f ? 0.f : c.z); }
static inline V3 btf_ycbcr(V3 y, bool normalized) {
  const float s = normalized ? 256.0f : 1.0f;
  V3 rgb;
  rgb.x = y.x * 1.1643828f + y.z * 1.5960273f - 222.921f / s;
  rgb.y = y.x * 1.1643828f - y.y * 0.39176172f - y.z * 0.81296875f + 135.576f / s;
  rgb.z = y.x * 1.1643828f + y.y * 2.0172344f - 276.836f / s;
  return btf_clamp0(rgb);
}
static inline float btf_srgb(float c) { return c > 0.0031308f ? 1.055f * powf(c, 1.f / 2.4f) - 0.055f : 12.92f * c; }
static inline V3 btf_logluv(V3 luv, bool normalized) {
  float Y;
  if (luv.x < 98.381f)
    Y = 0.056968f * luv.x;
  else if (luv.x < 1204.7f)
    Y = 7.3014e-30f * powf(luv.x + 884.17f, 9.9872f);
  else
    Y = 32.994f * expf(0.0047811f * luv.x);
  float X = 0.f, Z = 0.f;
  if (luv.z > 0.f) {
    X = 9.0f / 4.0f * luv.y / luv.z * Y;
    Z = Y * (3.0f * 410.f / luv.z - 5.0f) - X / 3.0f;
  }
  if (normalized) X /= 256.0f, Y /= 256.0f, Z /= 256.0f;
  V3 rgb;
  rgb.x = X * 3.2406f + Y * -1.5372f + Z * -0.4986f;
  rgb.y = X * -0.9689f + Y * 1.8758f + Z * 0.0415f;
  rgb.z = X * 0.0557f + Y * -0.2040f + Z * 1.0570f;
  rgb = btf_clamp0(rgb);
  return v3(btf_srgb(rgb.x), btf_srgb(rgb.y), btf_srgb(rgb.z));
}
static inline V3 btf_user_cm_to_rgb(const BTF& b, V3 cm) {  // UserCmToRgb, CIELab.cuh:168-185
  if (b.d.hdr) return btf_logluv(cm, b.d.code_btf_flag != 0);
  return btf_ycbcr(cm, b.d.code_btf_flag != 0);
}
// PDF2DSeperate::GetVal, PDF2D.cuh:121-141 (+ PDF2DColor::GetVal :46-60, PDF2DLuminance::GetVal :83-103)
static inline V3 btf_pdf2(const BTF& b, int idx, const BtfCoords& tc) {
  const int na = b.d.num_alpha, i = tc.i_alpha;
  const float w = tc.w_alpha;
  const int col = b.pdf2d_indices[(size_t)idx * b.d.pdf2d_slice_len + 1];
  const int lum = b.pdf2d_indices[(size_t)idx * b.d.pdf2d_slice_len + 0];
  V3 cm;
  float ab1[2], ab2[2];
  btf_iab(b, b.pdf2d_colors[(size_t)col * na + i], tc, ab1);
  btf_iab(b, b.pdf2d_colors[(size_t)col * na + i + 1], tc, ab2);
  cm.y = ab1[0] * (1.0f - w) + ab2[0] * w;
  cm.z = ab1[1] * (1.0f - w) + ab2[1] * w;
  const float l1 = b.pdf2d_scales[(size_t)lum * na + i] * btf_pdf1(b, b.pdf2d_slices[(size_t)lum * na + i], tc);
  const float l2 = b.pdf2d_scales[(size_t)lum * na + i + 1] * btf_pdf1(b, b.pdf2d_slices[(size_t)lum * na + i + 1], tc);
  cm.x = (1.f - w) * l1 + w * l2;
  return btf_user_cm_to_rgb(b, cm);
}
// PDF3D::GetVal, PDF3D.cuh:31-46
static inline V3 btf_pdf3(const BTF& b, int idx, const BtfCoords& tc) {
  const int nt = b.d.num_theta, i = tc.i_theta;
  const float w = tc.w_theta;
  const V3 o1 = btf_pdf2(b, b.pdf3d_slices[(size_t)idx * nt + i], tc);
  const V3 o2 = btf_pdf2(b, b.pdf3d_slices[(size_t)idx * nt + i + 1], tc);
  const float s1 = b.pdf3d_scales[(size_t)idx * nt + i] * (1.0f - w);
  const float s2 = b.pdf3d_scales[(size_t)idx * nt + i + 1] * w;
  return o1 * s1 + o2 * s2;
}
// PDF4D::GetVal, PDF4D.cuh:33-63 (phi wraps around)
static inline V3 btf_pdf4(const BTF& b, int idx, const BtfCoords& tc) {
  const int np = b.d.num_phi, lo = tc.i_phi, hi = lo != np - 1 ? lo + 1 : 0;
  const float w = tc.w_phi;
  const V3 o1 = btf_pdf3(b, b.pdf4d_slices[(size_t)idx * np + lo], tc);
  const V3 o2 = btf_pdf3(b, b.pdf4d_slices[(size_t)idx * np + hi], tc);
  const float s1 = b.pdf4d_scales[(size_t)idx * np + lo] * (1.0f - w);
  const float s2 = b.pdf4d_scales[(size_t)idx * np + hi] * w;
  return o1 * s1 + o2 * s2;
}
// ConvertThetaPhiToBetaAlpha, SharedCoordinates.cuh:126-165 (radians in, radians out).  With m_codeBtfFlag the
// reference subtracts tc.m_phi — the phi left in the by-value copy of BTFBase::m_sharedCoordinates, which is never
// set before this call (BTFBase.cuh:65-66): restated as 0.
static inline void btf_theta_phi_to_beta_alpha(const BTF& b, float theta, float phi, float* beta, float* alpha) {
  float st, ct, sp, cp;
  sincos_rad(theta, &st, &ct);
  sincos_rad(phi, &sp, &cp);
  if (b.d.code_btf_flag) {
    const float x = cp * st, y = sp * st;
    *beta = asinp(clampf_(y, -1.0f, 1.0f));
    float sb, cb;
    sincos_rad(*beta, &sb, &cb);
    if (cb < 0.001f) {
      *alpha = 0.0f;
      return;
    }
    *alpha = asinp(clampf_(-x / cb, -1.0f, 1.0f));
    return;
  }
  const float x = cp * st, y = sp * st;
  *beta = asinp(x);
  float sb, cb;
  sincos_rad(*beta, &sb, &cb);
  if (cb < 0.001f) {
    *alpha = 0.0f;
    return;
  }
  *alpha = asinp(clampf_(y / cb, -1.0f, 1.0f));
}
// BTFBase::GetValueDeg (BTFBase.cuh:49-81) -> PDF6D::GetValDeg2 (PDF6D.cuh:32-95)
static inline V3 btf_value_deg(const BTF& b, float u, float v, float illum_theta, float illum_phi, float view_theta,
                               float view_phi) {
  if (illum_theta > 90.0f || view_theta > 90.0f) return v3(0, 0, 0);
  V3 out = v3(1.0f, 1.0f, 1.0f);  // SurfaceCompressedBTF::GetValue sets 1 before the call (RDD.hpp:334)
  do {
    const rtf_btf_desc& d = b.d;
    int x = (int)((u * d.tex_coord_multiplier) * (float)d.pdf6_cols);
    int y = (int)((v * d.tex_coord_multiplier) * (float)d.pdf6_rows);
    x -= d.cols_offset;
    while (x < 0) x += d.pdf6_cols;
    y -= d.rows_offset;
    while (y < 0) y += d.pdf6_rows;
    x %= d.pdf6_cols, y %= d.pdf6_rows;
    view_phi = modf_glm(360.0f - view_phi, 360.0f);
    illum_phi = modf_glm((360.0f - illum_phi) - (90.0f + view_phi), 360.0f);
    float beta, alpha;
    btf_theta_phi_to_beta_alpha(b, radiansf(illum_theta), radiansf(illum_phi), &beta, &alpha);
    alpha = degreesf(alpha), beta = degreesf(beta);
    if (beta != beta || alpha != alpha || view_theta != view_theta || view_phi != view_phi) break;  // PDF6D.cuh:62-68
    BtfCoords tc;
    {  // SetForAngleBetaDeg, SharedCoordinates.cuh:49-72
      const float be = clampf_(beta, -90.0f, 90.0f);
      if (d.use_cos_beta) {
        float sb, cb;
        sincos_rad(radiansf(be), &sb, &cb);
        tc.i_beta = clampi((int)((sb + 1.0f) / 2.0f * (float)(d.num_beta - 1)), 0, d.num_beta - 2);
        tc.w_beta = (be - b.beta_angles[tc.i_beta]) / (b.beta_angles[tc.i_beta + 1] - b.beta_angles[tc.i_beta]);
      } else {
        const float step = 180.0f / (float)(d.num_beta - 1);
        tc.i_beta = clampi((int)((be + 90.0f) / step), 0, d.num_beta - 2);
        tc.w_beta = (be + 90.0f - (float)tc.i_beta * step) / step;
      }
    }
    {  // SetForAngleAlphaDeg, :75-83
      const float al = clampf_(alpha, -90.0f, 90.0f), step = 180.f / (float)(d.num_alpha - 1);
      tc.i_alpha = clampi((int)((al + 90.0f) / step), 0, d.num_alpha - 2);
      tc.w_alpha = (al + 90.f - (float)tc.i_alpha * step) / step;
    }
    {  // SetForAngleThetaDeg, :97-105
      const float th = clampf_(view_theta, 0.0f, 90.0f), step = 90.0f / (float)(d.num_theta - 1);
      tc.i_theta = clampi((int)(th / step), 0, d.num_theta - 2);
      tc.w_theta = (th - (float)tc.i_theta * step) / step;
    }
    {  // SetForAnglePhiDeg, :86-94
      const float ph = clampf_(view_phi, 0.0f, 360.0f), step = 360.0f / (float)d.num_phi;
      tc.i_phi = clampi((int)(ph / step), 0, d.num_phi - 1);
      tc.w_phi = (ph - (float)tc.i_phi * step) / step;
    }
    const size_t texel = (size_t)y * d.pdf6_cols + x;
    out = btf_pdf4(b, b.pdf6_slices[texel] - 1, tc) * b.pdf6_scales[texel];
  } while (false);
  out = out * b.d.multiplier;
  if (b.d.hdr) out = out * (1.0f / b.d.hdr_value);
  return v3(powf(out.x, b.d.gamma), powf(out.y, b.d.gamma), powf(out.z, b.d.gamma));
}
// SurfaceCompressedBTF::ComputeAngles, RDD.hpp:290-326 (degrees out)
static inline void btf_compute_angles(V3 dir, V3 normal, V3 tangent, float* theta, float* phi) {
  const V3 B = cross(normal, tangent);
  const float tx = dot(tangent, dir), ty = dot(B, dir), tz = dot(normal, dir);
  if (tx != tx) {
    *theta = 0.f, *phi = 0.f;
    return;
  }
  if (tz < 0.0f) {
    *phi = 0.0f, *theta = 90.0f;
    return;
  }
  *theta = degreesf(acosp(tz));
  float p = degreesf(atan2p(ty, tx)) + 360.0f;
  if (p > 360.f) p -= 360.f;
  *phi = p;
}
// SurfaceCompressedBTF::GetValue, RDD.hpp:329-354
static inline V3 btf_get_value(const BTF& b, float u, float v, V3 view_dir, V3 illum_dir, V3 normal, V3 tangent) {
  float it, ip, vt, vp;
  btf_compute_angles(neg(view_dir), normal, tangent, &vt, &vp);
  btf_compute_angles(illum_dir, normal, tangent, &it, &ip);
  return btf_value_deg(b, u, v, it, ip, vt, vp) * (1.0f / 256.0f);
}

}  // namespace orc
