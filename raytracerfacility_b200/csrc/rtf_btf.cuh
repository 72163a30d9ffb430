// rtf_btf.cuh — device side of the material inputs that are tables: images in linear memory (textures, cube-map
// faces) and the compressed BTF (MLVQ) evaluator.
//
// Reference (relative to /root/reference/include/RayTracerFacility/):
//   RayTracer/RayDataDefinations.hpp:240-284   SurfaceMaterial::GetAlbedo/GetRoughness/GetMetallic/ApplyNormalTexture
//   RayTracer/RayDataDefinations.hpp:286-354   SurfaceCompressedBTF::ComputeAngles / GetValue
//   RayTracer/RayTracerUtilities.cuh:54-80     SampleCubeMap
//   MLVQ/BTFBase.cuh:49-81 -> PDF6D.cuh:32-95 -> PDF4D.cuh:33-63 -> PDF3D.cuh:31-46 -> PDF2D.cuh:46-141 ->
//   PDF1D.cuh:24-69 (Hermite) / IndexAB.cuh:26-46 / VectorColor.cuh:22-28 -> CIELab.cuh:8-186
// The reference binds GL textures through CUDA interop (RT.cu:2233-2256, 270-347); there is no GL on this path, so the
// images live in plain device memory and tex2D's wrap + linear + normalised-coordinate lookup is written out.
// One BTF evaluation = 2 phi x 2 theta x (2 alpha luminance + 2 alpha colour) x 2..4 beta table reads: pure gather
// arithmetic, L2-resident tables (a few MB), no tensor-core shape anywhere.
#pragma once
#include "rtf_math.cuh"

namespace rtf {

// ---- portable inverse trigonometry (same polynomials as the CPU oracle; only + - * / sqrt) ----------------------
RTF_HD float atan01(float a) {
  float off = 0.0f;
  if (a > 0.4142135679721832275390625f) {
    a = (a - 1.0f) / (a + 1.0f);
    off = 0.785398185253143310546875f;
  }
  const float z = a * a;
  float p = 8.05374449538e-2f;
  p = p * z - 1.38776856032e-1f;
  p = p * z + 1.99777106478e-1f;
  p = p * z - 3.33329491539e-1f;
  return off + (a + a * (z * p));
}
RTF_HD float atan2p(float y, float x) {
  if (y != y || x != x) return u2f(0x7fc00000u);
  const float ax = fabsf(x), ay = fabsf(y);
  const float hi = ax > ay ? ax : ay, lo = ax > ay ? ay : ax;
  float r = hi > 0.0f ? atan01(lo / hi) : 0.0f;
  if (ay > ax) r = 1.57079637050628662109375f - r;
  if (x < 0.0f) r = 3.1415927410125732421875f - r;
  return y < 0.0f ? -r : r;
}
RTF_HD float asinp(float v) { return atan2p(v, sqrtf((1.0f - v) * (1.0f + v))); }
RTF_HD float acosp(float v) { return atan2p(sqrtf((1.0f - v) * (1.0f + v)), v); }
RTF_HD void sincos_rad(float a, float& s, float& c) { sincos2pi(a * 0.15915493667125701904296875f, s, c); }
RTF_HD float radiansf(float d) { return d * 0.01745329238474369049072265625f; }
RTF_HD float degreesf(float r) { return r * 57.295780181884765625f; }
RTF_HD float mod_glm(float x, float y) { return x - y * floorf(x / y); }

// ---- images ----------------------------------------------------------------------------------------------------------
struct DevTexture {
  const float4* texels = nullptr;  // width*height, row-major
  int w = 0, h = 0;
};
RTF_HD int wrapi(int i, int n) {
  const int m = i % n;
  return m < 0 ? m + n : m;
}
// tex2D<float4>: cudaAddressModeWrap, cudaFilterModeLinear, normalizedCoords (RT.cu:2246-2252)
RTF_D float4 tex2d(const DevTexture& t, float u, float v) {
  const float x = (u - floorf(u)) * (float)t.w - 0.5f, y = (v - floorf(v)) * (float)t.h - 0.5f;
  const float fx = floorf(x), fy = floorf(y);
  const float a = x - fx, b = y - fy;
  const int x0 = wrapi((int)fx, t.w), x1 = wrapi((int)fx + 1, t.w), y0 = wrapi((int)fy, t.h), y1 = wrapi((int)fy + 1, t.h);
  const float4 t00 = __ldg(t.texels + (size_t)y0 * t.w + x0), t10 = __ldg(t.texels + (size_t)y0 * t.w + x1);
  const float4 t01 = __ldg(t.texels + (size_t)y1 * t.w + x0), t11 = __ldg(t.texels + (size_t)y1 * t.w + x1);
  const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
  float4 o;
  o.x = ((w00 * t00.x + w10 * t10.x) + w01 * t01.x) + w11 * t11.x;
  o.y = ((w00 * t00.y + w10 * t10.y) + w01 * t01.y) + w11 * t11.y;
  o.z = ((w00 * t00.z + w10 * t10.z) + w01 * t01.z) + w11 * t11.z;
  o.w = ((w00 * t00.w + w10 * t10.w) + w01 * t01.w) + w11 * t11.w;
  return o;
}
// SampleCubeMap, RTU.cuh:54-80 (faces +X -X +Y -Y +Z -Z)
RTF_D float4 sample_cube_map(const DevTexture* faces, float3 d) {
  const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
  float ma, u, v;
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
  return tex2d(faces[face], u * ma + 0.5f, v * ma + 0.5f);
}

// ---- compressed BTF --------------------------------------------------------------------------------------------------
struct DevBTF {
  int use_cos_beta, num_beta, num_alpha, num_theta, num_phi, code_btf_flag, hdr;
  float hdr_value, multiplier, tex_coord_multiplier, gamma;
  int rows_offset, cols_offset, pdf6_rows, pdf6_cols, n_channels, pdf2d_slice_len;
  const int* pdf6_slices;
  const float* pdf6_scales;
  const float* pdf1d_basis;
  const float* colors;
  const int* index_ab;
  const int* pdf2d_colors;
  const int* pdf2d_slices;
  const float* pdf2d_scales;
  const int* pdf2d_indices;
  const float* pdf3d_scales;
  const int* pdf3d_slices;
  const float* pdf4d_scales;
  const int* pdf4d_slices;
  const float* beta_angles;  // num_beta, degrees (CompressedBTF.cpp:162-186)
};
struct BtfCoords {  // SharedCoordinates' per-query part (SharedCoordinates.cuh:21-44)
  int i_beta, i_alpha, i_theta, i_phi;
  float w_beta, w_alpha, w_theta, w_phi;
};
RTF_HD int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// PDF1D::GetVal, PDF1D.cuh:24-69: Fergusson cubic through p0,p1 with the tangents as the reference writes them.
RTF_D float btf_pdf1(const DevBTF& b, int slice, const BtfCoords& tc) {
  const int nb = b.num_beta, i = tc.i_beta;
  const float* row = b.pdf1d_basis + (size_t)slice * nb;
  const float w = tc.w_beta, p0 = __ldg(row + i), p1 = __ldg(row + i + 1);
  const float m0h = i == 0 ? p1 - p0 : 0.5f * (p1 - __ldg(row + i - 1));
  const float m1h = i == nb - 2 ? p1 - p0 : 0.5f * (p1 - p0);  // PDF1D.cuh:51-55 re-reads index i+1
  const float t2 = w * w, t3 = t2 * w;
  const float h01 = -2.0f * t3 + 3.0f * t2, h00 = 1.0f - h01, h11 = t3 - t2, h10 = h11 - t2 + w;
  return ((h00 * p0 + h10 * m0h) + h01 * p1) + h11 * m1h;
}
// IndexAB::GetVal, IndexAB.cuh:36-46 + VectorColor::Get, VectorColor.cuh:22-28
RTF_D float2 btf_iab(const DevBTF& b, int slice, const BtfCoords& tc) {
  const int* row = b.index_ab + (size_t)slice * b.num_beta;
  const int c0 = __ldg(row + tc.i_beta), c1 = __ldg(row + tc.i_beta + 1);
  const float* a0 = b.colors + (size_t)c0 * b.n_channels;
  const float* a1 = b.colors + (size_t)c1 * b.n_channels;
  float2 ab;
  ab.x = (1.f - tc.w_beta) * __ldg(a0) + tc.w_beta * __ldg(a1);
  ab.y = (1.f - tc.w_beta) * __ldg(a0 + 1) + tc.w_beta * __ldg(a1 + 1);
  return ab;
}
// CIELab.cuh:8-186
RTF_D float3 btf_clamp0(float3 c) { return f3(c.x < 0.f ? 0.f : c.x, c.y < 0.f ? 0.f : c.y, c.z < 0.f ? 0.f : c.z); }
RTF_D float3 btf_ycbcr(float3 y, bool normalized) {
  const float s = normalized ? 256.0f : 1.0f;
  float3 rgb;
  rgb.x = y.x * 1.1643828f + y.z * 1.5960273f - 222.921f / s;
  rgb.y = y.x * 1.1643828f - y.y * 0.39176172f - y.z * 0.81296875f + 135.576f / s;
  rgb.z = y.x * 1.1643828f + y.y * 2.0172344f - 276.836f / s;
  return btf_clamp0(rgb);
}
RTF_D float btf_srgb(float c) { return c > 0.0031308f ? 1.055f * powf(c, 1.f / 2.4f) - 0.055f : 12.92f * c; }
RTF_D float3 btf_logluv(float3 luv, bool normalized) {
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
  float3 rgb;
  rgb.x = X * 3.2406f + Y * -1.5372f + Z * -0.4986f;
  rgb.y = X * -0.9689f + Y * 1.8758f + Z * 0.0415f;
  rgb.z = X * 0.0557f + Y * -0.2040f + Z * 1.0570f;
  rgb = btf_clamp0(rgb);
  return f3(btf_srgb(rgb.x), btf_srgb(rgb.y), btf_srgb(rgb.z));
}
// PDF2DSeperate::GetVal, PDF2D.cuh:121-141 (colour :46-60, luminance :83-103) + UserCmToRgb, CIELab.cuh:168-185
RTF_D float3 btf_pdf2(const DevBTF& b, int idx, const BtfCoords& tc) {
  const int na = b.num_alpha, i = tc.i_alpha;
  const float w = tc.w_alpha;
  const int lum = __ldg(b.pdf2d_indices + (size_t)idx * b.pdf2d_slice_len);
  const int col = __ldg(b.pdf2d_indices + (size_t)idx * b.pdf2d_slice_len + 1);
  float3 cm;
  const float2 ab1 = btf_iab(b, __ldg(b.pdf2d_colors + (size_t)col * na + i), tc);
  const float2 ab2 = btf_iab(b, __ldg(b.pdf2d_colors + (size_t)col * na + i + 1), tc);
  cm.y = ab1.x * (1.0f - w) + ab2.x * w;
  cm.z = ab1.y * (1.0f - w) + ab2.y * w;
  const float l1 = __ldg(b.pdf2d_scales + (size_t)lum * na + i) * btf_pdf1(b, __ldg(b.pdf2d_slices + (size_t)lum * na + i), tc);
  const float l2 =
      __ldg(b.pdf2d_scales + (size_t)lum * na + i + 1) * btf_pdf1(b, __ldg(b.pdf2d_slices + (size_t)lum * na + i + 1), tc);
  cm.x = (1.f - w) * l1 + w * l2;
  if (b.hdr) return btf_logluv(cm, b.code_btf_flag != 0);
  return btf_ycbcr(cm, b.code_btf_flag != 0);
}
// PDF3D::GetVal, PDF3D.cuh:31-46
RTF_D float3 btf_pdf3(const DevBTF& b, int idx, const BtfCoords& tc) {
  const int nt = b.num_theta, i = tc.i_theta;
  const float w = tc.w_theta;
  const float3 o1 = btf_pdf2(b, __ldg(b.pdf3d_slices + (size_t)idx * nt + i), tc);
  const float3 o2 = btf_pdf2(b, __ldg(b.pdf3d_slices + (size_t)idx * nt + i + 1), tc);
  const float s1 = __ldg(b.pdf3d_scales + (size_t)idx * nt + i) * (1.0f - w);
  const float s2 = __ldg(b.pdf3d_scales + (size_t)idx * nt + i + 1) * w;
  return o1 * s1 + o2 * s2;
}
// PDF4D::GetVal, PDF4D.cuh:33-63 (phi wraps around)
RTF_D float3 btf_pdf4(const DevBTF& b, int idx, const BtfCoords& tc) {
  const int np = b.num_phi, lo = tc.i_phi, hi = lo != np - 1 ? lo + 1 : 0;
  const float w = tc.w_phi;
  const float3 o1 = btf_pdf3(b, __ldg(b.pdf4d_slices + (size_t)idx * np + lo), tc);
  const float3 o2 = btf_pdf3(b, __ldg(b.pdf4d_slices + (size_t)idx * np + hi), tc);
  const float s1 = __ldg(b.pdf4d_scales + (size_t)idx * np + lo) * (1.0f - w);
  const float s2 = __ldg(b.pdf4d_scales + (size_t)idx * np + hi) * w;
  return o1 * s1 + o2 * s2;
}
// The same value as btf_pdf4 — every float operation and its order unchanged — with the table reads issued level by
// level: the cascade is a chain of dependent gathers (PDF4D slice -> PDF3D slice -> PDF2D index pair -> alpha slices ->
// beta rows -> colour vectors), and walked depth-first a thread has two or three loads in flight while it waits an L2
// round trip per link, sixteen times over.  Here the 2 phi x 2 theta corners advance together (2, 4, 8 loads per level),
// then each phi pair finishes its 2 theta x 2 alpha leaves with 12 + 20 + 16 loads in flight: 9 round trips instead of
// ~30, at the cost of registers (the BTF shading kernel is compiled for 2 blocks of 256 threads per SM).
RTF_D float3 btf_pdf4_wide(const DevBTF& b, int idx, const BtfCoords& tc) {
  const int np = b.num_phi, lo = tc.i_phi, hi = lo != np - 1 ? lo + 1 : 0;
  const int nt = b.num_theta, it = tc.i_theta, na = b.num_alpha, ia = tc.i_alpha, nb = b.num_beta, ib = tc.i_beta;
  const int* p4 = b.pdf4d_slices + (size_t)idx * np;
  const float* q4 = b.pdf4d_scales + (size_t)idx * np;
  const int i4[2] = {__ldg(p4 + lo), __ldg(p4 + hi)};
  const float s4[2] = {__ldg(q4 + lo), __ldg(q4 + hi)};
  int i3[4];
  float s3[4];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    i3[k] = __ldg(b.pdf3d_slices + (size_t)i4[k >> 1] * nt + it + (k & 1));
    s3[k] = __ldg(b.pdf3d_scales + (size_t)i4[k >> 1] * nt + it + (k & 1));
  }
  int lum[4], col[4];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    lum[k] = __ldg(b.pdf2d_indices + (size_t)i3[k] * b.pdf2d_slice_len);
    col[k] = __ldg(b.pdf2d_indices + (size_t)i3[k] * b.pdf2d_slice_len + 1);
  }
  const float wa = tc.w_alpha, wb = tc.w_beta, wt = tc.w_theta;
  const float t2 = wb * wb, t3 = t2 * wb;
  const float h01 = -2.0f * t3 + 3.0f * t2, h00 = 1.0f - h01, h11 = t3 - t2, h10 = h11 - t2 + wb;
  float3 o3[2];
#pragma unroll
  for (int c = 0; c < 2; c++) {  // one phi corner: 2 theta x 2 alpha leaves, j = 2*theta + alpha
    int cs[4], ls[4];
    float lsc[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int k = 2 * c + (j >> 1), a = j & 1;
      cs[j] = __ldg(b.pdf2d_colors + (size_t)col[k] * na + ia + a);
      ls[j] = __ldg(b.pdf2d_slices + (size_t)lum[k] * na + ia + a);
      lsc[j] = __ldg(b.pdf2d_scales + (size_t)lum[k] * na + ia + a);
    }
    int c0[4], c1[4];
    float p0[4], p1[4], pm[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int* row = b.index_ab + (size_t)cs[j] * nb;
      c0[j] = __ldg(row + ib), c1[j] = __ldg(row + ib + 1);
      const float* br = b.pdf1d_basis + (size_t)ls[j] * nb;
      p0[j] = __ldg(br + ib), p1[j] = __ldg(br + ib + 1);
      pm[j] = ib == 0 ? 0.0f : __ldg(br + ib - 1);
    }
    float ax[4], ay[4], bx[4], by[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const float* a0 = b.colors + (size_t)c0[j] * b.n_channels;
      const float* a1 = b.colors + (size_t)c1[j] * b.n_channels;
      ax[j] = __ldg(a0), ay[j] = __ldg(a0 + 1), bx[j] = __ldg(a1), by[j] = __ldg(a1 + 1);
    }
    float3 o2[2];
#pragma unroll
    for (int t = 0; t < 2; t++) {
      float l[2];
      float2 ab[2];
#pragma unroll
      for (int a = 0; a < 2; a++) {
        const int j = 2 * t + a;
        const float m0h = ib == 0 ? p1[j] - p0[j] : 0.5f * (p1[j] - pm[j]);
        const float m1h = ib == nb - 2 ? p1[j] - p0[j] : 0.5f * (p1[j] - p0[j]);
        l[a] = lsc[j] * (((h00 * p0[j] + h10 * m0h) + h01 * p1[j]) + h11 * m1h);
        ab[a].x = (1.f - wb) * ax[j] + wb * bx[j];
        ab[a].y = (1.f - wb) * ay[j] + wb * by[j];
      }
      float3 cm;
      cm.y = ab[0].x * (1.0f - wa) + ab[1].x * wa;
      cm.z = ab[0].y * (1.0f - wa) + ab[1].y * wa;
      cm.x = (1.f - wa) * l[0] + wa * l[1];
      o2[t] = b.hdr ? btf_logluv(cm, b.code_btf_flag != 0) : btf_ycbcr(cm, b.code_btf_flag != 0);
    }
    const float s1 = s3[2 * c] * (1.0f - wt), s2 = s3[2 * c + 1] * wt;
    o3[c] = o2[0] * s1 + o2[1] * s2;
  }
  const float w = tc.w_phi;
  const float s1 = s4[0] * (1.0f - w), s2 = s4[1] * w;
  return o3[0] * s1 + o3[1] * s2;
}
// ConvertThetaPhiToBetaAlpha, SharedCoordinates.cuh:126-165.  With m_codeBtfFlag the reference subtracts tc.m_phi, a
// member of the by-value SharedCoordinates copy that nothing sets before the call (BTFBase.cuh:65-66): 0 here.
RTF_D void btf_theta_phi_to_beta_alpha(const DevBTF& b, float theta, float phi, float& beta, float& alpha) {
  float st, ct, sp, cp;
  sincos_rad(theta, st, ct);
  sincos_rad(phi, sp, cp);
  const float x = cp * st, y = sp * st;
  float sb, cb;
  if (b.code_btf_flag) {
    beta = asinp(clampf(y, -1.0f, 1.0f));
    sincos_rad(beta, sb, cb);
    if (cb < 0.001f) {
      alpha = 0.0f;
      return;
    }
    alpha = asinp(clampf(-x / cb, -1.0f, 1.0f));
    return;
  }
  beta = asinp(x);
  sincos_rad(beta, sb, cb);
  if (cb < 0.001f) {
    alpha = 0.0f;
    return;
  }
  alpha = asinp(clampf(y / cb, -1.0f, 1.0f));
}
// BTFBase::GetValueDeg (BTFBase.cuh:49-81) -> PDF6D::GetValDeg2 (PDF6D.cuh:32-95)
template <bool WIDE = false>
RTF_D float3 btf_value_deg(const DevBTF& b, float u, float v, float illum_theta, float illum_phi, float view_theta,
                           float view_phi) {
  if (illum_theta > 90.0f || view_theta > 90.0f) return f3(0, 0, 0);
  float3 out = f3(1.0f, 1.0f, 1.0f);  // SurfaceCompressedBTF::GetValue sets 1 before the call (RDD.hpp:334)
  do {
    int x = (int)((u * b.tex_coord_multiplier) * (float)b.pdf6_cols);
    int y = (int)((v * b.tex_coord_multiplier) * (float)b.pdf6_rows);
    x -= b.cols_offset;
    while (x < 0) x += b.pdf6_cols;
    y -= b.rows_offset;
    while (y < 0) y += b.pdf6_rows;
    x %= b.pdf6_cols, y %= b.pdf6_rows;
    view_phi = mod_glm(360.0f - view_phi, 360.0f);
    illum_phi = mod_glm((360.0f - illum_phi) - (90.0f + view_phi), 360.0f);
    float beta, alpha;
    btf_theta_phi_to_beta_alpha(b, radiansf(illum_theta), radiansf(illum_phi), beta, alpha);
    alpha = degreesf(alpha), beta = degreesf(beta);
    if (beta != beta || alpha != alpha || view_theta != view_theta || view_phi != view_phi) break;  // PDF6D.cuh:62-68
    BtfCoords tc;
    {  // SetForAngleBetaDeg, SharedCoordinates.cuh:49-72
      const float be = clampf(beta, -90.0f, 90.0f);
      if (b.use_cos_beta) {
        float sb, cb;
        sincos_rad(radiansf(be), sb, cb);
        tc.i_beta = clampi((int)((sb + 1.0f) / 2.0f * (float)(b.num_beta - 1)), 0, b.num_beta - 2);
        const float a0 = __ldg(b.beta_angles + tc.i_beta), a1 = __ldg(b.beta_angles + tc.i_beta + 1);
        tc.w_beta = (be - a0) / (a1 - a0);
      } else {
        const float step = 180.0f / (float)(b.num_beta - 1);
        tc.i_beta = clampi((int)((be + 90.0f) / step), 0, b.num_beta - 2);
        tc.w_beta = (be + 90.0f - (float)tc.i_beta * step) / step;
      }
    }
    {  // SetForAngleAlphaDeg, :75-83
      const float al = clampf(alpha, -90.0f, 90.0f), step = 180.f / (float)(b.num_alpha - 1);
      tc.i_alpha = clampi((int)((al + 90.0f) / step), 0, b.num_alpha - 2);
      tc.w_alpha = (al + 90.f - (float)tc.i_alpha * step) / step;
    }
    {  // SetForAngleThetaDeg, :97-105
      const float th = clampf(view_theta, 0.0f, 90.0f), step = 90.0f / (float)(b.num_theta - 1);
      tc.i_theta = clampi((int)(th / step), 0, b.num_theta - 2);
      tc.w_theta = (th - (float)tc.i_theta * step) / step;
    }
    {  // SetForAnglePhiDeg, :86-94
      const float ph = clampf(view_phi, 0.0f, 360.0f), step = 360.0f / (float)b.num_phi;
      tc.i_phi = clampi((int)(ph / step), 0, b.num_phi - 1);
      tc.w_phi = (ph - (float)tc.i_phi * step) / step;
    }
    const size_t texel = (size_t)y * b.pdf6_cols + x;
    const int idx4 = __ldg(b.pdf6_slices + texel) - 1;
    out = (WIDE ? btf_pdf4_wide(b, idx4, tc) : btf_pdf4(b, idx4, tc)) * __ldg(b.pdf6_scales + texel);
  } while (false);
  out = out * b.multiplier;
  if (b.hdr) out = out * (1.0f / b.hdr_value);
  return f3(powf(out.x, b.gamma), powf(out.y, b.gamma), powf(out.z, b.gamma));
}
// SurfaceCompressedBTF::ComputeAngles, RDD.hpp:290-326 (degrees out)
RTF_D void btf_compute_angles(float3 dir, float3 normal, float3 tangent, float& theta, float& phi) {
  const float3 B = cross3(normal, tangent);
  const float tx = dot3(tangent, dir), ty = dot3(B, dir), tz = dot3(normal, dir);
  if (tx != tx) {
    theta = 0.f, phi = 0.f;
    return;
  }
  if (tz < 0.0f) {
    phi = 0.0f, theta = 90.0f;
    return;
  }
  theta = degreesf(acosp(tz));
  float p = degreesf(atan2p(ty, tx)) + 360.0f;
  if (p > 360.f) p -= 360.f;
  phi = p;
}
// SurfaceCompressedBTF::GetValue, RDD.hpp:329-354
template <bool WIDE = false>
RTF_D float3 btf_get_value(const DevBTF& b, float u, float v, float3 view_dir, float3 illum_dir, float3 normal, float3 tangent) {
  float it, ip, vt, vp;
  btf_compute_angles(neg3(view_dir), normal, tangent, vt, vp);
  btf_compute_angles(illum_dir, normal, tangent, it, ip);
  return btf_value_deg<WIDE>(b, u, v, it, ip, vt, vp) * (1.0f / 256.0f);
}

}  // namespace rtf
