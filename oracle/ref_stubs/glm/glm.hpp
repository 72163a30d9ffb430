// oracle/ref_stubs/glm/glm.hpp — TEST INFRASTRUCTURE ONLY.
// A minimal stand-in for the parts of glm (absent from this image) that the reference's MLVQ headers use
// (include/RayTracerFacility/MLVQ/*.cuh: vec2, vec3, clamp, isnan, sin, cos, asin, radians, degrees, mod, pow), written
// from glm's documented semantics so that those headers compile UNMODIFIED on the host into oracle/_ref/libbtf_ref.so
// (and, with the additions at the end, RayTracer/{RayTracerUtilities,Environment,BSDF}.cuh into oracle/_ref/libenv_ref.so).
// Not a copy of glm; scalar functions forward to <cmath> exactly as glm's do.
#pragma once
#include <cmath>
#include <cstdio>
// The reference calls cos / sin / asin / abs / sqrt UNQUALIFIED on floats in places (e.g. SharedCoordinates.cuh:134-146).
// In CUDA device code those resolve to the float overloads; under g++ they do so only when the C++ <math.h> wrapper has
// brought std::'s overloads into the global namespace — with <cmath> alone they would silently run in double.
#include <math.h>
#include <stdlib.h>

namespace glm {
struct vec2 {
  float x, y;
  vec2() : x(0), y(0) {}
  vec2(float a, float b) : x(a), y(b) {}
  explicit vec2(float a) : x(a), y(a) {}
  float& operator[](int i) { return i == 0 ? x : y; }
  const float& operator[](int i) const { return i == 0 ? x : y; }
};
inline vec2 operator*(const vec2& a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2 operator*(float s, const vec2& a) { return vec2(a.x * s, a.y * s); }

struct vec4;
struct vec3 {
  float x, y, z;
  vec3() : x(0), y(0), z(0) {}
  vec3(const vec4& v);  // glm's vec<3>(vec<4>) is implicit unless GLM_FORCE_EXPLICIT_CTOR
  vec3(float a, float b, float c) : x(a), y(b), z(c) {}
  explicit vec3(float a) : x(a), y(a), z(a) {}
  vec3(int a, int b, int c) : x((float)a), y((float)b), z((float)c) {}
  float& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
  const float& operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
  vec3& operator+=(const vec3& o) { x += o.x, y += o.y, z += o.z; return *this; }
  vec3& operator-=(const vec3& o) { x -= o.x, y -= o.y, z -= o.z; return *this; }
  vec3& operator*=(const vec3& o) { x *= o.x, y *= o.y, z *= o.z; return *this; }
  vec3& operator*=(float s) { x *= s, y *= s, z *= s; return *this; }
  vec3& operator/=(float s) { x /= s, y /= s, z /= s; return *this; }
};
inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(const vec3& a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, const vec3& a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(const vec3& a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator-(const vec3& a) { return vec3(-a.x, -a.y, -a.z); }

template <class T>
inline T clamp(T v, T lo, T hi) { return v < lo ? lo : (hi < v ? hi : v); }  // min(max(v, lo), hi)
inline bool isnan(float v) { return std::isnan(v); }
inline float sin(float v) { return std::sin(v); }
inline float cos(float v) { return std::cos(v); }
inline float asin(float v) { return std::asin(v); }
inline float acos(float v) { return std::acos(v); }
inline float radians(float d) { return d * 0.01745329251994329576923690768489f; }
inline float degrees(float r) { return r * 57.295779513082320876798154814105f; }
inline float mod(float x, float y) { return x - y * std::floor(x / y); }
inline float pow(float a, float b) { return std::pow(a, b); }
inline vec3 pow(const vec3& a, const vec3& b) { return vec3(std::pow(a.x, b.x), std::pow(a.y, b.y), std::pow(a.z, b.z)); }
template <class T>
inline T pi() { return T(3.14159265358979323846264338327950288); }

// ---- additions for RayTracer/{RayTracer.hpp, RayTracerUtilities.cuh, Environment.cuh, BSDF.cuh} ------------------------
// Operation order follows glm's scalar (non-SIMD) code paths: dot = (x*x' + y*y') + z*z'; cross, normalize = v *
// inversesqrt(dot(v, v)) with inversesqrt(x) = 1 / sqrt(x); length = sqrt(dot(v, v)); mat3 * vec3 = columns scaled and
// summed left to right; angle(x, y) = acos(clamp(dot(x, y), -1, 1)).
typedef unsigned int uint;
struct ivec2 {
  int x, y;
  ivec2() : x(0), y(0) {}
  ivec2(int a, int b) : x(a), y(b) {}
};
struct uvec3 {
  unsigned x, y, z;
  uvec3() : x(0), y(0), z(0) {}
  uvec3(unsigned a, unsigned b, unsigned c) : x(a), y(b), z(c) {}
};
struct vec4 {
  float x, y, z, w;
  vec4() : x(0), y(0), z(0), w(0) {}
  vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
  explicit vec4(float a) : x(a), y(a), z(a), w(a) {}
  vec4(const vec3& v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
  float& operator[](int i) { return (&x)[i]; }
  const float& operator[](int i) const { return (&x)[i]; }
  vec4& operator/=(float s) { x /= s, y /= s, z /= s, w /= s; return *this; }
};
inline vec3::vec3(const vec4& v) : x(v.x), y(v.y), z(v.z) {}
inline vec4 operator+(const vec4& a, const vec4& b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(const vec4& a, const vec4& b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline vec4 operator*(const vec4& a, const vec4& b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline vec4 operator*(const vec4& a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator*(float s, const vec4& a) { return vec4(s * a.x, s * a.y, s * a.z, s * a.w); }
inline vec4 operator/(const vec4& a, float s) { return vec4(a.x / s, a.y / s, a.z / s, a.w / s); }
struct quat {
  float x, y, z, w;
  quat() : x(0), y(0), z(0), w(1) {}
};
struct mat4 {
  vec4 c[4];
  mat4() {}
  explicit mat4(float d) { c[0].x = d, c[1].y = d, c[2].z = d, c[3].w = d; }
  vec4& operator[](int i) { return c[i]; }
  const vec4& operator[](int i) const { return c[i]; }
};
struct mat3x3 {
  vec3 c[3];
  mat3x3() {}
  mat3x3(const vec3& a, const vec3& b, const vec3& d) { c[0] = a, c[1] = b, c[2] = d; }
  vec3& operator[](int i) { return c[i]; }
  const vec3& operator[](int i) const { return c[i]; }
};
typedef mat3x3 mat3;
// mat4 * scalar and mat4 += mat4, glm/detail/type_mat4x4.inl: column by column (CopySkinnedVerticesKernel's bone blend)
inline mat4 operator*(const mat4& m, float s) {
  mat4 r;
  r[0] = m[0] * s, r[1] = m[1] * s, r[2] = m[2] * s, r[3] = m[3] * s;
  return r;
}
inline mat4& operator+=(mat4& a, const mat4& b) {
  a[0] = a[0] + b[0], a[1] = a[1] + b[1], a[2] = a[2] + b[2], a[3] = a[3] + b[3];
  return a;
}
// mat4 * vec4, glm/detail/type_mat4x4.inl: (m[0] * x + m[1] * y) + (m[2] * z + m[3] * w)
inline vec4 operator*(const mat4& m, const vec4& v) { return (m[0] * v.x + m[1] * v.y) + (m[2] * v.z + m[3] * v.w); }
inline bool operator==(const vec3& a, const vec3& b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline bool operator!=(const vec3& a, const vec3& b) { return !(a == b); }
inline vec2 operator+(const vec2& a, const vec2& b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec3 operator*(const mat3x3& m, const vec3& v) {
  return vec3(m[0].x * v.x + m[1].x * v.y + m[2].x * v.z, m[0].y * v.x + m[1].y * v.y + m[2].y * v.z,
              m[0].z * v.x + m[1].z * v.y + m[2].z * v.z);
}
inline float sqrt(float v) { return std::sqrt(v); }
inline float exp(float v) { return std::exp(v); }
inline float abs(float v) { return std::fabs(v); }
inline float max(float a, float b) { return a < b ? b : a; }
inline float min(float a, float b) { return b < a ? b : a; }
inline vec3 max(const vec3& a, const vec3& b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline float dot(const vec3& a, const vec3& b) {
  const vec3 t(a * b);
  return t.x + t.y + t.z;
}
inline vec3 cross(const vec3& x, const vec3& y) {
  return vec3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
inline float inversesqrt(float v) { return 1.0f / std::sqrt(v); }
inline vec3 normalize(const vec3& v) { return v * inversesqrt(dot(v, v)); }
inline float length(const vec3& v) { return std::sqrt(dot(v, v)); }
inline float distance(const vec3& a, const vec3& b) { return length(b - a); }
inline float angle(const vec3& x, const vec3& y) { return std::acos(clamp(dot(x, y), -1.0f, 1.0f)); }
inline int clamp(int v, int lo, int hi) { return v < lo ? lo : (hi < v ? hi : v); }
// gtx/rotate_vector.hpp: rotate(v, angle, normal) = mat3(rotate(mat4(1), angle, normal)) * v, with the rotation matrix
// of gtc/matrix_transform (c + t_i * a_i on the diagonal, t_i * a_j +- s * a_k off it, t = (1 - c) * axis).
inline vec3 rotate(const vec3& v, float angle, const vec3& normal) {
  const float c = std::cos(angle), s = std::sin(angle);
  const vec3 a(normalize(normal));
  const vec3 t((1.0f - c) * a);
  mat3x3 r;
  r[0] = vec3(c + t.x * a.x, t.x * a.y + s * a.z, t.x * a.z - s * a.y);
  r[1] = vec3(t.y * a.x - s * a.z, c + t.y * a.y, t.y * a.z + s * a.x);
  r[2] = vec3(t.z * a.x + s * a.y, t.z * a.y - s * a.x, c + t.z * a.z);
  return r * v;
}
}  // namespace glm
