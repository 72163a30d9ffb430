// oracle/ref_stubs/glm/glm.hpp — TEST INFRASTRUCTURE ONLY.
// A minimal stand-in for the parts of glm (absent from this image) that the reference's MLVQ headers use
// (include/RayTracerFacility/MLVQ/*.cuh: vec2, vec3, clamp, isnan, sin, cos, asin, radians, degrees, mod, pow), written
// from glm's documented semantics so that those headers compile UNMODIFIED on the host into oracle/_ref/libbtf_ref.so.
// Not a copy of glm; scalar functions forward to <cmath> exactly as glm's do.
#pragma once
#include <cmath>
#include <cstdio>

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

struct vec3 {
  float x, y, z;
  vec3() : x(0), y(0), z(0) {}
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
}  // namespace glm
