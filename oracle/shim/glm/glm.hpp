// oracle/shim/glm/glm.hpp - TEST INFRASTRUCTURE ONLY.
// Minimal stand-in for the subset of glm that the reference's graph.cu /
// erosion.cu use (glm itself is an un-vendored system dependency that is not
// installed here).  Semantics follow glm: component-wise
// operators, length = sqrt(dot), float->int conversion truncates.
#pragma once
#include <cmath>
#include <cstddef>
#include <cuda_runtime.h>

#define SHIM_HD __host__ __device__ inline

namespace glm {

template <int N, typename T> struct vec;

template <typename T> struct vec<1, T> {  // silt::vec<1> of filter.cu / grad.cu: one float with vector operators
  T x;
  SHIM_HD vec() : x(0) {}
  SHIM_HD vec(T s) : x(s) {}
  SHIM_HD T& operator[](int) { return x; }
  SHIM_HD const T& operator[](int) const { return x; }
  SHIM_HD vec& operator+=(const vec& o) { x += o.x; return *this; }
};
template <typename T> SHIM_HD vec<1, T> operator+(const vec<1, T>& a, const vec<1, T>& b) { return vec<1, T>(a.x + b.x); }
template <typename T> SHIM_HD vec<1, T> operator-(const vec<1, T>& a, const vec<1, T>& b) { return vec<1, T>(a.x - b.x); }
template <typename T> SHIM_HD vec<1, T> operator*(const vec<1, T>& a, T s) { return vec<1, T>(a.x * s); }
template <typename T> SHIM_HD vec<1, T> operator*(T s, const vec<1, T>& a) { return vec<1, T>(s * a.x); }
template <typename T> SHIM_HD vec<1, T> operator/(const vec<1, T>& a, T s) { return vec<1, T>(a.x / s); }

template <typename T> struct vec<2, T> {
  T x, y;
  SHIM_HD vec() : x(0), y(0) {}
  SHIM_HD explicit vec(T s) : x(s), y(s) {}
  template <typename A, typename B> SHIM_HD vec(A a, B b) : x(static_cast<T>(a)), y(static_cast<T>(b)) {}
  template <typename U> SHIM_HD vec(const vec<2, U>& o) : x(static_cast<T>(o.x)), y(static_cast<T>(o.y)) {}
  template <typename U> SHIM_HD vec(const vec<3, U>& o);  // truncating, as glm
  SHIM_HD T& operator[](int i) { return i == 0 ? x : y; }
  SHIM_HD const T& operator[](int i) const { return i == 0 ? x : y; }
  SHIM_HD vec& operator+=(const vec& o) { x += o.x; y += o.y; return *this; }
};

template <typename T> struct vec<3, T> {
  T x, y, z;
  SHIM_HD vec() : x(0), y(0), z(0) {}
  SHIM_HD explicit vec(T s) : x(s), y(s), z(s) {}
  template <typename A, typename B, typename C>
  SHIM_HD vec(A a, B b, C c) : x(static_cast<T>(a)), y(static_cast<T>(b)), z(static_cast<T>(c)) {}
  template <typename U> SHIM_HD vec(const vec<3, U>& o) : x(static_cast<T>(o.x)), y(static_cast<T>(o.y)), z(static_cast<T>(o.z)) {}
  SHIM_HD T& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
  SHIM_HD const T& operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
  SHIM_HD vec& operator+=(const vec& o) { x += o.x; y += o.y; z += o.z; return *this; }
};

template <typename T> template <typename U> SHIM_HD vec<2, T>::vec(const vec<3, U>& o) : x(static_cast<T>(o.x)), y(static_cast<T>(o.y)) {}

using vec2 = vec<2, float>;
using vec3 = vec<3, float>;
using ivec2 = vec<2, int>;
using ivec3 = vec<3, int>;

#define SHIM_BINOP(op)                                                                                             \
  template <typename T> SHIM_HD vec<2, T> operator op(const vec<2, T>& a, const vec<2, T>& b) { return vec<2, T>(a.x op b.x, a.y op b.y); } \
  template <typename T> SHIM_HD vec<2, T> operator op(const vec<2, T>& a, T s) { return vec<2, T>(a.x op s, a.y op s); }                   \
  template <typename T> SHIM_HD vec<2, T> operator op(T s, const vec<2, T>& a) { return vec<2, T>(s op a.x, s op a.y); }                   \
  template <typename T> SHIM_HD vec<3, T> operator op(const vec<3, T>& a, const vec<3, T>& b) { return vec<3, T>(a.x op b.x, a.y op b.y, a.z op b.z); } \
  template <typename T> SHIM_HD vec<3, T> operator op(const vec<3, T>& a, T s) { return vec<3, T>(a.x op s, a.y op s, a.z op s); }         \
  template <typename T> SHIM_HD vec<3, T> operator op(T s, const vec<3, T>& a) { return vec<3, T>(s op a.x, s op a.y, s op a.z); }
SHIM_BINOP(+)
SHIM_BINOP(-)
SHIM_BINOP(*)
SHIM_BINOP(/)
#undef SHIM_BINOP

template <typename T> SHIM_HD vec<2, T> operator-(const vec<2, T>& a) { return vec<2, T>(-a.x, -a.y); }
template <typename T> SHIM_HD vec<3, T> operator-(const vec<3, T>& a) { return vec<3, T>(-a.x, -a.y, -a.z); }

SHIM_HD float max(float a, float b) { return a < b ? b : a; }  // glm::max(x, y) = (x < y) ? y : x
SHIM_HD float length(const vec<1, float>& a) { return sqrtf(a.x * a.x); }
SHIM_HD float dot(const vec2& a, const vec2& b) { return a.x * b.x + a.y * b.y; }
SHIM_HD float dot(const vec3& a, const vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
SHIM_HD float length(const vec2& a) { return sqrtf(dot(a, a)); }
SHIM_HD float length(const vec3& a) { return sqrtf(dot(a, a)); }
SHIM_HD vec2 floor(const vec2& a) { return vec2(floorf(a.x), floorf(a.y)); }
SHIM_HD vec3 normalize(const vec3& a) { return a * (1.0f / sqrtf(dot(a, a))); }

// sample.hpp: column-major 2x2 matrix, constexpr from four scalars, times a vec2
struct mat2 {
  float c0x, c0y, c1x, c1y;
  constexpr mat2(double a, double b, double c, double d) : c0x((float)a), c0y((float)b), c1x((float)c), c1y((float)d) {}
};
SHIM_HD vec2 operator*(const mat2& m, const vec2& v) { return vec2(m.c0x * v.x + m.c1x * v.y, m.c0y * v.x + m.c1y * v.y); }
template <typename T> struct vec<4, T> { T x, y, z, w; };
using vec4 = vec<4, float>;

}  // namespace glm
