// oracle/shim/glm/glm.hpp -- TEST INFRASTRUCTURE ONLY.
// glm is an un-vendored third-party dependency of the reference (header-only,
// version unpinned: include path C:/lib/glm in .vscode/c_cpp_properties.json:8).
// This is a minimal restatement of the handful of glm types/functions that
// src/globals.{hpp,cpp} and src/ray_caster.{hpp,cpp} use, with glm's published
// pure-C++ semantics: normalize(v) = v * (1/sqrt(dot(v,v))), dot = (x*x+y*y)+z*z,
// trunc/abs component-wise, converting constructors by static_cast.
#pragma once
#include <cmath>

namespace glm {

template <typename T>
struct tvec3 {
    T x, y, z;
    tvec3() : x(0), y(0), z(0) {}
    tvec3(T a, T b, T c) : x(a), y(b), z(c) {}
    explicit tvec3(T s) : x(s), y(s), z(s) {}
    template <typename A, typename B, typename C>
    tvec3(A a, B b, C c) : x(static_cast<T>(a)), y(static_cast<T>(b)), z(static_cast<T>(c)) {}
    template <typename U>
    tvec3(const tvec3<U>& o) : x(static_cast<T>(o.x)), y(static_cast<T>(o.y)), z(static_cast<T>(o.z)) {}
    T& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    const T& operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};

typedef tvec3<float> vec3;
typedef tvec3<double> dvec3;
typedef tvec3<int> ivec3;

template <typename T> inline tvec3<T> operator+(const tvec3<T>& a, const tvec3<T>& b) { return tvec3<T>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename T> inline tvec3<T> operator-(const tvec3<T>& a, const tvec3<T>& b) { return tvec3<T>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename T> inline tvec3<T> operator*(const tvec3<T>& a, const tvec3<T>& b) { return tvec3<T>(a.x * b.x, a.y * b.y, a.z * b.z); }
template <typename T> inline tvec3<T> operator*(const tvec3<T>& a, T s) { return tvec3<T>(a.x * s, a.y * s, a.z * s); }
template <typename T> inline bool operator==(const tvec3<T>& a, const tvec3<T>& b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

inline float dot(const vec3& a, const vec3& b) {
    vec3 t = a * b;
    return t.x + t.y + t.z;
}
inline vec3 normalize(const vec3& v) { return v * (1.0f / std::sqrt(dot(v, v))); }
inline vec3 trunc(const vec3& v) { return vec3(std::trunc(v.x), std::trunc(v.y), std::trunc(v.z)); }
inline vec3 floor(const vec3& v) { return vec3(std::floor(v.x), std::floor(v.y), std::floor(v.z)); }
inline double abs(double v) { return std::fabs(v); }
inline float abs(float v) { return std::fabs(v); }
inline int abs(int v) { return v < 0 ? -v : v; }

}  // namespace glm
