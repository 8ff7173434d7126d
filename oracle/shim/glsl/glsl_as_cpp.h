// oracle/shim/glsl/glsl_as_cpp.h -- TEST INFRASTRUCTURE ONLY.
//
// Enough of GLSL 4.60 (+ GL_ARB_gpu_shader_int64) as C++20 types and functions
// for the reference's fragment shader src/shaders/low_res.frag to be compiled
// by g++ AS IT STANDS (oracle/Makefile pipes it through three textual rewrites,
// listed in shim/glsl/frag_to_cpp.sed, none of which touches an expression).
// The result (oracle/_ref/libref_shader.so) is the reference's own traversal
// running on the host cores; the restated oracle (oracle_render.cpp) is checked
// against it pixel for pixel in tests/test_ref_shader_cpu.py.
//
// What a GL driver decides and this header therefore has to pick (the same
// picks as the oracle; DESIGN.md "parity status" lists them as unpinned):
//   * every float operation is IEEE binary32, round-to-nearest-even, evaluated
//     in source order, never contracted to FMA (build flag -ffp-contract=off);
//   * normalize(v) = v * (1.0f / sqrt(dot(v,v))), dot summed left to right;
//   * sin / cos / exp go through hooks that default to libm and that the test
//     harness can point at the oracle's deterministic versions.
//
// GLSL semantics that differ from C++ and are handled here:
//   * an unsuffixed literal such as 0.94 is a FLOAT in GLSL.  `float` is
//     therefore re-typed (by a macro in glsl_begin.h) to the class F below,
//     whose operators narrow any arithmetic operand to binary32 BEFORE the
//     operation, so `x * 0.94` is a binary32 product as in the shader.  Real
//     doubles in the shader (dvec3 ratios*, SCALED_COLOR) stay doubles.
//   * globals are re-initialised for every fragment: the shader body is
//     included inside a struct, one instance per pixel.
#ifndef GLSL_AS_CPP_H
#define GLSL_AS_CPP_H

#include <cmath>
#include <concepts>
#include <cstdint>
#include <type_traits>

namespace glslcpp {

using uint = unsigned int;

template <class T>
concept Num = std::is_arithmetic_v<T>;

extern float (*sin_hook)(float);
extern float (*cos_hook)(float);
extern float (*exp_hook)(float);

// ---------------------------------------------------------------- float
struct F {
    float v;
    F() = default;
    template <Num T>
    constexpr F(T a) : v((float)a) {}
    template <Num T>
    explicit constexpr operator T() const { return (T)v; }
};

#define GLSL_F_BINOP(op)                                                      \
    inline F operator op(F a, F b) { F r; r.v = a.v op b.v; return r; }       \
    template <Num T> inline F operator op(F a, T b) { return a op F(b); }     \
    template <Num T> inline F operator op(T a, F b) { return F(a) op b; }
GLSL_F_BINOP(+)
GLSL_F_BINOP(-)
GLSL_F_BINOP(*)
GLSL_F_BINOP(/)
#undef GLSL_F_BINOP
#define GLSL_F_CMP(op)                                                        \
    inline bool operator op(F a, F b) { return a.v op b.v; }                  \
    template <Num T> inline bool operator op(F a, T b) { return a.v op F(b).v; } \
    template <Num T> inline bool operator op(T a, F b) { return F(a).v op b.v; }
GLSL_F_CMP(<)
GLSL_F_CMP(>)
GLSL_F_CMP(<=)
GLSL_F_CMP(>=)
GLSL_F_CMP(==)
GLSL_F_CMP(!=)
#undef GLSL_F_CMP
#define GLSL_F_ASSIGN(op)                                                     \
    inline F& operator op##=(F& a, F b) { a = a op b; return a; }             \
    template <Num T> inline F& operator op##=(F& a, T b) { a = a op F(b); return a; }
GLSL_F_ASSIGN(+)
GLSL_F_ASSIGN(-)
GLSL_F_ASSIGN(*)
GLSL_F_ASSIGN(/)
#undef GLSL_F_ASSIGN
inline F operator-(F a) { F r; r.v = -a.v; return r; }

// --------------------------------------------- component type arithmetic
// GLSL implicit promotion: int -> uint -> float -> double.
template <class T> struct rank_of;
template <> struct rank_of<int> { static constexpr int v = 0; };
template <> struct rank_of<uint> { static constexpr int v = 1; };
template <> struct rank_of<F> { static constexpr int v = 2; };
template <> struct rank_of<double> { static constexpr int v = 3; };
template <int R> struct type_of_rank;
template <> struct type_of_rank<0> { using type = int; };
template <> struct type_of_rank<1> { using type = uint; };
template <> struct type_of_rank<2> { using type = F; };
template <> struct type_of_rank<3> { using type = double; };
template <class A, class B>
using prom = typename type_of_rank<(rank_of<A>::v > rank_of<B>::v ? rank_of<A>::v : rank_of<B>::v)>::type;
// a scalar operand written in the shader: integer literals / variables are int, anything floating is a GLSL float
template <Num T>
using lit = std::conditional_t<std::is_integral_v<T>, std::conditional_t<std::is_signed_v<T>, int, uint>, F>;

template <class To, class From>
inline To cv(From a) {
    if constexpr (std::is_same_v<To, From>) return a;
    else if constexpr (std::is_same_v<From, F>) return (To)a.v;       // float -> int truncates toward zero
    else if constexpr (std::is_same_v<To, F>) return F(a);
    else return (To)a;
}

template <class S> struct comp { using type = S; };
template <class S> using comp_t = typename comp<S>::type;

// ---------------------------------------------------------------- vectors
template <class S>
struct tvec2 {
    S x, y;
    tvec2() = default;
    template <class A> requires(Num<A> || std::is_same_v<A, F>)
    explicit tvec2(A a) : x(cv<S>(lit_or_f(a))), y(x) {}
    template <class A, class B>
    tvec2(A a, B b) : x(cv<S>(lit_or_f(a))), y(cv<S>(lit_or_f(b))) {}
    template <class O>
    tvec2(const tvec2<O>& o) : x(cv<S>(o.x)), y(cv<S>(o.y)) {}
    S& operator[](uint i) { return (&x)[i]; }
    const S& operator[](uint i) const { return (&x)[i]; }

  private:
    template <Num A> static lit<A> lit_or_f(A a) { return (lit<A>)a; }
    static F lit_or_f(F a) { return a; }
};

template <class S>
struct tvec3 {
    S x, y, z;
    tvec3() = default;
    template <class A> requires(Num<A> || std::is_same_v<A, F>)
    explicit tvec3(A a) : x(cv<S>(lit_or_f(a))), y(x), z(x) {}
    template <class A, class B, class C>
    tvec3(A a, B b, C c) : x(cv<S>(lit_or_f(a))), y(cv<S>(lit_or_f(b))), z(cv<S>(lit_or_f(c))) {}
    template <class O>
    tvec3(const tvec3<O>& o) : x(cv<S>(o.x)), y(cv<S>(o.y)), z(cv<S>(o.z)) {}
    S& operator[](uint i) { return (&x)[i]; }
    const S& operator[](uint i) const { return (&x)[i]; }

  private:
    template <Num A> static auto lit_or_f(A a) {
        if constexpr (std::is_same_v<A, double> && std::is_same_v<S, double>) return a;  // dvec3(1, f/f, f/f)
        else return (lit<A>)a;
    }
    static F lit_or_f(F a) { return a; }
};

using vec2 = tvec2<F>;
using uvec2 = tvec2<uint>;
using vec3 = tvec3<F>;
using ivec3 = tvec3<int>;
using dvec3 = tvec3<double>;

struct u64vec2 {
    uint64_t x, y;
    u64vec2() = default;
    explicit u64vec2(long long a) : x((uint64_t)a), y((uint64_t)a) {}
    u64vec2(long long a, long long b) : x((uint64_t)a), y((uint64_t)b) {}
};

// Only what the shader uses of a vec4: whole-value construction / assignment (FragColor) and the .xy swizzle
// (gl_FragCoord.xy); an anonymous {x,y,z,w} struct cannot hold class-typed members, so x and y live in xy.
struct vec4 {
    vec2 xy;
    F z, w;
    vec4() = default;
    template <class A, class B, class C, class D>
    vec4(A a, B b, C c, D d) : xy(F(a), F(b)), z(F(c)), w(F(d)) {}
    template <class S, class D>
    vec4(const tvec3<S>& v, D d) : xy(cv<F>(v.x), cv<F>(v.y)), z(cv<F>(v.z)), w(F(d)) {}
};

// vector (op) vector, component-wise with promotion
#define GLSL_V_BINOP(op)                                                                              \
    template <class A, class B>                                                                       \
    inline tvec3<prom<A, B>> operator op(const tvec3<A>& a, const tvec3<B>& b) {                      \
        using R = prom<A, B>;                                                                         \
        return tvec3<R>(cv<R>(a.x) op cv<R>(b.x), cv<R>(a.y) op cv<R>(b.y), cv<R>(a.z) op cv<R>(b.z)); \
    }                                                                                                 \
    template <class A, class B>                                                                       \
    inline tvec2<prom<A, B>> operator op(const tvec2<A>& a, const tvec2<B>& b) {                      \
        using R = prom<A, B>;                                                                         \
        return tvec2<R>(cv<R>(a.x) op cv<R>(b.x), cv<R>(a.y) op cv<R>(b.y));                          \
    }                                                                                                 \
    /* vector (op) scalar: F scalar */                                                                \
    template <class A>                                                                                \
    inline tvec3<prom<A, F>> operator op(const tvec3<A>& a, F s) {                                    \
        using R = prom<A, F>;                                                                         \
        return tvec3<R>(cv<R>(a.x) op cv<R>(s), cv<R>(a.y) op cv<R>(s), cv<R>(a.z) op cv<R>(s));      \
    }                                                                                                 \
    template <class A>                                                                                \
    inline tvec3<prom<A, F>> operator op(F s, const tvec3<A>& a) {                                    \
        using R = prom<A, F>;                                                                         \
        return tvec3<R>(cv<R>(s) op cv<R>(a.x), cv<R>(s) op cv<R>(a.y), cv<R>(s) op cv<R>(a.z));      \
    }                                                                                                 \
    /* vector (op) literal / plain arithmetic scalar */                                               \
    template <class A, Num T>                                                                         \
    inline tvec3<prom<A, lit<T>>> operator op(const tvec3<A>& a, T s) {                               \
        using R = prom<A, lit<T>>;                                                                    \
        const R r = cv<R>((lit<T>)s);                                                                 \
        return tvec3<R>(cv<R>(a.x) op r, cv<R>(a.y) op r, cv<R>(a.z) op r);                           \
    }                                                                                                 \
    template <class A, Num T>                                                                         \
    inline tvec3<prom<A, lit<T>>> operator op(T s, const tvec3<A>& a) {                               \
        using R = prom<A, lit<T>>;                                                                    \
        const R r = cv<R>((lit<T>)s);                                                                 \
        return tvec3<R>(r op cv<R>(a.x), r op cv<R>(a.y), r op cv<R>(a.z));                           \
    }                                                                                                 \
    template <class A, Num T>                                                                         \
    inline tvec2<prom<A, lit<T>>> operator op(T s, const tvec2<A>& a) {                               \
        using R = prom<A, lit<T>>;                                                                    \
        const R r = cv<R>((lit<T>)s);                                                                 \
        return tvec2<R>(r op cv<R>(a.x), r op cv<R>(a.y));                                            \
    }                                                                                                 \
    template <class A, class B>                                                                       \
    inline tvec3<A>& operator op##=(tvec3<A>& a, const B& b) { a = tvec3<A>(a op b); return a; }
GLSL_V_BINOP(+)
GLSL_V_BINOP(-)
GLSL_V_BINOP(*)
GLSL_V_BINOP(/)
#undef GLSL_V_BINOP

template <class A>
inline tvec3<A> operator-(const tvec3<A>& a) { return tvec3<A>(-a.x, -a.y, -a.z); }
// GLSL == on vectors: one bool, true when every component matches
template <class A>
inline bool operator==(const tvec3<A>& a, const tvec3<A>& b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

// ---------------------------------------------------------------- matrices (column major)
struct mat3 {
    vec3 c[3];
    mat3() = default;
    mat3(const vec3& a, const vec3& b, const vec3& d) : c{a, b, d} {}
    vec3& operator[](uint i) { return c[i]; }
    const vec3& operator[](uint i) const { return c[i]; }
};
struct mat4 {
    F m[16];
    mat4() = default;
    template <class... T> requires(sizeof...(T) == 16)
    mat4(T... a) : m{F(a)...} {}
};
struct sampler2D {};

// ---------------------------------------------------------------- built-ins
inline F abs(F a) { F r; r.v = std::fabs(a.v); return r; }
inline F floor(F a) { F r; r.v = std::floor(a.v); return r; }
inline F sqrt(F a) { F r; r.v = std::sqrt(a.v); return r; }  // IEEE correctly rounded
inline F sin(F a) { F r; r.v = sin_hook(a.v); return r; }
inline F cos(F a) { F r; r.v = cos_hook(a.v); return r; }
inline F exp(F a) { F r; r.v = exp_hook(a.v); return r; }
template <Num T> inline F sqrt(T a) { return sqrt(F(a)); }

// min(x,y) = y < x ? y : x ; max(x,y) = x < y ? y : x   (GLSL 4.60 spec 8.3)
inline int min(int a, int b) { return b < a ? b : a; }
inline int max(int a, int b) { return a < b ? b : a; }
inline F min(F a, F b) { return b < a ? b : a; }
inline F max(F a, F b) { return a < b ? b : a; }
template <Num T> inline F min(F a, T b) { return min(a, F(b)); }
template <Num T> inline F min(T a, F b) { return min(F(a), b); }
template <Num T> inline F max(F a, T b) { return max(a, F(b)); }
template <Num T> inline F max(T a, F b) { return max(F(a), b); }
template <class A, class B> requires(Num<A> || std::is_same_v<A, F>) && (Num<B> || std::is_same_v<B, F>)
inline F clamp(F x, A lo, B hi) { return min(max(x, F(lo)), F(hi)); }
inline ivec3 min(const ivec3& a, int b) { return ivec3(min(a.x, b), min(a.y, b), min(a.z, b)); }

inline vec3 abs(const vec3& a) { return vec3(abs(a.x), abs(a.y), abs(a.z)); }
inline vec3 floor(const vec3& a) { return vec3(floor(a.x), floor(a.y), floor(a.z)); }
inline F dot(const vec3& a, const vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline F dot(const vec2& a, const vec2& b) { return a.x * b.x + a.y * b.y; }
inline vec3 cross(const vec3& a, const vec3& b) {
    return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);  // GLSL 4.60 spec 8.5
}
inline F length(const vec3& a) { return sqrt(dot(a, a)); }
inline F distance(const vec3& a, const vec3& b) { const vec3 d = a - b; return sqrt(dot(d, d)); }
inline F distance(const vec2& a, const vec2& b) { const vec2 d = a - b; return sqrt(dot(d, d)); }
inline vec3 normalize(const vec3& a) { return a * (1.0f / sqrt(dot(a, a))); }

}  // namespace glslcpp

#endif
