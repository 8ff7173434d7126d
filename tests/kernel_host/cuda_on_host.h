// cuda_on_host.h -- TEST INFRASTRUCTURE.  Lets g++ compile the device code of raytracing_test_b200/csrc/kernels.cu for the
// host, so that the CUDA source itself -- not a restatement of it -- can be held to the oracle without a GPU
// (tests/test_kernel_source_cpu.py).  One host thread plays one CUDA thread: thread / block indices are plain globals the
// harness sets before it calls a __global__ function, every IEEE intrinsic is the corresponding C operation (this file is
// compiled with -ffp-contract=off -fno-fast-math: one rounded operation each, as -fmad=false gives on the device), atomics
// are plain read-modify-writes.  Warp-collective intrinsics are declared so that the warp-cooperative AO code compiles, but
// they only make sense for a "warp" of one lane: the harness runs the per-thread code (shade_pixel: primary rays, free
// flight, reflection / refraction, shading, shadow rays; the table kernels) and never the AO window code.
#pragma once
#include <cuda_runtime.h>  // vector types, make_*; no CUDA call is made
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <string>  // before __noinline__ below: libstdc++ spells attributes with that token
#include <vector>

#ifndef __launch_bounds__
#define __launch_bounds__(...)
#endif
// nvcc's keyword; undefined again by the harness right after the kernel source (libstdc++ uses the token in attributes)
#define __noinline__ __attribute__((noinline))

struct KhIdx {
    unsigned x = 0, y = 0, z = 0;
};
struct KhDim {
    unsigned x = 1, y = 1, z = 1;
};
static KhIdx threadIdx, blockIdx;
static KhDim blockDim, gridDim;

// ---- one IEEE operation each ---------------------------------------------------------------------------------------------
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
static inline float __fdividef(float a, float b) { return a / b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __dsub_rn(double a, double b) { return a - b; }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline float __int_as_float(int v) {
    float f;
    memcpy(&f, &v, 4);
    return f;
}
static inline float __uint_as_float(unsigned v) {
    float f;
    memcpy(&f, &v, 4);
    return f;
}
static inline unsigned __float_as_uint(float f) {
    unsigned v;
    memcpy(&v, &f, 4);
    return v;
}
static inline int __float_as_int(float f) {
    int v;
    memcpy(&v, &f, 4);
    return v;
}
static inline int __float2int_rd(float f) { return (int)floorf(f); }
static inline unsigned __float2uint_rd(float f) { return (unsigned)floorf(f); }

// ---- integer helpers -------------------------------------------------------------------------------------------------------
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline int __popcll(unsigned long long v) { return __builtin_popcountll(v); }
static inline int __ffs(int v) { return __builtin_ffs(v); }
static inline int __ffsll(long long v) { return __builtin_ffsll(v); }
static inline unsigned __brev(unsigned v) {
    unsigned r = 0;
    for (int i = 0; i < 32; i++) r |= ((v >> i) & 1u) << (31 - i);
    return r;
}
static inline unsigned __fns(unsigned mask, unsigned base, int offset) {
    int seen = 0;
    for (unsigned i = base; i < 32; i++)
        if ((mask >> i) & 1u)
            if (++seen == offset) return i;
    return 0xFFFFFFFFu;
}
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline unsigned min(unsigned a, unsigned b) { return a < b ? a : b; }
static inline unsigned max(unsigned a, unsigned b) { return a > b ? a : b; }
static inline long long clock64() { return 0; }

// ---- memory ----------------------------------------------------------------------------------------------------------------
template <typename T>
static inline T __ldg(const T* p) {
    return *p;
}
template <typename T>
static inline T atomicAdd(T* p, T v) {
    T old = *p;
    *p = old + v;
    return old;
}
static inline unsigned atomicOr(unsigned* p, unsigned v) {
    unsigned old = *p;
    *p = old | v;
    return old;
}
static inline int atomicMax(int* p, int v) {
    int old = *p;
    if (v > old) *p = v;
    return old;
}
static inline unsigned atomicMax_system(unsigned* p, unsigned v) {
    unsigned old = *p;
    if (v > old) *p = v;
    return old;
}
static inline void __threadfence_system() {}
static inline void __nanosleep(unsigned) {}
static inline void __syncwarp(unsigned = 0xffffffffu) {}
static inline void __syncthreads() {}
static inline void __trap() { abort(); }
static inline size_t __cvta_generic_to_shared(const void* p) { return (size_t)p; }

// ---- a warp of ONE lane ------------------------------------------------------------------------------------------------------
static inline unsigned __activemask() { return 1u; }
static inline unsigned __ballot_sync(unsigned, int pred) { return pred ? 1u : 0u; }
static inline int __any_sync(unsigned, int pred) { return pred != 0; }
static inline int __all_sync(unsigned, int pred) { return pred != 0; }
template <typename T>
static inline T __shfl_sync(unsigned, T v, int, int = 32) {
    return v;
}
template <typename T>
static inline T __shfl_xor_sync(unsigned, T v, int, int = 32) {
    return (T)0 * v;  // the other lanes of the butterfly do not exist: they contribute nothing to a sum
}
template <typename T>
static inline unsigned __match_any_sync(unsigned, T) {
    return 1u;
}
static inline unsigned __reduce_or_sync(unsigned, unsigned v) { return v; }
static inline unsigned __reduce_and_sync(unsigned, unsigned v) { return v; }
static inline unsigned __reduce_add_sync(unsigned, unsigned v) { return v; }
