// cuda_on_host.h -- TEST INFRASTRUCTURE.  Lets g++ compile the device code of raytracing_test_b200/csrc/kernels.cu for the
// host, so that the CUDA source itself -- not a restatement of it -- can be held to the oracle without a GPU
// (tests/test_kernel_source_cpu.py).  Thread / block indices are plain globals the harness sets before it calls a __global__
// function; every IEEE intrinsic is the corresponding C operation (compiled with -ffp-contract=off -fno-fast-math: one
// rounded operation each, as -fmad=false gives on the device; denormals are not flushed, the AO window code adds them like
// integers); atomics are plain read-modify-writes (one host thread).  A warp is 32 coroutines (KhWarp, ucontext): every
// *_sync intrinsic is a rendezvous at which the lanes exchange their arguments, and the emulator aborts if the lanes of a
// warp ever meet in different intrinsics.  Outside KhWarp::run a "warp" is a single lane (per-pixel harness).
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
static inline void __syncthreads() {}
static inline void __trap() { abort(); }
// shared memory of the emulated CTA: addresses are byte offsets into this array (the AO window code keeps them in 32 bits)
extern unsigned char* kh_shared_base;
static inline size_t __cvta_generic_to_shared(const void* p) { return (size_t)(static_cast<const unsigned char*>(p) - kh_shared_base); }
static inline unsigned kh_lds_u8(unsigned addr) { return kh_shared_base[addr]; }

// ---- CUDA runtime calls of upload_hemisphere(): plain memory on the host -------------------------------------------------------
template <typename T>
static inline cudaError_t kh_memcpy_to_symbol(T& symbol, const void* src, size_t count, size_t offset, cudaMemcpyKind, cudaStream_t) {
    memcpy(reinterpret_cast<char*>(&symbol) + offset, src, count);
    return cudaSuccess;
}
static inline cudaError_t kh_malloc(void** p, size_t n) {
    *p = malloc(n);
    return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}
static inline cudaError_t kh_memcpy(void* dst, const void* src, size_t n, cudaMemcpyKind, cudaStream_t) {
    memcpy(dst, src, n);
    return cudaSuccess;
}
static inline cudaError_t kh_stream_sync(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t kh_free(void* p) {
    free(p);
    return cudaSuccess;
}
#define cudaMemcpyToSymbolAsync kh_memcpy_to_symbol
#define cudaMalloc kh_malloc
#define cudaMemcpyAsync kh_memcpy
#define cudaStreamSynchronize kh_stream_sync
#define cudaFree kh_free

// ---- a warp: 32 lanes as coroutines on one host thread ---------------------------------------------------------------------------
// Every *_sync intrinsic is a rendezvous: the lane leaves its argument in the warp's exchange area and yields; when all
// lanes that are still running have arrived, each is resumed and reads what it needs.  Two exchange areas alternate (a lane
// can be at most one rendezvous ahead of the slowest).  Outside KhWarp::run (the per-pixel harness) a "warp" is one lane.
#include <stdio.h>
#include <ucontext.h>

struct KhWarp {
    static constexpr int N = 32;
    static constexpr size_t STACK = 512 * 1024;
    ucontext_t sched, ctx[N];
    char* stacks = nullptr;
    bool done[N], waiting[N];
    unsigned long long arg[2][N];
    unsigned part[2];       // lanes that took part in the rendezvous of each exchange area (a lane that has finished
                            // since must still count for the lanes that read their result later)
    int op[N];              // which intrinsic each waiting lane is in: a rendezvous of different intrinsics is a bug
    unsigned long count[N];  // rendezvous each lane has taken part in
    unsigned gen = 0;       // rendezvous being assembled
    int cur = -1;           // lane running now
    unsigned threadBase = 0;
    int order = 0;          // lane order between rendezvous (see run)
    unsigned pass = 0;
    void (*body)(void*) = nullptr;
    void* bodyArg = nullptr;

    static KhWarp*& active() {
        static KhWarp* w = nullptr;
        return w;
    }
    static void trampoline() {
        KhWarp* w = active();
        w->body(w->bodyArg);
        w->done[w->cur] = true;
        swapcontext(&w->ctx[w->cur], &w->sched);
    }
    // Runs body(arg) on 32 lanes; threadIdx.x = threadBase + lane while a lane runs.
    void run(unsigned base, void (*fn)(void*), void* a) {
        if (!stacks) stacks = static_cast<char*>(malloc(STACK * N));
        body = fn;
        bodyArg = a;
        threadBase = base;
        gen = 0;
        active() = this;
        for (int l = 0; l < N; l++) {
            done[l] = waiting[l] = false;
            op[l] = 0;
            count[l] = 0;
            getcontext(&ctx[l]);
            ctx[l].uc_stack.ss_sp = stacks + STACK * l;
            ctx[l].uc_stack.ss_size = STACK;
            ctx[l].uc_link = &sched;
            makecontext(&ctx[l], trampoline, 0);
        }
        for (;;) {
            bool any = false;
            pass++;
            for (int i = 0; i < N; i++) {
                // the order in which the lanes run between two rendezvous: 0 ascending, 1 descending, 2 shuffled anew in every
                // pass.  Results must not depend on it: whatever one lane writes for another to read has to be separated
                // from the read by a rendezvous (__syncwarp or any *_sync), or be an atomic whose order does not matter
                const int l = order == 0 ? i : (order == 1 ? N - 1 - i : (int)((i * 13u + pass * 7u + (pass >> 3) * 5u) & 31u));
                if (done[l] || waiting[l]) continue;
                any = true;
                cur = l;
                threadIdx.x = threadBase + (unsigned)l;
                swapcontext(&sched, &ctx[l]);  // until it finishes or arrives at a rendezvous
            }
            bool left = false;
            for (int l = 0; l < N; l++) left |= !done[l];
            if (!left) break;
            if (!any) {  // everyone still running has arrived: release them
                int first = -1;
                for (int l = 0; l < N; l++) {
                    if (done[l]) continue;
                    if (first < 0) first = l;
                    if (op[l] != op[first] || count[l] != count[first]) {
                        fprintf(stderr, "KhWarp: lanes %d and %d met in different intrinsics (%d #%lu vs %d #%lu): divergent *_sync\n", first, l,
                                op[first], count[first], op[l], count[l]);
                        abort();
                    }
                }
                part[gen & 1] = participants();
                gen++;
                for (int l = 0; l < N; l++) waiting[l] = false;
            }
        }
        cur = -1;
        active() = nullptr;
    }
    // lane side: contribute v, wait for the others; returns the exchange area of this rendezvous
    const unsigned long long* meet(unsigned long long v, int which) {
        const unsigned g = gen;
        arg[g & 1][cur] = v;
        op[cur] = which;
        count[cur]++;
        waiting[cur] = true;
        const int me = cur;
        swapcontext(&ctx[me], &sched);
        cur = me;
        met = part[g & 1];
        return arg[g & 1];
    }
    unsigned met = 0;  // participants of the rendezvous the running lane has just left
    unsigned participants() const {
        unsigned m = 0;
        for (int l = 0; l < N; l++)
            if (!done[l]) m |= 1u << l;
        return m;
    }
};

template <typename T>
static inline unsigned long long kh_bits(T v) {
    unsigned long long b = 0;
    memcpy(&b, &v, sizeof(T));
    return b;
}
template <typename T>
static inline T kh_from_bits(unsigned long long b) {
    T v;
    memcpy(&v, &b, sizeof(T));
    return v;
}
static inline int kh_lane() { return KhWarp::active() ? KhWarp::active()->cur : 0; }

static inline unsigned __activemask() { return 1u << kh_lane(); }  // converged lanes are not modelled: every lane acts alone
static inline unsigned __ballot_sync(unsigned, int pred) {
    KhWarp* w = KhWarp::active();
    if (!w) return pred ? 1u : 0u;
    const unsigned long long* a = w->meet(pred ? 1ull : 0ull, 1);
    unsigned m = 0, part = w->met;
    for (int l = 0; l < 32; l++)
        if (((part >> l) & 1u) && a[l]) m |= 1u << l;
    return m;
}
static inline int __any_sync(unsigned mask, int pred) { return __ballot_sync(mask, pred) != 0u; }
static inline int __all_sync(unsigned mask, int pred) {
    KhWarp* w = KhWarp::active();
    if (!w) return pred != 0;
    return __ballot_sync(mask, !pred) == 0u;
}
template <typename T>
static inline T __shfl_sync(unsigned, T v, int src, int = 32) {
    KhWarp* w = KhWarp::active();
    if (!w) return v;
    const unsigned long long* a = w->meet(kh_bits(v), 2);
    return kh_from_bits<T>(a[src & 31]);
}
template <typename T>
static inline T __shfl_xor_sync(unsigned, T v, int laneMask, int = 32) {
    KhWarp* w = KhWarp::active();
    if (!w) return (T)0 * v;  // the other lanes of the butterfly do not exist: they contribute nothing to a sum
    const int me = w->cur;
    const unsigned long long* a = w->meet(kh_bits(v), 3);
    return kh_from_bits<T>(a[(me ^ laneMask) & 31]);
}
template <typename T>
static inline unsigned __match_any_sync(unsigned, T v) {
    KhWarp* w = KhWarp::active();
    if (!w) return 1u;
    const int me = w->cur;
    const unsigned long long* a = w->meet(kh_bits(v), 4);
    unsigned m = 0, part = w->met;
    for (int l = 0; l < 32; l++)
        if (((part >> l) & 1u) && a[l] == a[me]) m |= 1u << l;
    return m;
}
static inline unsigned kh_reduce(unsigned v, int op) {
    KhWarp* w = KhWarp::active();
    if (!w) return v;
    const unsigned long long* a = w->meet(v, 5 + op);
    unsigned r = op == 1 ? 0xFFFFFFFFu : 0u, part = w->met;
    for (int l = 0; l < 32; l++) {
        if (!((part >> l) & 1u)) continue;
        const unsigned x = (unsigned)a[l];
        r = op == 0 ? (r | x) : (op == 1 ? (r & x) : r + x);
    }
    return r;
}
static inline unsigned __reduce_or_sync(unsigned, unsigned v) { return kh_reduce(v, 0); }
static inline unsigned __reduce_and_sync(unsigned, unsigned v) { return kh_reduce(v, 1); }
static inline unsigned __reduce_add_sync(unsigned, unsigned v) { return kh_reduce(v, 2); }
static inline void __syncwarp(unsigned = 0xffffffffu) {
    KhWarp* w = KhWarp::active();
    if (w) w->meet(0ull, 9);
}
