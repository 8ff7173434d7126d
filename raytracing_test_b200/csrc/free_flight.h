// free_flight.h -- the tail of a primary ray that has nothing left to hit (kernels.cu, loop A).
//
// The remaining steps of nextIntersectDDA (src/shaders/low_res.frag:446-491) only move the ray: per step the smallest of
// the three deltaPos fires (x if tx < ty && tx < tz, else y if ty < tz, else z) and grows by its absDelta, one rounded
// fp32 addition.  What the frame needs from them is the voxel the ray ends in, the axis of its last step and (to carry
// on) the three deltaPos -- bit for bit what the step-by-step loop leaves behind.
//
// One source for the device (primary_kernel) and for the host: tests/free_flight_fuzz.cpp compiles this header with g++
// and holds free_flight() to the plain loop over tens of millions of adversarial rays (ties, powers of two, half-way
// increments, axis-parallel directions, binade crossings).  No fused multiply-add anywhere: every operation below is one
// IEEE binary32 operation (nvcc -fmad=false / g++ -ffp-contract=off).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define RC_FF_FN __host__ __device__ __forceinline__
#else
#include <math.h>
#include <string.h>
#define RC_FF_FN static inline
#endif

namespace rc {

#if defined(__CUDA_ARCH__)
RC_FF_FN float ff_add(float a, float b) { return __fadd_rn(a, b); }
RC_FF_FN float ff_sub(float a, float b) { return __fsub_rn(a, b); }
RC_FF_FN float ff_mul(float a, float b) { return __fmul_rn(a, b); }
RC_FF_FN float ff_div(float a, float b) { return __fdiv_rn(a, b); }
RC_FF_FN float ff_div_approx(float a, float b) { return __fdividef(a, b); }  // any quotient within a few ulps will do
RC_FF_FN uint32_t ff_bits(float f) { return __float_as_uint(f); }
RC_FF_FN float ff_float(uint32_t u) { return __uint_as_float(u); }
RC_FF_FN uint32_t ff_floor_u32(float f) { return __float2uint_rd(f); }
#else
RC_FF_FN float ff_add(float a, float b) { return a + b; }
RC_FF_FN float ff_sub(float a, float b) { return a - b; }
RC_FF_FN float ff_mul(float a, float b) { return a * b; }
RC_FF_FN float ff_div(float a, float b) { return a / b; }
RC_FF_FN float ff_div_approx(float a, float b) { return a / b; }
RC_FF_FN uint32_t ff_bits(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
}
RC_FF_FN float ff_float(uint32_t u) {
    float f;
    memcpy(&f, &u, 4);
    return f;
}
RC_FF_FN uint32_t ff_floor_u32(float f) { return (uint32_t)floorf(f); }
#endif

// Free flight in closed form.  With nothing to test per step, the DDA is a merge of three independent sequences of event
// times, t_a, fl(t_a + absDelta_a), fl(fl(t_a + absDelta_a) + absDelta_a), ...: it always fires the smallest, so its first
// N steps are exactly the events earlier than some time T, whatever the tie order, and "every event before T has fired"
// is a state the step-by-step loop passes through.  leap_axis fires those events for one axis without walking them.
// Inside one binade [2^e, 2^(e+1)) t = m * u (u = one ulp, m its 24-bit mantissa) and absDelta = q + r / 2^sh ulps; as
// long as the exact sum stays below 2^(e+1) the rounded addition moves m by the same integer every time, c = q or q + 1
// (round to nearest; a half-way case r = 2^(sh-1) goes to the even mantissa, which is the same c every time once m is
// even -- an odd m takes one real step first).  So the k additions that keep m below min(T / u, 2^24 - q) are
// m += k * c, in 32-bit integers; the addition that crosses into the next binade is the real one.  Returns false outside
// the range this covers (the caller walks instead).  n += events fired.
RC_FF_FN bool leap_axis(float& t, const float a, const float T, uint32_t& n) {
    if (!(t < T)) return true;  // nothing fires (t = inf / NaN: an axis that never steps)
    if (!(a < 4096.0f)) {       // T < 2048: one event, the next one far beyond T (absDelta = inf included)
        t = ff_add(t, a);
        n++;
        return true;
    }
    if (!(t >= 0.5f && a >= 0.5f)) return false;
    const uint32_t abits = ff_bits(a);
    const int ea = (int)(abits >> 23);
    const uint32_t ma = (abits & 0x7FFFFFu) | 0x800000u;
    while (t < T) {
        const uint32_t bits = ff_bits(t);
        const int e = (int)(bits >> 23);  // biased; 126 <= e <= 137
        const int sh = e - ea;            // <= 11
        if (sh >= 1) {                    // else absDelta >= 2^23 ulps: the sum leaves the binade at once
            const uint32_t m = (bits & 0x7FFFFFu) | 0x800000u;
            const uint32_t q = ma >> sh, r = ma & ((1u << sh) - 1u), half = 1u << (sh - 1);
            const bool tie = r == half;
            if (!(tie && (m & 1u))) {
                const uint32_t c = q + ((r > half || (tie && (q & 1u))) ? 1u : 0u);  // >= 2^12
                // T in ulps, rounded up (m * u < T <=> m < ceil(T / u)), at most the top of the binade
                const float tm = fminf(ceilf(ff_mul(T, ff_float((uint32_t)(277 - e) << 23))), 16777216.0f);
                const int bm = (int)tm < (int)(0x1000000u - q) ? (int)tm : (int)(0x1000000u - q);
                if (bm > (int)m) {
                    // k = ceil((bm - m) / c): the fp32 quotient (operands < 2^24: exact, k < 2^12) floors to k - 2 .. k
                    uint32_t k = ff_floor_u32(ff_div_approx((float)(bm - (int)m), (float)c));
                    uint32_t m1 = m + k * c;
                    while ((int)m1 < bm) {
                        m1 += c;
                        k++;
                    }
                    t = ff_mul((float)m1, ff_float((uint32_t)(e - 23) << 23));  // m1 <= 2^24, u = 2^(e - 127 - 23): exact
                    n += k;
                }
            }
        }
        if (t < T) {
            t = ff_add(t, a);
            n++;
        }
    }
    return true;
}

// `total` (>= 1) steps of the DDA from deltaPos (tx, ty, tz) with nothing to test.  Out: deltaPos after them, the steps
// taken along each axis, the axis of the last one (lastX, lastY; neither: z).  leap = false walks every step.
RC_FF_FN void free_flight(float& tx, float& ty, float& tz, const float adx, const float ady, const float adz, const float dirx,
                          const float diry, const float dirz, const uint32_t total, const bool leap, int& nx, int& ny, int& nz,
                          bool& lastX, bool& lastY) {
    const float tx0 = tx, ty0 = ty;
    uint32_t left = total;
    if (leap && left >= 24u) {
        // T = a time by which at most left - 4 events have fired: an axis fires at most (T - T0) / absDelta * (1 + 2^-13)
        // + 1 events from T0 on (its additions are off by at most half an ulp of t < 2048 each), and the three
        // 1 / absDelta add up to |dir.x| + |dir.y| + |dir.z|
        const float T0 = fminf(fminf(tx, ty), tz);
        const float rate = ff_add(ff_add(fabsf(dirx), fabsf(diry)), fabsf(dirz));
        const float T = ff_add(T0, ff_div((float)(left - 7u), rate));
        // a NaN deltaPos (0 * inf in buildRay) on y or z makes the else-branch axis fire every time: not a merge
        if (rate >= 0.5f && T < 2048.0f && ty == ty && tz == tz) {
            float lx = tx, ly = ty, lz = tz;
            uint32_t fired = 0;
            if (leap_axis(lx, adx, T, fired) && leap_axis(ly, ady, T, fired) && leap_axis(lz, adz, T, fired) && fired < left) {
                tx = lx;
                ty = ly;
                tz = lz;
                left -= fired;
            }
        }
    }
    // same comparisons, same additions as nextIntersectDDA; nothing else per step
#if defined(__CUDA_ARCH__)
#pragma unroll 4
#endif
    for (uint32_t k = 0; k < left; k++) {
        lastX = tx < ty && tx < tz;
        lastY = !lastX && ty < tz;
        const bool az = !lastX && !lastY;
        if (lastX) tx = ff_add(tx, adx);
        if (lastY) ty = ff_add(ty, ady);
        if (az) tz = ff_add(tz, adz);
    }
    // Steps per axis from the growth of deltaPos: n rounded additions of absDelta moved it by n * absDelta give or take n
    // half-ulps of its final value -- with deltaPos <= 1.74 * max_steps + absDelta, absDelta >= 1 and max_steps <= 1024 at
    // most 0.11 * absDelta, so the quotient rounds to n.  An axis that never steps (absDelta = inf, deltaPos inf or NaN)
    // has a difference of 0 or NaN.
    const float gx = ff_sub(tx, tx0), gy = ff_sub(ty, ty0);
    nx = gx > 0.0f ? (int)rintf(ff_div(gx, adx)) : 0;
    ny = gy > 0.0f ? (int)rintf(ff_div(gy, ady)) : 0;
    nz = (int)total - nx - ny;
}

}  // namespace rc
