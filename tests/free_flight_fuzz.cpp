// Host fuzz of raytracing_test_b200/csrc/free_flight.h (the free-flight tail of primary_kernel): free_flight() with the
// leap on must leave bit for bit what the plain nextIntersectDDA loop (src/shaders/low_res.frag:446-491) leaves -- the
// three deltaPos, the steps per axis, the axis of the last step.  Built and run by tests/test_free_flight_cpu.py:
//   g++ -O2 -ffp-contract=off -fno-fast-math -I raytracing_test_b200/csrc tests/free_flight_fuzz.cpp -o fuzz && ./fuzz N seed
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "free_flight.h"

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t rnd() {  // splitmix64
    uint64_t z = (rng_state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static float unit() { return (float)(rnd() >> 40) * (1.0f / 16777216.0f); }  // [0, 1)
static uint32_t below(uint32_t n) { return (uint32_t)(rnd() % n); }

struct Ray {
    float dx, dy, dz, adx, ady, adz, tx, ty, tz;
};

// buildRay + the deltaPos set-up of low_res.frag:61-97,290-296, in binary32
static Ray make_ray(float vx, float vy, float vz, float ox, float oy, float oz, bool normalise) {
    Ray r;
    if (normalise) {
        const float d = (vx * vx + vy * vy) + vz * vz;
        const float s = 1.0f / sqrtf(d);
        vx *= s, vy *= s, vz *= s;
    }
    r.dx = vx, r.dy = vy, r.dz = vz;
    const float ix = 1.0f / vx, iy = 1.0f / vy, iz = 1.0f / vz;
    r.adx = fabsf(ix), r.ady = fabsf(iy), r.adz = fabsf(iz);
    const float ex = vx < 0 ? ox - 1.0f : ox, ey = vy < 0 ? oy - 1.0f : oy, ez = vz < 0 ? oz - 1.0f : oz;
    r.tx = r.adx - (ex - floorf(ox)) * ix;
    r.ty = r.ady - (ey - floorf(oy)) * iy;
    r.tz = r.adz - (ez - floorf(oz)) * iz;
    return r;
}

static void walk(float& tx, float& ty, float& tz, const Ray& r, uint32_t n, int* cnt, int& last) {
    for (uint32_t k = 0; k < n; k++) {
        if (tx < ty && tx < tz) {
            tx = tx + r.adx, cnt[0]++, last = 0;
        } else if (ty < tz) {
            ty = ty + r.ady, cnt[1]++, last = 1;
        } else {
            tz = tz + r.adz, cnt[2]++, last = 2;
        }
    }
}

static float pick_component(int kind) {
    switch (kind) {
        case 0: return unit() * 2.0f - 1.0f;
        case 1: return ldexpf(1.0f, -(int)below(12)) * (below(2) ? 1.0f : -1.0f);              // +-2^-k: absDelta a power of two
        case 2: return 0.0f;                                                                   // axis-parallel: absDelta = inf
        case 3: return (unit() * 2.0f - 1.0f) * 1.0e-4f;                                       // nearly parallel: huge absDelta
        case 4: return (float)(1 + below(64)) / 64.0f * (below(2) ? 1.0f : -1.0f);             // short mantissas: many ties
        default: return ldexpf((float)(2 * below(2048) + 1), -12) * (below(2) ? 1.0f : -1.0f); // half-way increments
    }
}

int main(int argc, char** argv) {
    const long n = argc > 1 ? atol(argv[1]) : 1000000;
    if (argc > 2) rng_state ^= (uint64_t)atol(argv[2]) * 0xD1342543DE82EF95ull;
    long leaps = 0, bad = 0;
    for (long it = 0; it < n; it++) {
        const int mode = (int)below(8);
        float v[3], o[3];
        for (int k = 0; k < 3; k++) {
            v[k] = pick_component(mode < 4 ? 0 : (int)below(6));
            const int ok = (int)below(4);
            o[k] = (float)below(900) + (ok == 0 ? 0.0f : (ok == 1 ? 0.5f : unit()));
        }
        if (v[0] == 0.0f && v[1] == 0.0f && v[2] == 0.0f) v[below(3)] = 1.0f;
        const bool normalise = mode != 7;  // mode 7: the components as they are (|dir| up to sqrt 3, absDelta exact powers of two)
        Ray r = make_ray(v[0], v[1], v[2], o[0], o[1], o[2], normalise);
        // some steps of the ordinary walk first (the kernel enters free flight after a few dozen steps), then the tail
        const uint32_t budget = below(4) == 0 ? 1024u : (below(2) ? 300u : 1u + below(1024));
        const uint32_t pre = below(budget);
        const uint32_t total = budget - pre;
        float tx = r.tx, ty = r.ty, tz = r.tz;
        int c0[3] = {0, 0, 0}, l0 = -1;
        walk(tx, ty, tz, r, pre, c0, l0);
        float wx = tx, wy = ty, wz = tz;
        int cw[3] = {0, 0, 0}, lw = -1;
        walk(wx, wy, wz, r, total, cw, lw);
        float fx = tx, fy = ty, fz = tz;
        int nx = -1, ny = -1, nz = -1;
        bool lastX = false, lastY = false;
        rc::free_flight(fx, fy, fz, r.adx, r.ady, r.adz, r.dx, r.dy, r.dz, total, true, nx, ny, nz, lastX, lastY);
        const int lf = lastX ? 0 : (lastY ? 1 : 2);
        // did the leap run? (walk-only gives the same answer by construction; count to make sure the fuzz exercises it)
        {
            float gx = tx, gy = ty, gz = tz;
            uint32_t fired = 0;
            const float T0 = fminf(fminf(tx, ty), tz), rate = (fabsf(r.dx) + fabsf(r.dy)) + fabsf(r.dz);
            const float T = T0 + (float)(total - 7u) / rate;
            if (total >= 24u && rate >= 0.5f && T < 2048.0f && ty == ty && tz == tz && rc::leap_axis(gx, r.adx, T, fired) &&
                rc::leap_axis(gy, r.ady, T, fired) && rc::leap_axis(gz, r.adz, T, fired) && fired < total)
                leaps += fired > 0;
        }
        const bool same_t = memcmp(&fx, &wx, 4) == 0 && memcmp(&fy, &wy, 4) == 0 && memcmp(&fz, &wz, 4) == 0;
        // the counts come from the growth of deltaPos: promised for the kernel's rays (normalised directions)
        const bool same_n = !normalise || (nx == cw[0] && ny == cw[1] && nz == cw[2]);
        if (!same_t || !same_n || lf != lw) {
            if (bad++ < 10)
                printf("MISMATCH it %ld mode %d: dir %.9g %.9g %.9g  t0 %.9g %.9g %.9g  ad %.9g %.9g %.9g total %u\n"
                       "  walk  t %.9g %.9g %.9g  n %d %d %d last %d\n  flight t %.9g %.9g %.9g  n %d %d %d last %d\n",
                       it, mode, r.dx, r.dy, r.dz, tx, ty, tz, r.adx, r.ady, r.adz, total, wx, wy, wz, cw[0], cw[1], cw[2], lw, fx, fy,
                       fz, nx, ny, nz, lf);
        }
    }
    printf("rays %ld leaps %ld mismatches %ld\n", n, leaps, bad);
    return bad ? 1 : 0;
}
