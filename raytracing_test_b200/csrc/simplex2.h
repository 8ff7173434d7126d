// simplex2.h -- 2-D OpenSimplex noise, shared by the host stand-in (voxel_host.cpp, plain C++) and the
// device world generator (worldgen.cu): one source, so the doubles agree bit for bit on both sides
// (no FMA contraction on either: -ffp-contract=off / -fmad=false).
//
// This file IS on a product path (rc_generate_world): it restates KdotJPG's public 2-D OpenSimplex with the
// reference's constants and expression order (include/OpenSimplexNoise.cpp, lines cited below) because the generated
// world has to equal genWorld()'s voxel for voxel, which needs identical doubles (tests/test_host_cpu.py::
// test_simplex_noise_bit_exact).  It is a compact re-expression, not a copy: one lattice walk instead of the vendored
// file's unrolled case tree.
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define RC_HD __host__ __device__
#else
#define RC_HD
#endif

namespace rc_noise {

// ---------------------------------------------------------------------------
// 2-D OpenSimplex noise (KdotJPG's algorithm as vendored by the reference:
// include/OpenSimplexNoise.cpp:52-75 permutation, :77-208 eval, :2518-2523 gradient).
// The expression order below is the reference's, so doubles agree bit for bit.
// ---------------------------------------------------------------------------
struct Simplex2 {
    static constexpr double kStretch = -0.211324865405187;
    static constexpr double kSquish = 0.366025403784439;
    short perm[256];

    Simplex2() = default;
    explicit Simplex2(int64_t seedIn) {
        short source[256];
        for (int i = 0; i < 256; i++) source[i] = (short)i;
        uint64_t s = (uint64_t)seedIn;
        auto lcg = [&s]() { s = s * 6364136223846793005ull + 1442695040888963407ull; };
        lcg();
        lcg();
        lcg();
        for (int i = 255; i >= 0; i--) {
            lcg();
            int r = (int)((int64_t)(s + 31) % (i + 1));
            if (r < 0) r += i + 1;
            perm[i] = source[r];
            source[r] = source[i];
        }
    }

    RC_HD double gradient(int xsb, int ysb, double dx, double dy) const {
        const signed char g[16] = {5, 2, 2, 5, -5, 2, -2, 5, 5, -2, 2, -5, -5, -2, -2, -5};
        const int idx = perm[(perm[xsb & 0xFF] + ysb) & 0xFF] & 0x0E;
        return g[idx] * dx + g[idx + 1] * dy;
    }

    RC_HD double eval(double x, double y) const {
        const double stretch = (x + y) * kStretch;
        const double xs = x + stretch, ys = y + stretch;
        int xsb = (int)floor(xs), ysb = (int)floor(ys);
        const double squish = (xsb + ysb) * kSquish;
        const double xins = xs - xsb, yins = ys - ysb;
        const double inSum = xins + yins;
        double dx0 = x - (xsb + squish), dy0 = y - (ysb + squish);

        double value = 0;
        // vertices (1,0) and (0,1) always contribute
        {
            const double dx1 = dx0 - 1 - kSquish, dy1 = dy0 - 0 - kSquish;
            double a = 2 - dx1 * dx1 - dy1 * dy1;
            if (a > 0) {
                a *= a;
                value += a * a * gradient(xsb + 1, ysb + 0, dx1, dy1);
            }
            const double dx2 = dx0 - 0 - kSquish, dy2 = dy0 - 1 - kSquish;
            double b = 2 - dx2 * dx2 - dy2 * dy2;
            if (b > 0) {
                b *= b;
                value += b * b * gradient(xsb + 0, ysb + 1, dx2, dy2);
            }
        }
        // the fourth ("extra") vertex depends on which half of the rhombus we are in
        int xe, ye;
        double dxe, dye;
        if (inSum <= 1) {
            const double zins = 1 - inSum;
            if (zins > xins || zins > yins) {
                if (xins > yins) { xe = xsb + 1; ye = ysb - 1; dxe = dx0 - 1; dye = dy0 + 1; }
                else             { xe = xsb - 1; ye = ysb + 1; dxe = dx0 + 1; dye = dy0 - 1; }
            } else {
                xe = xsb + 1; ye = ysb + 1;
                dxe = dx0 - 1 - 2 * kSquish; dye = dy0 - 1 - 2 * kSquish;
            }
        } else {
            const double zins = 2 - inSum;
            if (zins < xins || zins < yins) {
                if (xins > yins) { xe = xsb + 2; ye = ysb + 0; dxe = dx0 - 2 - 2 * kSquish; dye = dy0 + 0 - 2 * kSquish; }
                else             { xe = xsb + 0; ye = ysb + 2; dxe = dx0 + 0 - 2 * kSquish; dye = dy0 - 2 - 2 * kSquish; }
            } else {
                xe = xsb; ye = ysb; dxe = dx0; dye = dy0;
            }
            xsb += 1;
            ysb += 1;
            dx0 = dx0 - 1 - 2 * kSquish;
            dy0 = dy0 - 1 - 2 * kSquish;
        }
        {
            double a = 2 - dx0 * dx0 - dy0 * dy0;
            if (a > 0) {
                a *= a;
                value += a * a * gradient(xsb, ysb, dx0, dy0);
            }
            double e = 2 - dxe * dxe - dye * dye;
            if (e > 0) {
                e *= e;
                value += e * e * gradient(xe, ye, dxe, dye);
            }
        }
        return value / 47;
    }
};


}  // namespace rc_noise
