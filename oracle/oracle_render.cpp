// oracle/oracle_render.cpp -- TEST INFRASTRUCTURE ONLY (see oracle_render.h).
//
// CPU restatement of the reference's per-pixel renderer, src/shaders/low_res.frag,
// in IEEE fp32 (compile with -O2 -ffp-contract=off -fno-fast-math: every +,-,*,/
// and sqrt below is one correctly-rounded fp32 operation, exactly as written).
// Each function cites the shader lines it follows.  GLSL built-ins are restated as:
//   normalize(v) = v * (1/sqrt(dot(v,v)))   (glm's pure-C++ definition)
//   dot(a,b)     = (a.x*b.x + a.y*b.y) + a.z*b.z
//   cross(a,b)   = (a.y*b.z - b.y*a.z, a.z*b.x - b.z*a.x, a.x*b.y - b.x*a.y)
//   distance     = sqrt(dot(d,d)),  clamp(x,a,b) = min(max(x,a),b)
//   sin / exp    = the deterministic polynomials oracle_sinf / oracle_expf below
//                  (specified in DESIGN.md so the CUDA side evaluates the same
//                  sequence of fp32 operations; both are checked against libm).
//
// Two deliberate, documented divergences from what a GL driver would execute:
//   (1) F11: a child index of 0 (the memset value of child arrays, and what the
//       reference's aliased root array holds for root cell 63) is treated as
//       "no child" => empty.  The shader would recurse into the root again and
//       eventually index stack[] out of bounds (undefined behaviour).
//   (2) 8-bit framebuffer conversion is pinned to floor(clamp(c,0,1)*255 + 0.5).
//
// AO (ao_samples > 0) has no working counterpart in the reference
// (light_scattering.frag is dead code); its semantics are defined HERE, from
// light_scattering.frag:175-236 and the shadow ray of low_res.frag:373-391
// (SURVEY 8 row A10), and the CUDA kernel is held to them.
#include "oracle_render.h"

#include <math.h>
#include <string.h>

#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

typedef uint64_t u64;
typedef uint32_t u32;

struct V3 {
    float x, y, z;
    float& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
struct I3 {
    int x, y, z;
    int& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    int operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};

inline V3 v3(float a, float b, float c) { return V3{a, b, c}; }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
inline V3 operator+(V3 a, float s) { return v3(a.x + s, a.y + s, a.z + s); }
inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
inline V3 normalize(V3 v) { return v * (1.0f / sqrtf(dot(v, v))); }
inline float distance3(V3 a, V3 b) {
    V3 d = a - b;
    return sqrtf(dot(d, d));
}
inline float minf(float a, float b) { return b < a ? b : a; }   // GLSL min: y < x ? y : x
inline float maxf(float a, float b) { return a < b ? b : a; }   // GLSL max: x < y ? y : x
inline float clampf(float x, float lo, float hi) { return minf(maxf(x, lo), hi); }
inline V3 toV3(I3 i) { return v3((float)i.x, (float)i.y, (float)i.z); }

// ---------------------------------------------------------------------------
// Deterministic sin / exp (specification in DESIGN.md; mirrored op-for-op on the
// device). Only +,-,* and floor on fp32, so every IEEE machine agrees bit-for-bit.
// ---------------------------------------------------------------------------
float detSin(float x) {
    if (!(fabsf(x) <= 1.0e6f)) return 0.0f;  // out of the range reduction's reach (never hit by the path)
    // x = k*2pi + r, 2pi split into three pieces with short mantissas (Cody-Waite)
    float k = floorf(x * 0.15915494309189535f + 0.5f);
    float r = x - k * 6.28125f;
    r = r - k * 1.9350051879882812e-3f;
    r = r - k * 3.0199160691085e-7f;
    // fold [-pi,pi] onto [-pi/2,pi/2]
    const float pi = 3.14159265358979f;
    const float halfPi = 1.57079632679490f;
    if (r > halfPi) r = pi - r;
    if (r < -halfPi) r = -pi - r;
    float s = r * r;
    float p = -2.5052108385441720e-08f;           // -1/11!
    p = p * s + 2.7557319223985893e-06f;          //  1/9!
    p = p * s + -1.9841269841269841e-04f;         // -1/7!
    p = p * s + 8.3333333333333332e-03f;          //  1/5!
    p = p * s + -1.6666666666666666e-01f;         // -1/3!
    return r + (r * s) * p;
}

float detExp(float x) {
    if (x != x) return x;
    if (x > 88.0f) x = 88.0f;
    if (x < -87.0f) x = -87.0f;
    float k = floorf(x * 1.44269504088896341f + 0.5f);
    float r = x - k * 0.693359375f;
    r = r - k * -2.12194440e-4f;
    float p = 1.9875691500e-4f;
    p = p * r + 1.3981999507e-3f;
    p = p * r + 8.3334519073e-3f;
    p = p * r + 4.1665795894e-2f;
    p = p * r + 1.6666665459e-1f;
    p = p * r + 5.0000001201e-1f;
    float e = (p * (r * r) + r) + 1.0f;
    int ki = (int)k;  // in [-126, 127]
    u32 bits = (u32)(ki + 127) << 23;
    float scale;
    memcpy(&scale, &bits, 4);
    return e * scale;
}

// ---------------------------------------------------------------------------
// Morton position (src/main.cpp:45-66, low_res.frag:454-459): 2 bits per axis per
// 6-bit group, group g = ((z>>2g)&3)<<4 | ((y>>2g)&3)<<2 | ((x>>2g)&3).
// Only the low 2*(max_depth-1) bits of each coordinate survive => torus.
// ---------------------------------------------------------------------------
inline u64 spreadAxis(int v, int groups) {
    u32 u = (u32)v;
    if (groups == 5) {
        // the reference's literal bit spread (low_res.frag:454-459, main.cpp:48-51). Its first mask
        // 0x030000FF also keeps bits 24-25 of the RAW value, so for a negative coordinate (two's
        // complement: those bits are 1) the top group reads 3 whatever bits 8-9 are: x < 0 maps to
        // 768 | (x & 255), not to x & 1023.  Non-negative coordinates below 2^24 wrap modulo 1024.
        u32 n = u;
        n = (n | (n << 16)) & 0x030000FFu;
        n = (n | (n << 8)) & 0x0300F00Fu;
        n = (n | (n << 4)) & 0x030C30C3u;
        return n;
    }
    // declared extension (max_depth != 6): plain torus of 2*(max_depth-1) bits per axis
    u64 n = 0;
    for (int g = 0; g < groups; g++) n |= (u64)((u >> (2 * g)) & 3u) << (6 * g);
    return n;
}
inline u64 axisMask(int groups) {
    u64 m = 0;
    for (int g = 0; g < groups; g++) m |= (u64)3u << (6 * g);
    return m;
}

// ---------------------------------------------------------------------------
// Hemisphere table, gen_hemisphare_distrib.py:4-12 (double precision, then fp32).
// Stored as (x, y, up): the script prints (x, z, y) and the shader swizzles .xzy
// (light_scattering.frag:230), which lands cos(phi) in the third slot.
// ---------------------------------------------------------------------------
std::vector<float> g_hemi;
void buildHemisphere(u32 n, float* out) {
    const double pi = 3.141592653589793;
    for (u32 i = 0; i < n; i++) {
        double idx = (double)i + 0.5;
        double phi = acos(1.0 - idx * 0.85 / (double)n);
        double theta = pi * (1.0 + sqrt(5.0)) * idx;
        out[3 * i + 0] = (float)(cos(theta) * sin(phi));
        out[3 * i + 1] = (float)(sin(theta) * sin(phi));
        out[3 * i + 2] = (float)cos(phi);
    }
}

// ---------------------------------------------------------------------------
// Per-pixel state = the shader's globals (low_res.frag:41-59,99-119,170-176,196,254)
// ---------------------------------------------------------------------------
struct Ray {
    V3 dir;
    I3 step;
    V3 ratios[3];  // mat3 columns: ratios[i] = dir / dir[i], with ratios[i][i] = 1
    V3 delta;
    V3 absDelta;
};

// low_res.frag:61-97
Ray buildRay(V3 dir) {
    Ray ray;
    ray.dir = dir;
    ray.step = I3{1, 1, 1};
    if (dir.x < 0) ray.step.x = -1;
    if (dir.y < 0) ray.step.y = -1;
    if (dir.z < 0) ray.step.z = -1;
    ray.ratios[0] = v3(1.0f, dir.y / dir.x, dir.z / dir.x);
    ray.ratios[1] = v3(dir.x / dir.y, 1.0f, dir.z / dir.y);
    ray.ratios[2] = v3(dir.x / dir.z, dir.y / dir.z, 1.0f);
    ray.delta = v3(1.0f / dir.x, 1.0f / dir.y, 1.0f / dir.z);
    ray.absDelta = v3(fabsf(ray.delta.x), fabsf(ray.delta.y), fabsf(ray.delta.z));
    return ray;
}

struct Pixel {
    const OracleWorld* w;
    const OracleFrame* f;
    int groups;           // max_depth-1 (5 for the reference)
    u64 amask;            // axisMask(groups)

    // Pos pos (low_res.frag:99-107)
    I3 round;
    V3 exact;
    V3 deltaPos;
    Ray ray;

    // tree walk state (low_res.frag:112-115)
    u32 stack[8];
    u64 currentMortonPos;
    int depth;
    u64 mortonPos;

    u32 lastHit;
    V3 finalColorMod;
    bool rayReflected;
    float currentRefractiveIndex;
    u64 blockX, blockY;   // u64vec2 block

    // bookkeeping (not in the shader)
    int phase;
    u64 nodeReads[3], arrayReads[3], ddaSteps[3];

    void setMortonAxis(int a) {
        mortonPos = (mortonPos & ~(amask << (2 * a))) | (spreadAxis(round[a], groups) << (2 * a));
    }

    // low_res.frag:446-491
    void nextIntersectDDA() {
        V3 dst = toV3(ray.step) - (exact - toV3(round));
        if (deltaPos.x < deltaPos.y && deltaPos.x < deltaPos.z) {
            round.x += ray.step.x;
            deltaPos.x += ray.absDelta.x;
            setMortonAxis(0);
            lastHit = 0;
        } else if (deltaPos.y < deltaPos.z) {
            round.y += ray.step.y;
            deltaPos.y += ray.absDelta.y;
            setMortonAxis(1);
            lastHit = 1;
        } else {
            round.z += ray.step.z;
            deltaPos.z += ray.absDelta.z;
            setMortonAxis(2);
            lastHit = 2;
        }
        exact = exact + dst[lastHit] * ray.ratios[lastHit];
        ddaSteps[phase]++;
    }

    // low_res.frag:493-532
    void getBlock() {
        const int maxDepth = groups + 1;
        currentMortonPos ^= mortonPos;
        int n = maxDepth - 1;
        while (currentMortonPos != 0) {
            n--;
            currentMortonPos >>= 6;
        }
        currentMortonPos = mortonPos;
        depth = n < depth ? n : depth;
        u32 posOffset = (u32)(maxDepth - 1 - depth) * 6;
        while (depth < maxDepth) {
            blockX = w->nodes[2 * (u64)stack[depth]];
            blockY = w->nodes[2 * (u64)stack[depth] + 1];
            nodeReads[phase]++;
            if (((u32)blockY & 1u) == 1u) return;
            if (depth == maxDepth - 1) {  // divergence (1): a branch at voxel level is malformed => empty
                blockX = ~(u64)0;
                blockY = 0;
                return;
            }
            posOffset -= 6;
            u32 index = (u32)(mortonPos >> posOffset) & 0x3f;
            if (((blockX >> index) & 1) == 0) {
                blockX = ~(u64)0;
                blockY = 0;
                return;
            }
            u32 child = w->arrays[64 * (u64)(u32)(blockY >> 32) + index];
            arrayReads[phase]++;
            if (child == 0 || child >= w->n_nodes) {  // divergence (1): null / out-of-pool child => empty
                blockX = ~(u64)0;
                blockY = 0;
                return;
            }
            stack[++depth] = child;
        }
    }

    bool empty() const { return blockX == ~(u64)0; }

    // low_res.frag:178-194
    void reflectRay() {
        deltaPos[lastHit] -= ray.absDelta[lastHit];
        ray.step[lastHit] *= -1;
        ray.dir[lastHit] *= -1.0f;
        finalColorMod = finalColorMod * 0.94f;
        rayReflected = true;
    }

    // low_res.frag:198-210
    static V3 refractDir(V3 rayDir, V3 normal, float n1, float n2) {
        float r = n1 / n2;
        float c1 = dot(normal, rayDir);
        if (c1 < 0) {
            normal = -normal;
            c1 = dot(normal, rayDir);
        }
        float c2 = sqrtf(1.0f - r * r * (1.0f - c1 * c1));
        return r * rayDir + (r * c1 - c2) * normal;
    }

    // low_res.frag:212-240
    void refractRay(float newRefractIndex, u32 flags) {
        finalColorMod = finalColorMod * ((flags & 0x10) > 0 ? v3(0.94f, 0.97f, 1.0f) : v3(0.95f, 0.95f, 0.95f));
        if (currentRefractiveIndex == newRefractIndex) return;
        if (lastHit != 0 && ray.step.x < 0) exact.x += 1;
        if (lastHit != 1 && ray.step.y < 0) exact.y += 1;
        if (lastHit != 2 && ray.step.z < 0) exact.z += 1;
        V3 normal = v3(0, 0, 0);
        normal[lastHit] = (float)ray.step[lastHit];
        if ((flags & 0x10) > 0) {
            normal.x += detSin((f->delta_time + exact.x * 0.2f - exact.z * 0.1f) * 10.0f) * 0.2f;
            normal = normalize(normal);
        }
        V3 nd = refractDir(ray.dir, normal, currentRefractiveIndex, newRefractIndex);
        ray = buildRay(nd);
        exact = exact + v3((float)(ray.step.x < 0 ? ray.step.x : 0), (float)(ray.step.y < 0 ? ray.step.y : 0),
                           (float)(ray.step.z < 0 ? ray.step.z : 0));
        deltaPos = ray.absDelta - (exact - toV3(round)) * ray.delta;
        currentRefractiveIndex = newRefractIndex;
    }

    // low_res.frag:242-244
    bool facingLightDir(V3 lightDir, u32 index) const { return lightDir[index] * (float)(-ray.step[index]) > 0; }

    // low_res.frag:246-252
    V3 calcLightIntensity(V3 color, V3 lightDir, u32 index) const {
        float mod = facingLightDir(lightDir, index) ? 0.15f : 0.0f;
        float intensity = minf(maxf(0.0f, lightDir[index] * (float)(-ray.step[index])) + 0.4f + mod, 1.0f);
        return color * intensity;
    }

    // low_res.frag:153-155
    static float sigmoid(float x, float scale, float dropOffSteepness) {
        return 1.0f / (1.0f + detExp(-x * dropOffSteepness)) * scale;
    }

    // low_res.frag:157-168
    V3 genSkyBox(V3 sunDir) {
        if (ray.dir.y < 0) ray.dir.y *= 1.4f;
        float haze = (0.1f - fabsf(clampf(ray.dir.y, -.3f, .3f))) * 0.8f + 0.1f;
        float modifier = clampf(sigmoid(1.0f - (haze * 2.0f), 1.0f, 2.0f), 0.0f, 1.0f);
        V3 skyDefaultColor = v3(0.2f, 0.4f, 1.0f);
        float b = distance3(ray.dir, sunDir) * 50.0f;
        V3 sun = v3(1, 1, 0) * (sigmoid(1.5f - b, 1.0f, 1.6f));
        return (skyDefaultColor + clampf(haze, 0.0f, 1.0f) * 3.0f) * modifier + sun;
    }
};

// low_res.frag:137-147 (21 bits per channel; every value is exact in fp32)
V3 colorIntToVec3(u64 color) {
    const double scaled = 1.0 / (double)(1 << 21);
    const u64 mask = (1u << 21) - 1;
    return v3((float)((double)(color >> 42) * scaled), (float)((double)((color >> 21) & mask) * scaled),
              (float)((double)(color & mask) * scaled));
}

inline uint8_t toUnorm8(float c) {
    float cl = c > 0.0f ? (c < 1.0f ? c : 1.0f) : 0.0f;  // NaN -> 0
    return (uint8_t)floorf(cl * 255.0f + 0.5f);
}

struct PixelOut {
    V3 color;
    OracleHit hit;
    bool shadowRay;
};

void writeOut(const OracleFrame* f, u32 px, u32 py, const PixelOut& o, OracleHit* hits, uint8_t* rgba8, float* rgba32f) {
    size_t idx = (size_t)py * f->width + px;
    if (hits) hits[idx] = o.hit;
    if (rgba8) {
        rgba8[4 * idx + 0] = toUnorm8(o.color.x);
        rgba8[4 * idx + 1] = toUnorm8(o.color.y);
        rgba8[4 * idx + 2] = toUnorm8(o.color.z);
        rgba8[4 * idx + 3] = 0;  // FragColor.a = 0 everywhere in the shader
    }
    if (rgba32f) {
        rgba32f[4 * idx + 0] = o.color.x;
        rgba32f[4 * idx + 1] = o.color.y;
        rgba32f[4 * idx + 2] = o.color.z;
        rgba32f[4 * idx + 3] = 0.0f;
    }
}

// AO, SURVEY 8 A10 (semantics defined by this build):
//  * normal axis a = primary face, sign sg = -step[a] of the primary ray at the hit
//  * table direction t = (tx, ty, up) maps to world by the signed permutation
//      d[a] = sg*up, d[(a+1)%3] = tx, d[(a+2)%3] = ty
//  * the ray starts exactly like the shadow ray (low_res.frag:373-380): from the
//    hit voxel, un-biased exact position re-biased for the new step signs,
//    deltaPos = absDelta - (exact-round)*delta, deltaPos[a] -= absDelta[a]
//  * at most ao_steps DDA steps while the block is empty or LIQUID
//    (low_res.frag:384); occluded iff the block is non-empty when the loop stops
//  * colour *= 1 - ao_strength * occluded/ao_samples
u32 castAo(Pixel& P, const I3& hitRound, const V3& hitExact, u32 face, int sg) {
    const OracleFrame* f = P.f;
    const float* tab = g_hemi.data();
    u32 occluded = 0;
    P.phase = 2;
    for (u32 s = 0; s < f->ao_samples; s++) {
        V3 d;
        d[(int)face] = (float)sg * tab[3 * s + 2];
        d[(int)(face + 1) % 3] = tab[3 * s + 0];
        d[(int)(face + 2) % 3] = tab[3 * s + 1];
        P.ray = buildRay(d);
        P.round = hitRound;
        for (int a = 0; a < 3; a++) P.setMortonAxis(a);
        P.exact = hitExact;
        if (P.ray.step.x < 0) P.exact.x -= 1;
        if (P.ray.step.y < 0) P.exact.y -= 1;
        if (P.ray.step.z < 0) P.exact.z -= 1;
        P.deltaPos = P.ray.absDelta - (P.exact - toV3(P.round)) * P.ray.delta;
        P.deltaPos[(int)face] -= P.ray.absDelta[(int)face];
        u32 i = f->ao_steps;
        P.blockX = ~(u64)0;
        P.blockY = ~(u64)0;
        while ((P.empty() || ((u32)P.blockY & 0x10) > 0) && i-- != 0) {
            P.nextIntersectDDA();
            P.getBlock();
        }
        if (!P.empty()) occluded++;
    }
    return occluded;
}

// low_res.frag:256-393
PixelOut shadePixel(Pixel& P, u32 px, u32 py) {
    const OracleFrame* f = P.f;
    PixelOut out;
    memset(&out.hit, 0, sizeof(out.hit));
    out.color = v3(0, 0, 0);
    out.shadowRay = false;
    const V3 origin = v3(f->origin[0], f->origin[1], f->origin[2]);
    const V3 cameraDir = v3(f->camera_dir[0], f->camera_dir[1], f->camera_dir[2]);
    const V3 sunDir = v3(f->sun_dir[0], f->sun_dir[1], f->sun_dir[2]);
    const float fragX = (float)px + 0.5f, fragY = (float)py + 0.5f;  // gl_FragCoord.xy

    // :259-262
    if ((f->flags & ORACLE_F_CROSSHAIR) && f->render_pos_data == 0) {
        float dx = fragX - f->mouse_pos[0], dy = fragY - f->mouse_pos[1];
        if (sqrtf(dx * dx + dy * dy) <= 3.0f) {
            out.color = v3(1, 1, 1);
            out.hit.kind = ORACLE_KIND_CROSSHAIR;
            return out;
        }
    }

    // :264-265
    const float pixelSizeX = 1.0f / (float)f->width, pixelSizeY = 1.0f / (float)f->height;
    const float fcx = fragX * pixelSizeX, fcy = fragY * pixelSizeY;

    // :267-277
    for (int i = 0; i < 8; i++) P.stack[i] = 0;
    P.depth = 0;
    P.currentMortonPos = 0;
    P.round = I3{(int)floorf(origin.x), (int)floorf(origin.y), (int)floorf(origin.z)};
    P.mortonPos = 0;
    for (int a = 0; a < 3; a++) P.setMortonAxis(a);  // originMortonPos (src/main.cpp:45-66,86)
    P.lastHit = 0;
    P.finalColorMod = v3(1, 1, 1);
    P.rayReflected = false;
    P.currentRefractiveIndex = 1.0f;
    P.phase = 0;

    // :279-285
    V3 planeLeft = cross(cameraDir, v3(0, 1, 0));
    V3 intersect = normalize(cameraDir + (planeLeft * -(f->proj_plane[0] * (fcx - 0.5f))) +
                             cross(cameraDir, planeLeft) * (-fcy + 0.5f) * f->proj_plane[1]);

    // :288-296
    P.ray = buildRay(intersect);
    P.exact = origin;
    if (P.ray.step.x < 0) P.exact.x -= 1;
    if (P.ray.step.y < 0) P.exact.y -= 1;
    if (P.ray.step.z < 0) P.exact.z -= 1;
    P.deltaPos = P.ray.absDelta - (P.exact - toV3(P.round)) * P.ray.delta;

    // :299-307
    P.getBlock();
    if (!P.empty()) {
        if ((P.blockY & 0x4) > 0) {
            P.currentRefractiveIndex = 1.1f;
        } else {
            out.color = colorIntToVec3(P.blockX);
            out.hit.kind = ORACLE_KIND_CAMERA_INSIDE;
            out.hit.x = P.round.x;
            out.hit.y = P.round.y;
            out.hit.z = P.round.z;
            return out;
        }
    }

    // :309-333
    u32 i = 0;
    const u32 numSteps = f->max_steps;
    while (P.empty() && i++ < numSteps) {
        P.nextIntersectDDA();
        P.getBlock();
    }
    while (i++ < numSteps) {
        if (!P.empty()) {
            u32 flags = (u32)P.blockY & 0x7;
            if (flags == 0x3) {
                P.reflectRay();
            } else if (flags == 0x5) {
                P.refractRay(1.1f, (u32)P.blockY);
            } else
                break;
        }
        P.nextIntersectDDA();
        P.getBlock();
    }

    out.hit.x = P.round.x;
    out.hit.y = P.round.y;
    out.hit.z = P.round.z;
    out.hit.steps = (uint16_t)P.ddaSteps[0];
    out.hit.face = (uint8_t)P.lastHit;
    out.hit.flags = (uint8_t)((P.rayReflected ? ORACLE_HF_REFLECTED : 0) |
                              (P.currentRefractiveIndex != 1.0f ? ORACLE_HF_REFRACTED : 0) |
                              (P.ray.step.x < 0 ? ORACLE_HF_STEPX_NEG : 0) | (P.ray.step.y < 0 ? ORACLE_HF_STEPY_NEG : 0) |
                              (P.ray.step.z < 0 ? ORACLE_HF_STEPZ_NEG : 0));

    // :335-341
    if (f->render_pos_data == 1) {
        out.color = P.exact;
        out.hit.kind = ORACLE_KIND_DEBUG;
        return out;
    } else if (f->render_pos_data == 2) {
        out.color = toV3(P.round);
        out.hit.kind = ORACLE_KIND_DEBUG;
        return out;
    }

    // :343-345
    if (P.lastHit != 0 && P.ray.step.x < 0) P.exact.x += 1;
    if (P.lastHit != 1 && P.ray.step.y < 0) P.exact.y += 1;
    if (P.lastHit != 2 && P.ray.step.z < 0) P.exact.z += 1;

    // :347-350
    if (f->looking_at[0] == P.round.x && f->looking_at[1] == P.round.y && f->looking_at[2] == P.round.z) {
        out.color = colorIntToVec3(P.blockX) * 2.0f + 0.3f;
        out.hit.kind = ORACLE_KIND_HIGHLIGHT;
        return out;
    }

    // :352-355
    if (P.empty()) {
        out.color = P.genSkyBox(sunDir) * P.finalColorMod;
        out.hit.kind = ORACLE_KIND_SKY;
        return out;
    }

    // :357-371
    out.hit.kind = ORACLE_KIND_HIT;
    const u32 origLastHit = P.lastHit;
    const I3 origPos = P.round;
    const V3 origExact = P.exact;
    const int normalSign = -P.ray.step[(int)origLastHit];
    V3 color = colorIntToVec3(P.blockX);
    out.color = P.calcLightIntensity(color, sunDir, P.lastHit) * P.finalColorMod;
    bool done = false;
    if (P.rayReflected) done = true;
    if (!done && !P.facingLightDir(sunDir, P.lastHit)) {
        out.color = color * 0.3f * P.finalColorMod;
        done = true;
    }
    if (!done) out.hit.flags |= ORACLE_HF_FACING_SUN;

    // :373-391
    if (!done && (f->flags & ORACLE_F_SHADOWS)) {
        P.phase = 1;
        out.shadowRay = true;
        P.ray = buildRay(sunDir);
        if (P.ray.step.x < 0) P.exact.x -= 1;
        if (P.ray.step.y < 0) P.exact.y -= 1;
        if (P.ray.step.z < 0) P.exact.z -= 1;
        P.deltaPos = P.ray.absDelta - (P.exact - toV3(P.round)) * P.ray.delta;
        P.deltaPos[(int)P.lastHit] -= P.ray.absDelta[(int)P.lastHit];
        u32 k = f->shadow_steps;
        P.blockX = ~(u64)0;
        P.blockY = ~(u64)0;
        while ((P.empty() || ((u32)P.blockY & 0x10) > 0) && k-- != 0) {
            P.nextIntersectDDA();
            P.getBlock();
        }
        out.hit.shadow_steps = (uint16_t)P.ddaSteps[1];
        if (!P.empty()) {
            out.color = color * 0.3f * P.finalColorMod;
            out.hit.flags |= ORACLE_HF_SHADOWED;
        }
    }

    if (f->ao_samples > 0) {
        u32 occ = castAo(P, origPos, origExact, origLastHit, normalSign);
        out.hit.ao_occluded = (uint8_t)occ;
        float ao = 1.0f - f->ao_strength * ((float)occ / (float)f->ao_samples);
        out.color = out.color * ao;
    }
    return out;
}

}  // namespace

static int g_lastThreads = 1;

extern "C" {

int oracle_render(const OracleWorld* world, const OracleFrame* frame, OracleHit* hits, uint8_t* rgba8, float* rgba32f,
                  OracleCounters* counters) {
    if (!world || !frame || !world->nodes || !world->arrays) return -1;
    if (world->max_depth < 2 || world->max_depth > 8) return -2;
    if (frame->width == 0 || frame->height == 0) return -3;
    if (frame->ao_samples > 255) return -4;
    u32 y0 = frame->row_begin, y1 = frame->row_end;
    if (y1 > frame->height) y1 = frame->height;
    if (y0 > y1) y0 = y1;
    if (frame->ao_samples > 0 && g_hemi.size() != 3 * (size_t)frame->ao_samples) {
        g_hemi.resize(3 * (size_t)frame->ao_samples);
        buildHemisphere(frame->ao_samples, g_hemi.data());
    }
    OracleCounters total;
    memset(&total, 0, sizeof(total));
#ifdef _OPENMP
    int nthreads = frame->threads > 0 ? frame->threads : omp_get_max_threads();
#pragma omp parallel num_threads(nthreads)
#endif
    {
        OracleCounters local;
        memset(&local, 0, sizeof(local));
#ifdef _OPENMP
        if (omp_get_thread_num() == 0) g_lastThreads = omp_get_num_threads();  // what the run really got (bench.py reports it)
#pragma omp for schedule(dynamic, 2)
#endif
        for (int64_t py = (int64_t)y0; py < (int64_t)y1; py++) {
            if (frame->band_stride > 1 && frame->band_rows > 0 &&
                ((u32)py / frame->band_rows) % frame->band_stride != frame->band_offset)
                continue;
            for (u32 px = 0; px < frame->width; px++) {
                Pixel P;
                P.w = world;
                P.f = frame;
                P.groups = world->max_depth - 1;
                P.amask = axisMask(P.groups);
                for (int k = 0; k < 3; k++) P.nodeReads[k] = P.arrayReads[k] = P.ddaSteps[k] = 0;
                PixelOut o = shadePixel(P, px, (u32)py);
                writeOut(frame, px, (u32)py, o, hits, rgba8, rgba32f);
                for (int k = 0; k < 3; k++) {
                    local.node_reads[k] += P.nodeReads[k];
                    local.array_reads[k] += P.arrayReads[k];
                    local.dda_steps[k] += P.ddaSteps[k];
                }
                local.kind_pixels[o.hit.kind]++;
                bool traced = o.hit.kind != ORACLE_KIND_CROSSHAIR;
                if (traced) local.rays[0]++;
                if (o.shadowRay) local.rays[1]++;
                if (o.hit.kind == ORACLE_KIND_HIT) local.rays[2] += frame->ao_samples;
            }
        }
#ifdef _OPENMP
#pragma omp critical
#endif
        {
            for (int k = 0; k < 3; k++) {
                total.node_reads[k] += local.node_reads[k];
                total.array_reads[k] += local.array_reads[k];
                total.dda_steps[k] += local.dda_steps[k];
                total.rays[k] += local.rays[k];
            }
            for (int k = 0; k < 6; k++) total.kind_pixels[k] += local.kind_pixels[k];
        }
    }
    if (counters) *counters = total;
    return 0;
}

void oracle_get_block(const OracleWorld* world, int32_t x, int32_t y, int32_t z, uint64_t* out) {
    Pixel P;
    P.w = world;
    P.f = nullptr;
    P.groups = world->max_depth - 1;
    P.amask = axisMask(P.groups);
    for (int i = 0; i < 8; i++) P.stack[i] = 0;
    P.depth = 0;
    P.currentMortonPos = 0;
    P.round = I3{x, y, z};
    P.mortonPos = 0;
    P.phase = 0;
    for (int k = 0; k < 3; k++) P.nodeReads[k] = P.arrayReads[k] = P.ddaSteps[k] = 0;
    for (int a = 0; a < 3; a++) P.setMortonAxis(a);
    P.getBlock();
    out[0] = P.blockX;
    out[1] = P.blockY;
}

void oracle_get_block_batch(const OracleWorld* world, const int32_t* xyz, uint64_t n, uint64_t* out) {
    for (uint64_t i = 0; i < n; i++) oracle_get_block(world, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], out + 2 * i);
}

uint64_t oracle_morton(int32_t x, int32_t y, int32_t z, int32_t max_depth) {
    int groups = max_depth - 1;
    return spreadAxis(x, groups) | (spreadAxis(y, groups) << 2) | (spreadAxis(z, groups) << 4);
}

// src/main.cpp:45-66 with its literal masks.
uint32_t oracle_morton_ref10(int32_t x, int32_t y, int32_t z) {
    int32_t c[3] = {x, y, z};
    uint32_t s[3];
    for (int a = 0; a < 3; a++) {
        int n = c[a];
        n = (n | (int)((uint32_t)n << 16)) & 0x030000FF;
        n = (n | (n << 8)) & 0x0300F00F;
        n = (n | (n << 4)) & 0x030C30C3;
        s[a] = (uint32_t)n;
    }
    return (s[2] << 4) | (s[1] << 2) | s[0];
}

void oracle_dda_walk(const float* origin, const float* dir, uint32_t n, int use_double, int32_t* out) {
    int step[3];
    int rnd[3];
    uint32_t last = 0;
    if (use_double) {
        // src/ray_caster.cpp:19-44,54-80 and test.cpp:78-117
        double delta[3], ad[3], ex[3], dp[3];
        for (int a = 0; a < 3; a++) {
            step[a] = dir[a] < 0 ? -1 : 1;
            delta[a] = (double)(1.0f / dir[a]);
            ad[a] = fabs(delta[a]);
            rnd[a] = (int)truncf(origin[a]);
            ex[a] = origin[a];
            if (step[a] < 0) ex[a] -= 1;
            dp[a] = ad[a] - (ex[a] - (double)rnd[a]) * delta[a];
        }
        while (n--) {
            if (dp[0] < dp[1] && dp[0] < dp[2]) last = 0;
            else if (dp[1] < dp[2]) last = 1;
            else last = 2;
            rnd[last] += step[last];
            dp[last] += ad[last];
        }
    } else {
        // low_res.frag:275,290-296,446-488
        float delta[3], ad[3], ex[3], dp[3];
        for (int a = 0; a < 3; a++) {
            step[a] = dir[a] < 0 ? -1 : 1;
            delta[a] = 1.0f / dir[a];
            ad[a] = fabsf(delta[a]);
            rnd[a] = (int)floorf(origin[a]);
            ex[a] = origin[a];
            if (step[a] < 0) ex[a] -= 1;
            dp[a] = ad[a] - (ex[a] - (float)rnd[a]) * delta[a];
        }
        while (n--) {
            if (dp[0] < dp[1] && dp[0] < dp[2]) last = 0;
            else if (dp[1] < dp[2]) last = 1;
            else last = 2;
            rnd[last] += step[last];
            dp[last] += ad[last];
        }
    }
    out[0] = rnd[0];
    out[1] = rnd[1];
    out[2] = rnd[2];
    out[3] = (int32_t)last;
}

float oracle_sinf(float x) { return detSin(x); }
float oracle_expf(float x) { return detExp(x); }

void oracle_hemisphere(uint32_t n, float* out) { buildHemisphere(n, out); }

/* OpenMP threads the last oracle_render really ran on (OMP_NUM_THREADS, cgroup limits and nested launchers such as
 * torchrun can give fewer than asked for). */
int oracle_last_threads(void) { return g_lastThreads; }

void oracle_set_hemisphere(uint32_t n, const float* dirs) { g_hemi.assign(dirs, dirs + 3 * (size_t)n); }

}  // extern "C"
