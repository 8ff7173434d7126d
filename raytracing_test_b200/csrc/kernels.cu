// kernels.cu -- hand-written sm_100a traversal kernels of the voxel ray caster.
//
// What the reference does per pixel (src/shaders/low_res.frag:256-393): a 3-D DDA
// that advances ONE voxel per step (nextIntersectDDA, :446-491) and asks the tree
// for the new voxel after EVERY step (getBlock, :493-532: a 16-byte node read per
// level still shared with the previous voxel + a 4-byte child-index read per level
// descended).  Hit voxel, face and step count depend on the exact sequence of fp32
// additions `deltaPos[a] += absDelta[a]`, so that arithmetic is kept as is, in
// registers, with explicit round-to-nearest intrinsics (no FMA contraction).
//
// What is redesigned is the memory side:
//   * the world is the packed breadth-first form of packed_world.h: one 128-bit
//     __ldg per level, popcount child addressing, 4^3 bricks at the bottom;
//   * the tree is consulted only when the ray LEAVES the uniform cell it is in.
//     A lookup that answers "empty" also answers "empty for the whole 4^k cell",
//     so the kernel counts down the voxels left to each face of that cell and
//     steps in registers until one counter hits zero (the reference re-reads
//     nodes[stack[depth]] on every step: ~178 node reads per primary ray in the
//     default scene, of which only ~7.5 are real descents);
//   * a re-descent starts from the deepest cached node that still contains the
//     voxel (the reference's stack, collapsed to one cached node in registers);
//   * pos.exact (needed only to seed the shadow/AO ray and for refraction) is
//     carried with the same per-step arithmetic as the shader.
//
// Rays are issued in 8x4-pixel warp tiles (coherent neighbours share cells, so
// their lookups hit the same L1 lines).
#include "kernels.h"

#include "free_flight.h"
#include "packed_world.h"

namespace rc {

// ---------------------------------------------------------------------------
// exact fp32 helpers: one IEEE round-to-nearest operation each, never contracted
// ---------------------------------------------------------------------------
#define FADD(a, b) __fadd_rn((a), (b))
#define FSUB(a, b) __fsub_rn((a), (b))
#define FMUL(a, b) __fmul_rn((a), (b))
#define FDIV(a, b) __fdiv_rn((a), (b))
#define FSQRT(a) __fsqrt_rn((a))

__constant__ float c_hemi[3 * RC_MAX_AO_SAMPLES];      // (x, y, up) per sample
__constant__ float c_hemi_inv[3 * RC_MAX_AO_SAMPLES];  // (1/x, 1/y, 1/up)
// The window kernel's oriented table lists the samples grouped by the signs of their two tangent components:
// c_aoSeg[g] .. c_aoSeg[g + 1] is group g = (x < 0) | (y < 0) << 1.  Inside a group every ray of a pixel steps the same
// way along all three axes (the normal component's sign is the pixel's), so the per-axis choice between the two
// (exact - round) values leaves the ray loop.  The order of the samples does not matter: occlusion is a count.
__constant__ uint32_t c_aoSeg[5];

struct F3 {
    float x, y, z;
};
struct I3v {
    int x, y, z;
};

__device__ __forceinline__ float sel3(int a, float x, float y, float z) { return a == 0 ? x : (a == 1 ? y : z); }
__device__ __forceinline__ int sel3i(int a, int x, int y, int z) { return a == 0 ? x : (a == 1 ? y : z); }

__device__ __forceinline__ float glsl_min(float a, float b) { return b < a ? b : a; }
__device__ __forceinline__ float glsl_max(float a, float b) { return a < b ? b : a; }
__device__ __forceinline__ float glsl_clamp(float x, float lo, float hi) { return glsl_min(glsl_max(x, lo), hi); }

__device__ __forceinline__ float dot3(F3 a, F3 b) { return FADD(FADD(FMUL(a.x, b.x), FMUL(a.y, b.y)), FMUL(a.z, b.z)); }
__device__ __forceinline__ F3 scale3(F3 a, float s) { return F3{FMUL(a.x, s), FMUL(a.y, s), FMUL(a.z, s)}; }
__device__ __forceinline__ F3 add3(F3 a, F3 b) { return F3{FADD(a.x, b.x), FADD(a.y, b.y), FADD(a.z, b.z)}; }
__device__ __forceinline__ F3 normalize3(F3 v) { return scale3(v, FDIV(1.0f, FSQRT(dot3(v, v)))); }

// Deterministic sin / exp: the same sequence of fp32 operations as the oracle's
// detSin / detExp (DESIGN.md "deterministic transcendentals").
__device__ __forceinline__ float det_sin(float x) {
    if (!(fabsf(x) <= 1.0e6f)) return 0.0f;
    float k = floorf(FADD(FMUL(x, 0.15915494309189535f), 0.5f));
    float r = FSUB(x, FMUL(k, 6.28125f));
    r = FSUB(r, FMUL(k, 1.9350051879882812e-3f));
    r = FSUB(r, FMUL(k, 3.0199160691085e-7f));
    const float pi = 3.14159265358979f;
    const float halfPi = 1.57079632679490f;
    if (r > halfPi) r = FSUB(pi, r);
    if (r < -halfPi) r = FSUB(-pi, r);
    float s = FMUL(r, r);
    float p = -2.5052108385441720e-08f;
    p = FADD(FMUL(p, s), 2.7557319223985893e-06f);
    p = FADD(FMUL(p, s), -1.9841269841269841e-04f);
    p = FADD(FMUL(p, s), 8.3333333333333332e-03f);
    p = FADD(FMUL(p, s), -1.6666666666666666e-01f);
    return FADD(r, FMUL(FMUL(r, s), p));
}

__device__ __forceinline__ float det_exp(float x) {
    if (x != x) return x;
    if (x > 88.0f) x = 88.0f;
    if (x < -87.0f) x = -87.0f;
    float k = floorf(FADD(FMUL(x, 1.44269504088896341f), 0.5f));
    float r = FSUB(x, FMUL(k, 0.693359375f));
    r = FSUB(r, FMUL(k, -2.12194440e-4f));
    float p = 1.9875691500e-4f;
    p = FADD(FMUL(p, r), 1.3981999507e-3f);
    p = FADD(FMUL(p, r), 8.3334519073e-3f);
    p = FADD(FMUL(p, r), 4.1665795894e-2f);
    p = FADD(FMUL(p, r), 1.6666665459e-1f);
    p = FADD(FMUL(p, r), 5.0000001201e-1f);
    float e = FADD(FADD(FMUL(p, FMUL(r, r)), r), 1.0f);
    int ki = (int)k;
    return FMUL(e, __int_as_float((ki + 127) << 23));
}

// ---------------------------------------------------------------------------
// Tree cursor: one cached PackedNode (never a PK_LEAF) + the last voxel looked up.
// ---------------------------------------------------------------------------
struct Cursor {
    unsigned long long mask;
    uint32_t base;
    uint32_t kind;
    int shift;       // cell shift of the cached node's children
    int lx, ly, lz;  // wrapped coordinates of a voxel inside the cached node
    // SCHED_PERSISTENT_STAGED: the first ntop nodes (breadth-first prefix = the top tree levels) are also in shared memory.
    // Everywhere else ntop is the constant 0 and the test below folds away.
    const uint4* top;
    uint32_t ntop;
};

__device__ __forceinline__ uint4 load_node(const DeviceWorld& W, const Cursor& c, uint32_t idx) {
    if (idx < c.ntop) return c.top[idx];
    return __ldg(&W.nodes[idx]);
}

// Where a descent starts when the voxel lies outside the cached node.  With the top table (DeviceWorld::top_table): one
// 4-byte load names the node of the 2^top_shift cell around the voxel -- or says that the cell is empty, and how big the empty
// cell around it is -- instead of a walk down from the root (one dependent 16-byte load per level: 3 of them to a 16^3 node
// of the reference's 1024^3 torus).  Returns true when r is already the answer (an empty or uniform cell); else the cursor
// holds a node that contains the voxel.  Without the table (rc_debug_set_option "top_table" 0, or the top levels staged in
// shared memory: SCHED_PERSISTENT_STAGED) the descent starts at the root.
__device__ __forceinline__ void cursor_to_root(const DeviceWorld& W, Cursor& c) {
    const uint4 n = load_node(W, c, 0u);
    c.mask = (unsigned long long)n.x | ((unsigned long long)n.y << 32);
    c.base = n.z;
    c.kind = n.w;
    c.shift = W.root_shift;
}

// No node cached yet: the first lookup starts at the top (lx is a coordinate no voxel has).
__device__ __forceinline__ void cursor_reset(const DeviceWorld& W, Cursor& c, const uint4* top = nullptr, uint32_t ntop = 0u) {
    c.top = top;
    c.ntop = ntop;
    c.mask = 0ull;
    c.base = 0u;
    c.kind = 0u;
    c.shift = W.root_shift;
    c.lx = c.ly = c.lz = 0x40000000;
}

// Top-table entry of the cell around wrapped voxel (wx, wy, wz): bit 31 set = index of the node of that cell (its children
// are cells of shift top_shift - 2; a PK_LEAF node = the cell is uniform), else the shift of the empty aligned cell around it.
__device__ __forceinline__ uint32_t top_entry(const DeviceWorld& W, int wx, int wy, int wz) {
    const uint32_t cx = (uint32_t)wx >> W.top_shift, cy = (uint32_t)wy >> W.top_shift, cz = (uint32_t)wz >> W.top_shift;
    return __ldg(&W.top_table[(((cz << W.top_bits) | cy) << W.top_bits) | cx]);
}

// Voxel coordinate -> tree coordinate. The reference spreads a coordinate into its Morton code with
// n = (n | n << 16) & 0x030000FF ... (low_res.frag:454-459): bits 0-9 survive (a 1024-torus), and
// bits 24-25 of the RAW value are OR-ed onto bits 8-9, so a negative coordinate lands in
// 768 | (x & 255).  W.leak = 0x300 reproduces that for max_depth 6; deeper trees (extension) use a
// plain torus (leak 0).  The low 8 bits are never touched, so aligned cells up to the root's
// children (256^3) keep their extent in raw coordinates and the countdown of count_cell holds.
__device__ __forceinline__ int wrap_coord(const DeviceWorld& W, int v) {
    return (v & (int)W.wrap) | ((int)((unsigned)v >> 16) & (int)W.leak);
}

// Material word of voxel (x,y,z) (0 = empty) and the shift of the uniform cell it
// lies in. Equivalent of getBlock (low_res.frag:493-532) on the packed form.
__device__ __forceinline__ uint32_t lookup(const DeviceWorld& W, Cursor& c, int x, int y, int z, int& cellShift) {
    const int wx = wrap_coord(W, x), wy = wrap_coord(W, y), wz = wrap_coord(W, z);
    if ((((wx ^ c.lx) | (wy ^ c.ly) | (wz ^ c.lz)) >> (c.shift + 2)) != 0) {
        if (W.top_table && c.ntop == 0u) {
            const uint32_t e = top_entry(W, wx, wy, wz);
            if (!(e & 0x80000000u)) {  // empty: the cursor keeps its node
                cellShift = (int)e;
                return 0u;
            }
            const uint4 n = __ldg(&W.nodes[e & 0x7FFFFFFFu]);
            if ((n.w & PK_KIND_MASK) == PK_LEAF) {
                cellShift = W.top_shift;
                return n.z;
            }
            c.mask = (unsigned long long)n.x | ((unsigned long long)n.y << 32);
            c.base = n.z;
            c.kind = n.w;
            c.shift = W.top_shift - 2;
        } else {
            cursor_to_root(W, c);
        }
    }
    c.lx = wx;
    c.ly = wy;
    c.lz = wz;
    for (;;) {
        const int idx = (((wz >> c.shift) & 3) << 4) | (((wy >> c.shift) & 3) << 2) | ((wx >> c.shift) & 3);
        if (!((c.mask >> idx) & 1ull)) {
            cellShift = c.shift;
            return 0u;
        }
        const uint32_t rank = (uint32_t)__popcll(c.mask & ((1ull << idx) - 1ull));
        if ((c.kind & PK_KIND_MASK) == PK_BRICK) {
            cellShift = 0;
            return __ldg(&W.voxels[c.base + rank]);
        }
        const uint4 n = load_node(W, c, c.base + rank);
        if ((n.w & PK_KIND_MASK) == PK_LEAF) {
            cellShift = c.shift;
            return n.z;
        }
        c.mask = (unsigned long long)n.x | ((unsigned long long)n.y << 32);
        c.base = n.z;
        c.kind = n.w;
        c.shift -= 2;
    }
}

__device__ __forceinline__ uint32_t word_flags(uint32_t word) { return word >> MATERIAL_FLAG_SHIFT; }

// ---------------------------------------------------------------------------
// Region: the largest aligned piece of the world the tree describes with ONE node around a
// voxel -- an empty 4^k cell (mask 0), a uniform solid 4^k cell (mask ~0, base = its material
// word), or a 4^3 brick (mask = occupancy, base = first voxel word).  While a ray stays inside
// a region the tree is not touched: emptiness is a bit test in registers.
// ---------------------------------------------------------------------------
constexpr uint32_t BR_UNIFORM = 0x40000000u;    // the whole cell is one material, base = its word
constexpr uint32_t BR_LIQUID = 0x80000000u;     // some voxel of the cell is LIQUID: the word decides

struct Region {
    unsigned long long mask;  // occupancy of the 4^3 cell, bit z<<4|y<<2|x (0 / ~0 for uniform cells)
    uint32_t base;            // voxels[base + rank], or the material word (BR_UNIFORM)
    uint32_t info;            // BR_* (the AO kernel keeps its cache key in the low 30 bits)
    int shift;                // log2 of the region's edge (2 for a brick)
};

// Region around wrapped voxel (wx,wy,wz); descends from the deepest cached node that contains it.
__device__ __forceinline__ void fetch_region(const DeviceWorld& W, Cursor& c, int wx, int wy, int wz, Region& r) {
    if ((((wx ^ c.lx) | (wy ^ c.ly) | (wz ^ c.lz)) >> (c.shift + 2)) != 0) {
        if (W.top_table && c.ntop == 0u) {
            const uint32_t e = top_entry(W, wx, wy, wz);
            if (!(e & 0x80000000u)) {  // empty: the cursor keeps its node
                r.mask = 0ull;
                r.base = 0u;
                r.info = 0u;
                r.shift = (int)e;
                return;
            }
            const uint4 n = __ldg(&W.nodes[e & 0x7FFFFFFFu]);
            if ((n.w & PK_KIND_MASK) == PK_LEAF) {
                r.mask = ~0ull;
                r.base = n.z;
                r.info = BR_UNIFORM | ((word_flags(n.z) & 0x10u) ? BR_LIQUID : 0u);
                r.shift = W.top_shift;
                return;
            }
            c.mask = (unsigned long long)n.x | ((unsigned long long)n.y << 32);
            c.base = n.z;
            c.kind = n.w;
            c.shift = W.top_shift - 2;
        } else {
            cursor_to_root(W, c);
        }
    }
    c.lx = wx;
    c.ly = wy;
    c.lz = wz;
    for (;;) {
        if (c.shift == 0) {  // the cached node is itself the brick around the voxel
            r.mask = c.mask;
            r.base = c.base;
            r.info = (c.kind & PK_HAS_LIQUID) ? BR_LIQUID : 0u;
            r.shift = 2;
            return;
        }
        const int idx = (((wz >> c.shift) & 3) << 4) | (((wy >> c.shift) & 3) << 2) | ((wx >> c.shift) & 3);
        if (!((c.mask >> idx) & 1ull)) {
            r.mask = 0ull;
            r.base = 0u;
            r.info = 0u;
            r.shift = c.shift;
            return;
        }
        const uint32_t rank = (uint32_t)__popcll(c.mask & ((1ull << idx) - 1ull));
        const uint4 n = load_node(W, c, c.base + rank);
        if ((n.w & PK_KIND_MASK) == PK_LEAF) {
            r.mask = ~0ull;
            r.base = n.z;
            r.info = BR_UNIFORM | ((word_flags(n.z) & 0x10u) ? BR_LIQUID : 0u);
            r.shift = c.shift;
            return;
        }
        if (c.shift == 2) {  // n is the brick node of this 4^3 cell
            r.mask = (unsigned long long)n.x | ((unsigned long long)n.y << 32);
            r.base = n.z;
            r.info = (n.w & PK_HAS_LIQUID) ? BR_LIQUID : 0u;
            r.shift = 2;
            return;
        }
        c.mask = (unsigned long long)n.x | ((unsigned long long)n.y << 32);
        c.base = n.z;
        c.kind = n.w;
        c.shift -= 2;
    }
}

__device__ __forceinline__ int brick_bit(int x, int y, int z) { return ((z & 3) << 4) | ((y & 3) << 2) | (x & 3); }

// Material word of an occupied voxel (bit idx of r.mask is set).
__device__ __forceinline__ uint32_t region_word(const DeviceWorld& W, const Region& r, int idx) {
    return (r.info & BR_UNIFORM) ? r.base : __ldg(&W.voxels[r.base + (uint32_t)__popcll(r.mask & ((1ull << idx) - 1ull))]);
}

// State of a shadow / AO ray after entering voxel (x,y,z) of region r: 0 empty, 1 LIQUID (the
// ray goes on, low_res.frag:384), 2 opaque.
__device__ __forceinline__ uint32_t occlusion_state(const DeviceWorld& W, const Region& r, int x, int y, int z) {
    if (r.mask == 0ull) return 0u;
    const int idx = brick_bit(x, y, z);
    if (!((r.mask >> idx) & 1ull)) return 0u;
    if (!(r.info & BR_LIQUID)) return 2u;
    return (word_flags(region_word(W, r, idx)) & 0x10u) ? 1u : 2u;
}

// ---------------------------------------------------------------------------
// Ray + position state (the shader's `ray` and `pos` globals)
// ---------------------------------------------------------------------------
struct Ray {
    float dx, dy, dz;     // ray.dir
    int sx, sy, sz;       // ray.step
    float adx, ady, adz;  // ray.absDelta
    float ddx, ddy, ddz;  // ray.delta
    // ratios columns (low_res.frag:75-91): ratios[a] = dir / dir[a], diagonal = 1
    float r0y, r0z, r1x, r1z, r2x, r2y;
};

struct Pos {
    int x, y, z;       // pos.round
    float ex, ey, ez;  // pos.exact
    float tx, ty, tz;  // pos.deltaPos
};

// low_res.frag:61-97
__device__ __forceinline__ void build_ray(Ray& r, float dx, float dy, float dz, bool withRatios) {
    r.dx = dx;
    r.dy = dy;
    r.dz = dz;
    r.sx = dx < 0.0f ? -1 : 1;
    r.sy = dy < 0.0f ? -1 : 1;
    r.sz = dz < 0.0f ? -1 : 1;
    if (withRatios) {
        r.r0y = FDIV(dy, dx);
        r.r0z = FDIV(dz, dx);
        r.r1x = FDIV(dx, dy);
        r.r1z = FDIV(dz, dy);
        r.r2x = FDIV(dx, dz);
        r.r2y = FDIV(dy, dz);
    }
    r.ddx = FDIV(1.0f, dx);
    r.ddy = FDIV(1.0f, dy);
    r.ddz = FDIV(1.0f, dz);
    r.adx = fabsf(r.ddx);
    r.ady = fabsf(r.ddy);
    r.adz = fabsf(r.ddz);
}

// pos.deltaPos = ray.absDelta - (pos.exact - pos.round) * ray.delta   (low_res.frag:296,379)
__device__ __forceinline__ void init_delta_pos(const Ray& r, Pos& p) {
    p.tx = FSUB(r.adx, FMUL(FSUB(p.ex, (float)p.x), r.ddx));
    p.ty = FSUB(r.ady, FMUL(FSUB(p.ey, (float)p.y), r.ddy));
    p.tz = FSUB(r.adz, FMUL(FSUB(p.ez, (float)p.z), r.ddz));
}

// Steps left before the ray leaves the aligned 2^shift cell around (x,y,z), per axis.
struct CellCount {
    int nx, ny, nz;
};
__device__ __forceinline__ void count_cell(CellCount& n, const Ray& r, const Pos& p, int shift) {
    const int m = (1 << shift) - 1;
    const int fx = p.x & m, fy = p.y & m, fz = p.z & m;
    n.nx = r.sx > 0 ? (m + 1 - fx) : (fx + 1);
    n.ny = r.sy > 0 ? (m + 1 - fy) : (fy + 1);
    n.nz = r.sz > 0 ? (m + 1 - fz) : (fz + 1);
}

// nextIntersectDDA (low_res.frag:446-491). Returns the axis stepped.
// TRACK_EXACT: also advance pos.exact (only the primary ray needs it).
template <bool TRACK_EXACT>
__device__ __forceinline__ int dda_step(const Ray& r, Pos& p, CellCount& n, bool& left) {
    int a;
    if (p.tx < p.ty && p.tx < p.tz) {
        if (TRACK_EXACT) {
            const float dst = FSUB((float)r.sx, FSUB(p.ex, (float)p.x));
            p.ex = FADD(p.ex, dst);
            p.ey = FADD(p.ey, FMUL(dst, r.r0y));
            p.ez = FADD(p.ez, FMUL(dst, r.r0z));
        }
        p.x += r.sx;
        p.tx = FADD(p.tx, r.adx);
        left = (--n.nx == 0);
        a = 0;
    } else if (p.ty < p.tz) {
        if (TRACK_EXACT) {
            const float dst = FSUB((float)r.sy, FSUB(p.ey, (float)p.y));
            p.ex = FADD(p.ex, FMUL(dst, r.r1x));
            p.ey = FADD(p.ey, dst);
            p.ez = FADD(p.ez, FMUL(dst, r.r1z));
        }
        p.y += r.sy;
        p.ty = FADD(p.ty, r.ady);
        left = (--n.ny == 0);
        a = 1;
    } else {
        if (TRACK_EXACT) {
            const float dst = FSUB((float)r.sz, FSUB(p.ez, (float)p.z));
            p.ex = FADD(p.ex, FMUL(dst, r.r2x));
            p.ey = FADD(p.ey, FMUL(dst, r.r2y));
            p.ez = FADD(p.ez, dst);
        }
        p.z += r.sz;
        p.tz = FADD(p.tz, r.adz);
        left = (--n.nz == 0);
        a = 2;
    }
    return a;
}

// nextIntersectDDA without pos.exact and without divergence: the three-way choice becomes
// predicates.  Same comparisons, same additions as dda_step.
struct StepAxis {
    bool x, y;  // neither: z
};
__device__ __forceinline__ StepAxis dda_step_bl(const Ray& r, Pos& p) {
    StepAxis a;
    a.x = p.tx < p.ty && p.tx < p.tz;
    a.y = !a.x && p.ty < p.tz;
    const bool az = !a.x && !a.y;
    if (a.x) { p.x += r.sx; p.tx = FADD(p.tx, r.adx); }
    if (a.y) { p.y += r.sy; p.ty = FADD(p.ty, r.ady); }
    if (az) { p.z += r.sz; p.tz = FADD(p.tz, r.adz); }
    return a;
}
__device__ __forceinline__ int axis_index(StepAxis a) { return a.x ? 0 : (a.y ? 1 : 2); }

// The same step with pos.exact carried along (low_res.frag:447,489: dst is taken BEFORE the
// update, exact += dst[a] * ratios[a] with ratios[a][a] = 1), still branch-free.
__device__ __forceinline__ StepAxis dda_step_exact_bl(const Ray& r, Pos& p) {
    StepAxis a;
    a.x = p.tx < p.ty && p.tx < p.tz;
    a.y = !a.x && p.ty < p.tz;
    const bool az = !a.x && !a.y;
    const float ea = a.x ? p.ex : (a.y ? p.ey : p.ez);
    const int ra = a.x ? p.x : (a.y ? p.y : p.z);
    const int sa = a.x ? r.sx : (a.y ? r.sy : r.sz);
    const float dst = FSUB((float)sa, FSUB(ea, (float)ra));
    const float rx = a.x ? 1.0f : (a.y ? r.r1x : r.r2x);
    const float ry = a.x ? r.r0y : (a.y ? 1.0f : r.r2y);
    const float rz = a.x ? r.r0z : (a.y ? r.r1z : 1.0f);
    p.ex = FADD(p.ex, FMUL(dst, rx));
    p.ey = FADD(p.ey, FMUL(dst, ry));
    p.ez = FADD(p.ez, FMUL(dst, rz));
    if (a.x) { p.x += r.sx; p.tx = FADD(p.tx, r.adx); }
    if (a.y) { p.y += r.sy; p.ty = FADD(p.ty, r.ady); }
    if (az) { p.z += r.sz; p.tz = FADD(p.tz, r.adz); }
    return a;
}

// Region that holds the ray's voxel + where it was fetched (raw coordinates: aligned cells up to
// the root's children keep their extent under wrap_coord).
struct RegionAt {
    Region r;
    int x, y, z;
};
__device__ __forceinline__ bool left_region(const RegionAt& g, const Pos& p) {
    return ((((p.x ^ g.x) | (p.y ^ g.y) | (p.z ^ g.z)) >> g.r.shift) != 0);
}
__device__ __forceinline__ void enter_region(const DeviceWorld& W, Cursor& c, RegionAt& g, const Pos& p) {
    g.x = p.x;
    g.y = p.y;
    g.z = p.z;
    fetch_region(W, c, wrap_coord(W, p.x), wrap_coord(W, p.y), wrap_coord(W, p.z), g.r);
}

// Loop A's region entry: the column above the terrain as one empty region when the ray is above its column block's top
// (FrameArgs::col_tops), else the tree.  colNy = steps along y before the ray leaves that region (down: to the top; up: to
// the edge of the torus), 0 = a tree region.  Raw coordinates must be on the torus (negative ones go through the
// reference's Morton quirk, wrap_coord: the tree decides there).
__device__ __forceinline__ void enter_region_a(const FrameArgs& A, Cursor& c, RegionAt& g, const Pos& p, int sy, int& colNy) {
    const DeviceWorld& W = A.world;
    colNy = 0;
    if (A.col_tops && (uint32_t)p.x <= W.wrap && (uint32_t)p.y <= W.wrap && (uint32_t)p.z <= W.wrap) {
        const int top = __ldg(&A.col_tops[(((uint32_t)p.z >> A.sky_shift) * SKY_GRID) + ((uint32_t)p.x >> A.sky_shift)]);
        if (p.y > top) {
            g.x = p.x;
            g.y = p.y;
            g.z = p.z;
            g.r.mask = 0ull;
            g.r.base = 0u;
            g.r.info = 0u;
            g.r.shift = A.sky_shift;  // extent along x and z; along y: colNy
            colNy = sy > 0 ? (int)W.wrap + 1 - p.y : p.y - top;
#ifdef RC_MUTATE_COLUMN  // mutation check of tests/test_sky_gates_random_gpu.py (tools/mutation_check.sh): one step too many
            if (sy <= 0) colNy++;
#endif
            return;
        }
    }
    enter_region(W, c, g, p);
}

// After a lookup: refresh the countdown. A solid voxel forces a lookup on the next step.
__device__ __forceinline__ void after_lookup(CellCount& n, const Ray& r, const Pos& p, uint32_t word, int cellShift) {
    if (word == 0u) {
        count_cell(n, r, p, cellShift);
    } else {
        n.nx = n.ny = n.nz = 1;
    }
}

// low_res.frag:137-147
__device__ __forceinline__ F3 decode_color(unsigned long long color) {
    const float s = 1.0f / 2097152.0f;
    F3 c;
    c.x = FMUL((float)(uint32_t)(color >> 42), s);
    c.y = FMUL((float)(uint32_t)((color >> 21) & 0x1FFFFFull), s);
    c.z = FMUL((float)(uint32_t)(color & 0x1FFFFFull), s);
    return c;
}

__device__ __forceinline__ unsigned char to_unorm8(float c) {
    const float cl = c > 0.0f ? (c < 1.0f ? c : 1.0f) : 0.0f;
    return (unsigned char)floorf(FADD(FMUL(cl, 255.0f), 0.5f));
}

// low_res.frag:153-155
__device__ __forceinline__ float sigmoid(float x, float scale, float steep) {
    return FMUL(FDIV(1.0f, FADD(1.0f, det_exp(FMUL(-x, steep)))), scale);
}

// low_res.frag:157-168 (dirY is ray.dir.y, which the function scales in place)
__device__ __forceinline__ F3 gen_sky(float dirX, float dirY, float dirZ, F3 sun) {
    if (dirY < 0.0f) dirY = FMUL(dirY, 1.4f);
    const float haze = FADD(FMUL(FSUB(0.1f, fabsf(glsl_clamp(dirY, -.3f, .3f))), 0.8f), 0.1f);
    const float modifier = glsl_clamp(sigmoid(FSUB(1.0f, FMUL(haze, 2.0f)), 1.0f, 2.0f), 0.0f, 1.0f);
    const F3 d = F3{FSUB(dirX, sun.x), FSUB(dirY, sun.y), FSUB(dirZ, sun.z)};
    const float b = FMUL(FSQRT(dot3(d, d)), 50.0f);
    const float s = sigmoid(FSUB(1.5f, b), 1.0f, 1.6f);
    const float h3 = FMUL(glsl_clamp(haze, 0.0f, 1.0f), 3.0f);
    F3 out;
    out.x = FADD(FMUL(FADD(0.2f, h3), modifier), FMUL(1.0f, s));
    out.y = FADD(FMUL(FADD(0.4f, h3), modifier), FMUL(1.0f, s));
    out.z = FADD(FMUL(FADD(1.0f, h3), modifier), FMUL(0.0f, s));
    return out;
}

// Sky gates (FrameArgs::sky_gate).  A ray at voxel (x, y, z) that steps UP, heading (sxNeg, szNeg) horizontally, is above
// every occupied voxel of the columns it can still reach when the value returned is > 0 (clear by that much; <= 0: that
// many steps up, plus one, to go): near blocks as they are, each farther ring less what the ray must have climbed before
// it can be there.  Climbed per voxel of horizontal (Chebyshev) distance: at least min(absDelta.x, absDelta.z) /
// absDelta.y steps up per step sideways (events fire in time order), give or take 2 for the phases and 1 for the
// rounding of the sums; a ring whose nearest block is d blocks away is at least (d - 1) * B voxels away.  merge = false
// (a NaN deltaPos on y or z -- 0 * inf in buildRay -- makes z fire every time, whatever the times): no climb to count on;
// fmaxf / fminf also turn a NaN product into 0.  The caller makes sure x and z are on the torus and y + the steps left is.
__device__ __forceinline__ int sky_gate_distance(const FrameArgs& A, int x, int y, int z, bool sxNeg, bool szNeg, float adx, float adz,
                                                 float dy, bool merge) {
    const uint32_t q = (sxNeg ? 1u : 0u) | (szNeg ? 2u : 0u);
    const uint4 g = __ldg(&A.sky_gate[(q * SKY_GRID + ((uint32_t)z >> A.sky_shift)) * SKY_GRID + ((uint32_t)x >> A.sky_shift)]);
    const float slope = merge ? fminf(fmaxf(FMUL(fminf(adx, adz), fabsf(dy)), 0.0f), 16.0f) * (float)(1 << A.sky_shift) : 0.0f;
#ifdef RC_MUTATE_GATE  // mutation check (tools/mutation_check.sh): credit the climb without the slack for phases / rounding
    const int y1 = y + 1 + 3;
#else
    const int y1 = y + 1;
#endif
    int d = y1 - (int)(g.x & 0xFFFFu);
    d = min(d, y1 + max(__float2int_rd(slope) - 3, 0) - (int)(g.x >> 16));
    d = min(d, y1 + max(__float2int_rd(slope * 3.0f) - 3, 0) - (int)(g.y & 0xFFFFu));
    d = min(d, y1 + max(__float2int_rd(slope * 7.0f) - 3, 0) - (int)(g.y >> 16));
    d = min(d, y1 + max(__float2int_rd(slope * 15.0f) - 3, 0) - (int)g.z);
    return d;
}

// The same question for a ray that steps DOWN (FrameArgs::drop_gate): a grazing ray that loses only a few voxels of height
// over its whole step budget is above everything it can still reach long before it has taken its steps -- the 6 % of the
// benchmark frame's pixels just under the horizon walk more than half of all the steps the kernel walks.  `left` events are
// still to fire.  Events fire in time order (merge of three arithmetic sequences, see free_flight.h), so by the time tau of
// the last of them every axis a has fired at least (tau - now) / absDelta[a] - 1: left >= (tau - now) * (|dx| + |dy| + |dz|)
// - 3, and y has fired at most (tau - now) / absDelta.y + 1 <= (left + 3) * |dy| / (|dx| + |dy| + |dz|) + 1 -- the ray's
// whole drop (`total`, with 1 more for the rounding of the sums: the additions are off by half an ulp of t < 2048 each,
// far less than one event over 1024 steps).  And while the ray is still within d blocks (Chebyshev) of this one it has
// stepped fewer than (d + 1) * B times along x and along z, so less time than (d + 1) * B * min(absDelta.x, absDelta.z) has
// passed since its next x / z event was due (itself at most one absDelta after now), in which y fires at most
// (d + 1) * B * min(absDelta.x, absDelta.z) / absDelta.y + 1 times: the drop inside ring 0-1 uses d + 1 = 2, ring 2-3 4,
// ring 4-7 8, ring 8-15 16, the last ring the whole drop.  > 0: y less that drop clears every ring's gate (and stays on the
// torus: gates are >= 0).  The caller makes sure x, y and z are on the torus and that no deltaPos is NaN.
__device__ __forceinline__ int sky_drop_distance(const FrameArgs& A, int x, int y, int z, bool sxNeg, bool szNeg, float adx, float adz,
                                                 float dx, float dy, float dz, uint32_t left) {
    const uint32_t q = (sxNeg ? 1u : 0u) | (szNeg ? 2u : 0u);
    const uint4 g = __ldg(&A.sky_gate[(q * SKY_GRID + ((uint32_t)z >> A.sky_shift)) * SKY_GRID + ((uint32_t)x >> A.sky_shift)]);
    const float rate = FDIV(fabsf(dy), FADD(FADD(fabsf(dx), fabsf(dy)), fabsf(dz)));
    const float total = floorf(FMUL((float)(left + 3u), rate)) + 2.0f;
    // per block of horizontal travel; inf / NaN (an axis that never steps) fall back to the whole drop in fminf below
    const float slope = FMUL(fminf(adx, adz), fabsf(dy)) * (float)(1 << A.sky_shift);
    const int y1 = y + 1;
#ifdef RC_MUTATE_DROP  // mutation check (tools/mutation_check.sh): the drop up to the NEAR edge of each ring, no slack
    int d = y1 - (int)(g.x & 0xFFFFu);
    d = min(d, y1 - (int)fminf(total, floorf(slope * 2.0f)) - (int)(g.x >> 16));
    d = min(d, y1 - (int)fminf(total, floorf(slope * 4.0f)) - (int)(g.y & 0xFFFFu));
    d = min(d, y1 - (int)fminf(total, floorf(slope * 8.0f)) - (int)(g.y >> 16));
    d = min(d, y1 - (int)fminf(total, floorf(slope * 16.0f)) - (int)g.z);
#else
    int d = y1 - (int)fminf(total, floorf(slope * 2.0f) + 2.0f) - (int)(g.x & 0xFFFFu);
    if (d <= 0) return d;  // rays on their way into the terrain leave here
    d = min(d, y1 - (int)fminf(total, floorf(slope * 4.0f) + 2.0f) - (int)(g.x >> 16));
    d = min(d, y1 - (int)fminf(total, floorf(slope * 8.0f) + 2.0f) - (int)(g.y & 0xFFFFu));
    d = min(d, y1 - (int)fminf(total, floorf(slope * 16.0f) + 2.0f) - (int)(g.y >> 16));
    d = min(d, y1 - (int)total - (int)g.z);
#endif
    return d;
}

// The copy of the frame statistics this CTA adds to (FrameCounterBlock: spread over lines so that the atomics of 65 k warps
// do not queue on one).
__device__ __forceinline__ FrameCounters& counter_slot(const FrameArgs& A) {
    return A.counters->slot[(blockIdx.x + 13u * blockIdx.y) & (COUNTER_SLOTS - 1u)];
}

struct PixelResult {
    F3 color;
    rc_hit hit;
    bool traced;
    bool shadowRay;
    bool aoSeed;
    float sx_, sy_, sz_;  // un-biased exact at the hit (AO seed)
#ifdef RC_COST_MAP
    long long mark[3];    // developer probe: clock64 after loop A / after the replay / before the shadow ray (0: not reached)
#endif
};
#ifdef RC_COST_MAP
#define RC_MARK(i) out.mark[i] = clock64()
#else
#define RC_MARK(i)
#endif

// One pixel of low_res.frag main() (:256-393).
// Called by ALL 32 lanes of a warp (valid = the lane has a pixel): the primary loop agrees on its
// batch length with warp votes.
__device__ __forceinline__ void shade_pixel(const FrameArgs& A, bool valid, uint32_t px, uint32_t py, PixelResult& out,
                                            const uint4* top = nullptr, uint32_t ntop = 0u) {
    const rc_frame_params& P = A.p;
    const DeviceWorld& W = A.world;
    out.color = F3{0.0f, 0.0f, 0.0f};
    out.hit.x = out.hit.y = out.hit.z = 0;
    out.hit.steps = out.hit.shadow_steps = 0;
    out.hit.face = out.hit.kind = out.hit.flags = out.hit.ao_occluded = 0;
    out.traced = false;
    out.shadowRay = false;
    out.aoSeed = false;

    const float fragX = FADD((float)px, 0.5f), fragY = FADD((float)py, 0.5f);

    // :259-262
    bool live = valid;  // this lane still traces a ray
    if (live && (P.flags & RC_F_CROSSHAIR) && P.render_pos_data == 0) {
        const float mx = FSUB(fragX, P.mouse_pos[0]), my = FSUB(fragY, P.mouse_pos[1]);
        if (FSQRT(FADD(FMUL(mx, mx), FMUL(my, my))) <= 3.0f) {
            out.color = F3{1.0f, 1.0f, 1.0f};
            out.hit.kind = RC_KIND_CROSSHAIR;
            live = false;
        }
    }
    out.traced = live;

    // :264-265
    const float fcx = FMUL(fragX, A.inv_width);   // 1.0 / resolution, divided once on the host (IEEE fp32)
    const float fcy = FMUL(fragY, A.inv_height);

    const F3 origin = F3{P.origin[0], P.origin[1], P.origin[2]};
    const F3 camDir = F3{P.camera_dir[0], P.camera_dir[1], P.camera_dir[2]};
    const F3 sun = F3{P.sun_dir[0], P.sun_dir[1], P.sun_dir[2]};

    // :279-285
    // left = cross(cameraDir, (0,1,0)), up = cross(cameraDir, left): the same for every pixel, from the host (FrameArgs)
    const F3 left = F3{A.cam_left[0], A.cam_left[1], A.cam_left[2]};
    const F3 camUp = F3{A.cam_up[0], A.cam_up[1], A.cam_up[2]};
    const F3 t1 = scale3(left, -FMUL(P.proj_plane[0], FSUB(fcx, 0.5f)));
    const F3 t2 = scale3(scale3(camUp, FADD(-fcy, 0.5f)), P.proj_plane[1]);
    const F3 dir = normalize3(add3(add3(camDir, t1), t2));

    // :275,288-296
    Ray ray;
    build_ray(ray, dir.x, dir.y, dir.z, false);  // the six ratios only serve pos.exact: rays that hit divide for them below
    Pos pos;
    pos.x = (int)floorf(origin.x);
    pos.y = (int)floorf(origin.y);
    pos.z = (int)floorf(origin.z);
    pos.ex = ray.sx < 0 ? FSUB(origin.x, 1.0f) : origin.x;
    pos.ey = ray.sy < 0 ? FSUB(origin.y, 1.0f) : origin.y;
    pos.ez = ray.sz < 0 ? FSUB(origin.z, 1.0f) : origin.z;
    init_delta_pos(ray, pos);

    Cursor cur;
    cursor_reset(W, cur, top, ntop);
    CellCount cnt;
    int cellShift;
    int face = 0;
    F3 colorMod = F3{1.0f, 1.0f, 1.0f};
    bool reflected = false;
    float refrIndex = 1.0f;

    // :299-307
    RegionAt rg;
    int colNy;  // > 0: rg is the column above the terrain (enter_region_a), this many steps along y left in it
    enter_region_a(A, cur, rg, pos, ray.sy, colNy);
    uint32_t word = 0u;
    {
        const int idx = brick_bit(pos.x, pos.y, pos.z);
        if ((rg.r.mask >> idx) & 1ull) word = region_word(W, rg.r, idx);
    }
    if (live && word != 0u) {
        if (word_flags(word) & 0x4u) {
            refrIndex = 1.1f;
        } else {
            const uint4 pe = __ldg(&W.palette[word & MATERIAL_INDEX_MASK]);
            out.color = decode_color((unsigned long long)pe.x | ((unsigned long long)pe.y << 32));
            out.hit.kind = RC_KIND_CAMERA_INSIDE;
            out.hit.x = pos.x;
            out.hit.y = pos.y;
            out.hit.z = pos.z;
            live = false;
        }
    }

    // :309-315  loop A: to the first non-empty voxel.  pos.exact is not needed to FIND it, so this
    // pass leaves it alone (sky rays, ~300 steps each, never need it); rays that do hit replay their
    // steps below with pos.exact carried, in registers, without touching the tree.
    // Inside an empty 4^k cell a step is the DDA arithmetic plus one XOR test against the cell; the
    // tree is consulted only when the ray leaves the cell.  Inside a brick a step adds a bit test.
    uint32_t i = 0;
    uint32_t steps = 0;
    bool tailSkipped = false;
    const uint32_t numSteps = P.max_steps;
    {
        StepAxis ax;
        ax.x = ax.y = false;
        bool act = live && word == 0u && numSteps != 0u;
        bool freeFlight = false;
        while (act) {
            if (rg.r.mask == 0ull) {
                // Free flight: a ray that steps up and is above every occupied voxel of the columns it can still reach has
                // nothing left to hit: the remaining steps only move it, and run after this loop without cells, countdowns
                // or the tree.
                int gateDist = -0x10000;
                if (A.sky_gate && (uint32_t)pos.x <= W.wrap && (uint32_t)pos.z <= W.wrap) {
                    if (ray.sy > 0) {
                        gateDist = sky_gate_distance(A, pos.x, pos.y, pos.z, ray.sx < 0, ray.sz < 0, ray.adx, ray.adz, ray.dy,
                                                     pos.ty == pos.ty && pos.tz == pos.tz);
                        if (gateDist > 0) {
                            freeFlight = true;
                            break;
                        }
                    } else if (A.drop_gate && (uint32_t)pos.y <= W.wrap && pos.tx == pos.tx && pos.ty == pos.ty && pos.tz == pos.tz &&
                               sky_drop_distance(A, pos.x, pos.y, pos.z, ray.sx < 0, ray.sz < 0, ray.adx, ray.adz, ray.dx, ray.dy, ray.dz,
                                                 numSteps - steps) > 0) {
                        // a ray that steps down, but not as far down as anything it can still reach
                        freeFlight = true;
                        break;
                    }
                }
                // Empty cell.  The integer side of a step (position, cell test, step budget) is kept
                // as float countdowns -- voxels left to the cell's faces per axis, steps left -- so the
                // loop body is FADD/FMUL on the FMA pipes plus five compares, instead of integer/logic
                // work on the half-rate ALU pipe.  All values are small integers: exact in fp32.
                CellCount n;
                count_cell(n, ray, pos, rg.r.shift);
                if (colNy) n.ny = colNy;  // a column region: its extent along y is not the aligned cell's
                // a big cell that reaches above the gate: the countdown stops where the ray rises above it, the region is
                // entered again there and the test above fires (4^3 cells keep their full count: the sibling walk below
                // relies on it)
                if (rg.r.shift > 2) n.ny = min(n.ny, 1 - gateDist);
                float fnx = (float)n.nx, fny = (float)n.ny, fnz = (float)n.nz;
                const float rem0 = (float)(numSteps - steps);
                float rem = rem0;
                float tx = pos.tx, ty = pos.ty, tz = pos.tz;
                // An empty 4^3 cell is a child of the cursor's 16^3 node, whose mask says which of its 64 cells are empty:
                // when the ray crosses into a sibling cell that is empty too, the walk goes on without a region change
                // (no descent, no count_cell, no position recovery) -- five of a ray's nine region changes are of this kind.
                // The cell's coordinates in the node (0..3) are kept biased by 4 in byte fields: a step out of the node shows
                // up as a 3 or an 8 in its field and never borrows from the next one.
                const bool walk = rg.r.shift == 2 && colNy == 0;
                const uint32_t cell0 = 0x040404u + (uint32_t)((pos.x >> 2) & 3) + ((uint32_t)((pos.y >> 2) & 3) << 8) +
                                       ((uint32_t)((pos.z >> 2) & 3) << 16);
                uint32_t cell = cell0;
                bool outOfSteps = false;
                for (;;) {
                    do {
                        ax.x = tx < ty && tx < tz;
                        ax.y = !ax.x && ty < tz;
                        const bool az = !ax.x && !ax.y;
                        if (ax.x) { tx = FADD(tx, ray.adx); fnx -= 1.0f; }
                        if (ax.y) { ty = FADD(ty, ray.ady); fny -= 1.0f; }
                        if (az) { tz = FADD(tz, ray.adz); fnz -= 1.0f; }
                        rem -= 1.0f;
                    } while ((fnx * fny) * (fnz * rem) != 0.0f);
                    // keeps the compiler from folding the code below into the 16-instruction loop above (it would carry
                    // pre-decrement copies of the countdowns through it: 22 instructions per step)
                    asm volatile("" : "+f"(fnx), "+f"(fny), "+f"(fnz), "+f"(rem));
                    if ((fnx * fny) * fnz != 0.0f) {  // out of steps inside the empty cell
                        outOfSteps = true;
                        break;
                    }
                    if (!walk) break;
                    // the axis that stepped last reached its cell face: the ray is in the next cell along it
                    const bool wz = !ax.x && !ax.y;  // the loop's own three predicates
                    if (ax.x) { cell += (uint32_t)ray.sx; fnx = 4.0f; }
                    if (ax.y) { cell += (uint32_t)ray.sy << 8; fny = 4.0f; }
                    if (wz) { cell += (uint32_t)ray.sz << 16; fnz = 4.0f; }
                    if (rem == 0.0f) break;                             // that was the last step of the budget
                    if ((cell & 0x0C0C0Cu) != 0x040404u) break;         // left the 16^3 node
                    const uint32_t ci = (cell & 3u) | ((cell >> 6) & 0xCu) | ((cell >> 12) & 0x30u);
                    if ((cur.mask >> ci) & 1ull) break;                 // the next cell is not empty
                }
                pos.tx = tx;
                pos.ty = ty;
                pos.tz = tz;
                // steps along an axis = first countdown + 4 per cell crossed - what is left of the current one
                pos.x += ray.sx * (n.nx - (int)fnx) + 4 * ((int)(cell & 0xFFu) - (int)(cell0 & 0xFFu));
                pos.y += ray.sy * (n.ny - (int)fny) + 4 * ((int)((cell >> 8) & 0xFFu) - (int)((cell0 >> 8) & 0xFFu));
                pos.z += ray.sz * (n.nz - (int)fnz) + 4 * ((int)((cell >> 16) & 0xFFu) - (int)((cell0 >> 16) & 0xFFu));
                steps += (uint32_t)(rem0 - rem);
                if (outOfSteps) break;
                enter_region_a(A, cur, rg, pos, ray.sy, colNy);
            } else {
                // Brick (or uniform solid cell): step until the ray leaves it, enters an occupied voxel or runs out of
                // steps -- all inside this one pass of the outer loop, so that a lane walking a brick does not make the
                // rest of the warp (empty-cell lanes wait through every pass) pay a pass per step.
                bool leftBrick;
                do {
                    ax = dda_step_bl(ray, pos);
                    steps++;
                    leftBrick = left_region(rg, pos);
                } while (!leftBrick && !((rg.r.mask >> brick_bit(pos.x, pos.y, pos.z)) & 1ull) && steps < numSteps);
                if (leftBrick) enter_region_a(A, cur, rg, pos, ray.sy, colNy);
            }
            if (rg.r.mask != 0ull) {
                const int idx = brick_bit(pos.x, pos.y, pos.z);
                if ((rg.r.mask >> idx) & 1ull) word = region_word(W, rg.r, idx);
            }
            act = word == 0u && steps < numSteps;
        }
        if (freeFlight) {
            // same comparisons, same additions (nextIntersectDDA, low_res.frag:446-491); nothing else per step
            {
                const unsigned m = __activemask();
                if ((int)(threadIdx.x & 31u) == __ffs(m) - 1) atomicAdd(&counter_slot(A).free_flight_rays, (unsigned)__popc(m));
            }
            if (A.sky_tail) {
                // free_flight.h: most of the remaining steps in closed form, the last few walked; the voxel the ray ends in
                int nx, ny, nz;
                free_flight(pos.tx, pos.ty, pos.tz, ray.adx, ray.ady, ray.adz, ray.dx, ray.dy, ray.dz, numSteps - steps,
                            !(P.flags & RC_F_NO_LEAP), nx, ny, nz, ax.x, ax.y);
                pos.x += ray.sx * nx;
                pos.y += ray.sy * ny;
                pos.z += ray.sz * nz;
            } else {
                // Nobody will see where this ray ends: no hit records, no position debug output, and the looked-at voxel is
                // not max_steps steps from the camera's (every step moves one voxel along one axis, the signs never
                // change in loop A), so the highlight test below cannot fire.  The pixel is sky either way.
                tailSkipped = true;
            }
            steps = numSteps;
        }
        if (!live) return;
        // the shader's counter: `while (empty && i++ < numSteps)` leaves numSteps + 1 behind when it runs out
        i = word != 0u ? steps : numSteps + 1u;
        if (steps) face = axis_index(ax);
    }
    RC_MARK(0);
    if (word != 0u || P.render_pos_data == 1) {
        // ratios columns (low_res.frag:75-91) for pos.exact, now that the ray is known to need it
        ray.r0y = FDIV(ray.dy, ray.dx);
        ray.r0z = FDIV(ray.dz, ray.dx);
        ray.r1x = FDIV(ray.dx, ray.dy);
        ray.r1z = FDIV(ray.dz, ray.dy);
        ray.r2x = FDIV(ray.dx, ray.dz);
        ray.r2y = FDIV(ray.dy, ray.dz);
    }
    if (steps != 0u && (word != 0u || P.render_pos_data == 1)) {
        // replay: same ray, same comparisons, now with pos.exact (low_res.frag:447,489)
        Pos rp;
        rp.x = (int)floorf(origin.x);
        rp.y = (int)floorf(origin.y);
        rp.z = (int)floorf(origin.z);
        rp.ex = pos.ex;
        rp.ey = pos.ey;
        rp.ez = pos.ez;
        init_delta_pos(ray, rp);
        for (uint32_t n = 0; n < steps; n++) dda_step_exact_bl(ray, rp);
        pos.ex = rp.ex;
        pos.ey = rp.ey;
        pos.ez = rp.ez;
    }
    cnt.nx = cnt.ny = cnt.nz = 1;
    cellShift = 0;
    RC_MARK(1);
    // :317-333  loop B
    while (i++ < numSteps) {
        if (word != 0u) {
            const uint32_t fl = word_flags(word);
            if ((fl & 0x7u) == 0x3u) {
                // reflectRay :178-194
                if (face == 0) {
                    pos.tx = FSUB(pos.tx, ray.adx);
                    ray.sx = -ray.sx;
                    ray.dx = -ray.dx;
                } else if (face == 1) {
                    pos.ty = FSUB(pos.ty, ray.ady);
                    ray.sy = -ray.sy;
                    ray.dy = -ray.dy;
                } else {
                    pos.tz = FSUB(pos.tz, ray.adz);
                    ray.sz = -ray.sz;
                    ray.dz = -ray.dz;
                }
                colorMod = scale3(colorMod, 0.94f);
                reflected = true;
            } else if ((fl & 0x7u) == 0x5u) {
                // refractRay :212-240
                const bool liquid = (fl & 0x10u) != 0u;
                if (liquid) {
                    colorMod.x = FMUL(colorMod.x, 0.94f);
                    colorMod.y = FMUL(colorMod.y, 0.97f);
                    colorMod.z = FMUL(colorMod.z, 1.0f);
                } else {
                    colorMod = scale3(colorMod, 0.95f);
                }
                if (refrIndex != 1.1f) {
                    if (face != 0 && ray.sx < 0) pos.ex = FADD(pos.ex, 1.0f);
                    if (face != 1 && ray.sy < 0) pos.ey = FADD(pos.ey, 1.0f);
                    if (face != 2 && ray.sz < 0) pos.ez = FADD(pos.ez, 1.0f);
                    F3 nrm = F3{0.0f, 0.0f, 0.0f};
                    if (face == 0) nrm.x = (float)ray.sx;
                    else if (face == 1) nrm.y = (float)ray.sy;
                    else nrm.z = (float)ray.sz;
                    if (liquid) {
                        const float arg = FMUL(FSUB(FADD(P.delta_time, FMUL(pos.ex, 0.2f)), FMUL(pos.ez, 0.1f)), 10.0f);
                        nrm.x = FADD(nrm.x, FMUL(det_sin(arg), 0.2f));
                        nrm = normalize3(nrm);
                    }
                    // refractRay(vec3,vec3,float,float) :198-210
                    const F3 rd = F3{ray.dx, ray.dy, ray.dz};
                    const float rr = FDIV(refrIndex, 1.1f);
                    float c1 = dot3(nrm, rd);
                    if (c1 < 0.0f) {
                        nrm = F3{-nrm.x, -nrm.y, -nrm.z};
                        c1 = dot3(nrm, rd);
                    }
                    const float c2 = FSQRT(FSUB(1.0f, FMUL(FMUL(rr, rr), FSUB(1.0f, FMUL(c1, c1)))));
                    const float k2 = FSUB(FMUL(rr, c1), c2);
                    const F3 nd = add3(scale3(rd, rr), scale3(nrm, k2));
                    build_ray(ray, nd.x, nd.y, nd.z, true);
                    pos.ex = FADD(pos.ex, ray.sx < 0 ? -1.0f : 0.0f);
                    pos.ey = FADD(pos.ey, ray.sy < 0 ? -1.0f : 0.0f);
                    pos.ez = FADD(pos.ez, ray.sz < 0 ? -1.0f : 0.0f);
                    init_delta_pos(ray, pos);
                    refrIndex = 1.1f;
                }
            } else {
                break;
            }
            cnt.nx = cnt.ny = cnt.nz = 1;  // step signs may have changed: re-derive after the next lookup
        }
        bool leftCell;
        face = dda_step<true>(ray, pos, cnt, leftCell);
        steps++;
        if (leftCell) {
            word = lookup(W, cur, pos.x, pos.y, pos.z, cellShift);
            after_lookup(cnt, ray, pos, word, cellShift);
        }
    }

    out.hit.x = pos.x;
    out.hit.y = pos.y;
    out.hit.z = pos.z;
    out.hit.steps = (uint16_t)steps;
    out.hit.face = (uint8_t)face;
    out.hit.flags = (uint8_t)((reflected ? RC_HF_REFLECTED : 0u) | (refrIndex != 1.0f ? RC_HF_REFRACTED : 0u) |
                              (ray.sx < 0 ? RC_HF_STEPX_NEG : 0u) | (ray.sy < 0 ? RC_HF_STEPY_NEG : 0u) |
                              (ray.sz < 0 ? RC_HF_STEPZ_NEG : 0u));

    // :335-341
    if (P.render_pos_data == 1) {
        out.color = F3{pos.ex, pos.ey, pos.ez};
        out.hit.kind = RC_KIND_DEBUG;
        return;
    } else if (P.render_pos_data == 2) {
        out.color = F3{(float)pos.x, (float)pos.y, (float)pos.z};
        out.hit.kind = RC_KIND_DEBUG;
        return;
    }

    // :343-345
    if (face != 0 && ray.sx < 0) pos.ex = FADD(pos.ex, 1.0f);
    if (face != 1 && ray.sy < 0) pos.ey = FADD(pos.ey, 1.0f);
    if (face != 2 && ray.sz < 0) pos.ez = FADD(pos.ez, 1.0f);

    unsigned long long packedColor = ~0ull;
    if (word != 0u) {
        const uint4 pe = __ldg(&W.palette[word & MATERIAL_INDEX_MASK]);
        packedColor = (unsigned long long)pe.x | ((unsigned long long)pe.y << 32);
    }

    // :347-350
    if (!tailSkipped && P.looking_at[0] == pos.x && P.looking_at[1] == pos.y && P.looking_at[2] == pos.z) {
        const F3 c = decode_color(packedColor);
        out.color = F3{FADD(FMUL(c.x, 2.0f), 0.3f), FADD(FMUL(c.y, 2.0f), 0.3f), FADD(FMUL(c.z, 2.0f), 0.3f)};
        out.hit.kind = RC_KIND_HIGHLIGHT;
        return;
    }

    // :352-355
    if (word == 0u) {
        const F3 sky = gen_sky(ray.dx, ray.dy, ray.dz, sun);
        out.color = F3{FMUL(sky.x, colorMod.x), FMUL(sky.y, colorMod.y), FMUL(sky.z, colorMod.z)};
        out.hit.kind = RC_KIND_SKY;
        return;
    }

    // :357-371
    out.hit.kind = RC_KIND_HIT;
    const F3 color = decode_color(packedColor);
    const float sunA = sel3(face, sun.x, sun.y, sun.z);
    const int stepA = sel3i(face, ray.sx, ray.sy, ray.sz);
    const float facingTerm = FMUL(sunA, (float)(-stepA));
    const bool facing = facingTerm > 0.0f;
    {
        const float mod = facing ? 0.15f : 0.0f;
        const float intensity = glsl_min(FADD(FADD(glsl_max(0.0f, facingTerm), 0.4f), mod), 1.0f);
        out.color = F3{FMUL(FMUL(color.x, intensity), colorMod.x), FMUL(FMUL(color.y, intensity), colorMod.y),
                       FMUL(FMUL(color.z, intensity), colorMod.z)};
    }
    const F3 dim = F3{FMUL(FMUL(color.x, 0.3f), colorMod.x), FMUL(FMUL(color.y, 0.3f), colorMod.y),
                      FMUL(FMUL(color.z, 0.3f), colorMod.z)};
    out.aoSeed = P.ao_samples > 0;
    out.sx_ = pos.ex;
    out.sy_ = pos.ey;
    out.sz_ = pos.ez;
    // normal sign for AO = -step[face] of the primary ray
    out.hit.ao_occluded = 0;

    bool done = reflected;
    if (!done && !facing) {
        out.color = dim;
        done = true;
    }
    if (!done) out.hit.flags |= RC_HF_FACING_SUN;

    RC_MARK(2);
    // :373-391 shadow ray
    if (!done && (P.flags & RC_F_SHADOWS)) {
        out.shadowRay = true;
        Ray sr;  // buildRay(sunDir): the same for every pixel, its three divisions are done once on the host (FrameArgs::sun_inv)
        sr.dx = sun.x; sr.dy = sun.y; sr.dz = sun.z;
        sr.sx = sun.x < 0.0f ? -1 : 1; sr.sy = sun.y < 0.0f ? -1 : 1; sr.sz = sun.z < 0.0f ? -1 : 1;
        sr.ddx = A.sun_inv[0]; sr.ddy = A.sun_inv[1]; sr.ddz = A.sun_inv[2];
        sr.adx = fabsf(sr.ddx); sr.ady = fabsf(sr.ddy); sr.adz = fabsf(sr.ddz);
        Pos sp;
        sp.x = pos.x;
        sp.y = pos.y;
        sp.z = pos.z;
        sp.ex = sr.sx < 0 ? FSUB(pos.ex, 1.0f) : pos.ex;
        sp.ey = sr.sy < 0 ? FSUB(pos.ey, 1.0f) : pos.ey;
        sp.ez = sr.sz < 0 ? FSUB(pos.ez, 1.0f) : pos.ez;
        init_delta_pos(sr, sp);
        if (face == 0) sp.tx = FSUB(sp.tx, sr.adx);
        else if (face == 1) sp.ty = FSUB(sp.ty, sr.ady);
        else sp.tz = FSUB(sp.tz, sr.adz);
        uint32_t k = P.shadow_steps;
        uint32_t state = 0u;  // block = (-1,-1): "empty"; 1 = LIQUID (the ray goes on), 2 = opaque
        uint32_t ssteps = 0;
        enter_region(W, cur, rg, sp);
        // the sun is up: a shadow ray that has risen above everything it can still reach (sky gates, as in loop A) stays
        // unoccluded and runs out of steps -- asked when it enters an empty region and every 8th step inside one
        const bool gated = A.shadow_gate && sr.sy > 0 && sp.ty == sp.ty && sp.tz == sp.tz;
        while (state != 2u && k-- != 0u) {
            dda_step_bl(sr, sp);
            ssteps++;
            const bool left = left_region(rg, sp);
            if (left) enter_region(W, cur, rg, sp);
            state = occlusion_state(W, rg.r, sp.x, sp.y, sp.z);
            if (gated && rg.r.mask == 0ull && (left || (k & 7u) == 0u) && (uint32_t)sp.x <= W.wrap && (uint32_t)sp.z <= W.wrap &&
                (uint32_t)sp.y + k <= W.wrap &&
                sky_gate_distance(A, sp.x, sp.y, sp.z, sr.sx < 0, sr.sz < 0, sr.adx, sr.adz, sr.dy, true) > 0) {
                ssteps += k;  // the steps it would still take, all empty
                break;
            }
        }
        const uint32_t sw = state;
        out.hit.shadow_steps = (uint16_t)ssteps;
        if (sw != 0u) {
            out.color = dim;
            out.hit.flags |= RC_HF_SHADOWED;
        }
    }
}

// Pixel (px,py) of the frame -> index into this rank's compact buffers, or -1 if
// the row belongs to another rank.
__device__ __forceinline__ uint32_t local_index(const FrameArgs& A, uint32_t px, uint32_t localRow) {
    return localRow * A.p.width + px;
}

// idx: this rank's compact buffers; framePixel: the pixel's place in the whole frame, for the direct
// multi-GPU frame (rc_frame_target_*: a peer-mapped buffer on the composing GPU, written over NVLink)
__device__ __forceinline__ uint32_t frame_pixel_of(const FrameArgs& A, uint32_t px, uint32_t localRow) {
    const uint32_t row = ((localRow / BAND_ROWS) * A.world_size + A.rank) * BAND_ROWS + (localRow % BAND_ROWS);
    return row * A.p.width + px;
}
__device__ __forceinline__ void store_pixel(const FrameArgs& A, uint32_t idx, uint32_t framePixel, F3 c) {
    const uchar4 v = make_uchar4(to_unorm8(c.x), to_unorm8(c.y), to_unorm8(c.z), 0);
#ifdef RC_CHECKED
    // a pixel outside this rank's buffer, or a peer store outside the composing GPU's frame
    if (idx >= A.local_rows * A.p.width || (A.frame_target && framePixel >= A.p.width * A.p.height)) __trap();
#endif
    A.rgba8[idx] = v;
    if (A.frame_target) A.frame_target[framePixel] = v;
    if (A.rgba32f) A.rgba32f[idx] = make_float4(c.x, c.y, c.z, 0.0f);
}

// ---------------------------------------------------------------------------
// AO kernel (semantics = oracle castAo): one thread per hit pixel,
// all lanes of a warp walk the same table entry at the same time, so the
// direction comes from __constant__ memory as a broadcast.
//
// The 64 rays of a pixel take <= ao_steps steps each from the SAME voxel: together
// they visit at most a few dozen 4^3 bricks, over and over.  So the tree is not
// walked per step: each thread keeps the occupancy of the brick it is in
// (64-bit mask) in registers, and the bricks it has seen in a private,
// direct-mapped cache in shared memory ([slot][thread] columns: bank = lane, no
// conflicts whatever the slots).  A step inside the current brick costs a bit
// test; a step into a cached brick one 16-byte shared-memory read; only a miss
// descends the packed tree (from the deepest cached node that contains it).
// The DDA arithmetic is the shader's, operation for operation.
// ---------------------------------------------------------------------------
constexpr int AO_THREADS = 128;
constexpr int AO_SLOTS = 16;
constexpr uint32_t AO_KEY_MASK = 0x3FFFFFFFu;   // 10 bits of brick coordinate per axis (max_depth <= 7)

__global__ void __launch_bounds__(AO_THREADS) ao_cached_kernel(const __grid_constant__ FrameArgs A) {
    __shared__ uint32_t sInfo[AO_SLOTS][AO_THREADS];
    __shared__ uint32_t sBase[AO_SLOTS][AO_THREADS];
    __shared__ uint2 sMask[AO_SLOTS][AO_THREADS];

    const uint32_t nSeeds = A.counters->seed_cursor.next;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nSeeds) return;
    const uint32_t tid = threadIdx.x;
    const rc_frame_params& P = A.p;
    const DeviceWorld& W = A.world;
    const AoSeed s = A.ao_seeds[t];
    const int face = (int)(s.face_sign & 3u);
    const bool negNormal = (s.face_sign & 4u) != 0u;
    const bool f0 = face == 0, f1 = face == 1;
    const bool cacheable = W.wrap <= 4095u;  // brick coordinates fit 10 bits per axis

    // Empty slots hold a key that can never be looked up there: brick (1023,1023,1023) maps to slot
    // 15, brick (0,0,0) to slot 0.
#pragma unroll
    for (int k = 0; k < AO_SLOTS; k++) sInfo[k][tid] = (k == AO_SLOTS - 1) ? 0u : 0xFFFFFFFFu;

    Cursor cur;
    cursor_reset(W, cur);
    Region br;
    br.mask = 0ull;
    br.base = 0u;
    br.info = 0u;
    br.shift = 2;
    int cx = s.x ^ 0x40000000, cy = s.y, cz = s.z;  // a voxel of the region held in br: none yet (far away)

    const float fx = (float)s.x, fy = (float)s.y, fz = (float)s.z;
    uint32_t occluded = 0;
    const uint32_t nSamples = P.ao_samples, nSteps = P.ao_steps;
    for (uint32_t k = 0; k < nSamples; k++) {
        const float tx = c_hemi[3 * k + 0], ty = c_hemi[3 * k + 1], up = c_hemi[3 * k + 2];
        const float itx = c_hemi_inv[3 * k + 0], ity = c_hemi_inv[3 * k + 1], iup = c_hemi_inv[3 * k + 2];
        // d[face] = sg*up, d[(face+1)%3] = tx, d[(face+2)%3] = ty ; 1/(-up) == -(1/up) exactly
        const float su = negNormal ? -up : up, siu = negNormal ? -iup : iup;
        const float dx = f0 ? su : (f1 ? ty : tx), ix = f0 ? siu : (f1 ? ity : itx);
        const float dy = f1 ? su : (f0 ? tx : ty), iy = f1 ? siu : (f0 ? itx : ity);
        const float dz = f0 ? ty : (f1 ? tx : su), iz = f0 ? ity : (f1 ? itx : siu);
        const bool nx = dx < 0.0f, ny = dy < 0.0f, nz = dz < 0.0f;
        const int sx = nx ? -1 : 1, sy = ny ? -1 : 1, sz = nz ? -1 : 1;
        const float adx = fabsf(ix), ady = fabsf(iy), adz = fabsf(iz);
        // pos.exact re-biased for the new signs; deltaPos = absDelta - (exact - round) * delta
        const float ex = nx ? FSUB(s.ex, 1.0f) : s.ex;
        const float ey = ny ? FSUB(s.ey, 1.0f) : s.ey;
        const float ez = nz ? FSUB(s.ez, 1.0f) : s.ez;
        float px = FSUB(adx, FMUL(FSUB(ex, fx), ix));
        float py = FSUB(ady, FMUL(FSUB(ey, fy), iy));
        float pz = FSUB(adz, FMUL(FSUB(ez, fz), iz));
        if (f0) px = FSUB(px, adx);
        else if (f1) py = FSUB(py, ady);
        else pz = FSUB(pz, adz);
        int x = s.x, y = s.y, z = s.z;

        // while ((block empty || LIQUID) && i-- != 0) { DDA step; getBlock }  -- state: 0 empty, 1 liquid, 2 opaque
        uint32_t state = 0u;
        uint32_t i = nSteps;
        while (state != 2u && i-- != 0u) {
            // nextIntersectDDA (low_res.frag:446-491), branch-free
            const bool bx = px < py && px < pz;
            const bool by = !bx && py < pz;
            const bool bz = !bx && !by;
            if (bx) { x += sx; px = FADD(px, adx); }
            if (by) { y += sy; py = FADD(py, ady); }
            if (bz) { z += sz; pz = FADD(pz, adz); }

            if ((((x ^ cx) | (y ^ cy) | (z ^ cz)) >> br.shift) != 0) {
                cx = x;
                cy = y;
                cz = z;
                const int wx = wrap_coord(W, x), wy = wrap_coord(W, y), wz = wrap_coord(W, z);
                const uint32_t bxk = (uint32_t)wx >> 2, byk = (uint32_t)wy >> 2, bzk = (uint32_t)wz >> 2;
                const uint32_t key = bxk | (byk << 10) | (bzk << 20);
                const uint32_t slot = (bxk & 1u) | ((byk & 1u) << 1) | (bzk & 3u) << 2;
                const uint32_t info = sInfo[slot][tid];
                if (cacheable && (info & AO_KEY_MASK) == key) {
                    const uint2 m = sMask[slot][tid];
                    br.mask = (unsigned long long)m.x | ((unsigned long long)m.y << 32);
                    br.base = sBase[slot][tid];
                    br.info = info;
                    br.shift = 2;
                } else {
                    fetch_region(W, cur, wx, wy, wz, br);
                    br.info |= key;
                    if (cacheable) {
                        sInfo[slot][tid] = br.info;
                        sBase[slot][tid] = br.base;
                        sMask[slot][tid] = make_uint2((uint32_t)br.mask, (uint32_t)(br.mask >> 32));
                    }
                }
            }
            state = occlusion_state(W, br, x, y, z);
        }
        if (state != 0u) occluded++;
    }
    const float ao = FSUB(1.0f, FMUL(P.ao_strength, FDIV((float)occluded, (float)P.ao_samples)));
    store_pixel(A, s.pixel, s.frame_pixel, F3{FMUL(s.r, ao), FMUL(s.g, ao), FMUL(s.b, ao)});
    if (A.hits) A.hits[s.pixel].ao_occluded = (uint8_t)occluded;
}

// ---------------------------------------------------------------------------
// Window AO: ao_steps == AOW_STEPS (the reference's 5, light_scattering.frag:231).
//
// The first step of an AO ray leaves the hit voxel through the hit face (deltaPos[face] was lowered by absDelta[face]);
// after that its normal coordinate never comes back.  So all rays of a pixel stay inside a box of 2 bricks along the normal
// and 3 bricks along each tangent axis: 12 x 12 x 8 voxels.  The warp copies the occupancy of those 18 bricks out of the
// tree into a dense window in shared memory -- ONE window per distinct (hit voxel, face, normal sign) among its 32 pixels,
// filled cooperatively -- laid out so that a DDA step is one add on the index (q += +-1 | +-strideY | +-strideZ) and the
// emptiness test one shared-memory byte.  The rays then run without touching the tree.
//   face x: strides 1, 8, 96   (8 x 12 x 12 voxels)
//   face y: strides 1, 16, 128 (12 x 8 x 12, rows padded to 16)
//   face z: strides 1, 16, 192 (12 x 12 x 8, rows padded to 16)
// A 4-voxel x-row of a brick lands at a multiple of 4 and never straddles a word.
//
// Where a pixel's rays start.  The primary ray leaves pos.exact exactly on the hit face along the face axis (stepping an
// axis snaps that coordinate to an integer, and the shader un-biases only the other two: low_res.frag:343-345), so
// (exact - round)[face] = 0 for a positive AO step and -1 for a negative one, in exact arithmetic:
//   normal -  (primary step +): deltaPos[face] = absDelta - (-1)(-absDelta) - absDelta = -absDelta: the first step is through
//             the face whatever the tangents do; it leaves deltaPos[face] = 0, and with both tangent deltaPos > 0 the
//             SECOND step is along the normal too (0 wins every comparison of low_res.frag:450-488 for its own axis and
//             loses none), leaving deltaPos[face] = absDelta.  Rays start 2 voxels out with 3 free steps.
//   normal +  (primary step -): deltaPos[face] = absDelta - 0 - absDelta = 0: the first step is through the face when both
//             tangent deltaPos > 0, leaving deltaPos[face] = absDelta.  Rays start 1 voxel out with 4 free steps.
// Tangent deltaPos = absDelta - (exact - round) * delta > 0 strictly when (exact - round) lies in [2^-20, 1 - 2^-20] (for
// the step sign used): the rounded product stays below absDelta (1 - 2^-21).  A pixel that meets all of this and has no
// LIQUID in its window (rays pass through liquid, low_res.frag:384, a bit cannot say that) is FAST: every sample starts
// from the same voxel with the same deltaPos[face] = absDelta -- obtained from the common formula with (exact - round)[face]
// set to 0 -- and the occupancy of the one or two voxels already crossed is read once per pixel.  Everything else (a
// pixel exactly on a voxel edge, pos.exact drifted by a mirror bounce, liquid nearby) is cast on the tree itself, the
// pixel's samples dealt to the 32 lanes of the warp (ao_ray_general).
// ---------------------------------------------------------------------------
constexpr uint32_t AOW_STEPS = 5;
constexpr int AOW_WORDS = 48;
constexpr uint32_t AOW_BYTE_SLOTS = 4;    // windows per warp in byte mode ...
constexpr uint32_t AOW_BYTE_WORDS = 384;  // ... of 16 x 12 x 8 bytes: 4 * 384 = 32 * 48 words, the warp's share of window memory
constexpr uint32_t AOW_WARP_WORDS = 32u * AOW_WORDS;  // a warp's share of window memory

// One table entry per (orientation = face*2 + negative normal, sample): what buildRay derives
// from the direction, precomputed on the host with the same IEEE operations.
struct AoDir {
    float ix, iy, iz;   // ray.delta = 1 / dir
    float qx, qy, qz;   // window stride of one step along each axis, sign(dir) * stride(face), as the fp32 DENORMAL whose
                        // bit pattern is that integer: denormals add like integers, and the sum's bits are the index itself
    float lx, ly;       // absDelta[face] when the face is x (lx) or y (ly), else 0: what deltaPos[face] is lowered by
};
static_assert(sizeof(AoDir) == 32, "AoDir is two 128-bit loads");
#ifdef RC_AO_TABLE_CONSTANT
// A/B variant (tools/build_variants.sh): the oriented table in __constant__ memory.  The entry a lane needs
// depends on ITS pixel's face and normal sign, so a warp whose pixels see more than one orientation reads several
// addresses per load -- which the constant cache serialises, where the L1 serves a 128-bit __ldg of any 32 addresses at once.
__constant__ AoDir c_aoTable[6 * RC_MAX_AO_SAMPLES];
#define AO_DIR_V0(dirs, orient, k) (*reinterpret_cast<const float4*>(&c_aoTable[(orient) * RC_MAX_AO_SAMPLES + (k)]))
#define AO_DIR_FIRST(dirs, orient, k) (c_aoTable[(orient) * RC_MAX_AO_SAMPLES + (k)])
#else
#define AO_DIR_V0(dirs, orient, k) __ldg(reinterpret_cast<const float4*>((dirs) + (k)))
#define AO_DIR_FIRST(dirs, orient, k) ((dirs)[k])
#endif

__device__ __forceinline__ int aow_stride_y(int face) { return face == 0 ? 8 : 16; }
__device__ __forceinline__ int aow_stride_z(int face) { return face == 0 ? 96 : (face == 1 ? 128 : 192); }

// The shader's loop for one AO ray on the tree itself (no window): semantics of castAo.
// (x,y,z) / (px,py,pz) = the ray's state before its first step.
__device__ __noinline__ uint32_t ao_ray_general(const DeviceWorld& W, int hx, int hy, int hz, float px, float py, float pz,
                                                float adx, float ady, float adz, int sx, int sy, int sz, uint32_t nSteps) {
    Cursor cur;
    cursor_reset(W, cur);
    RegionAt g;
    Pos p;
    p.x = hx; p.y = hy; p.z = hz;
    p.tx = px; p.ty = py; p.tz = pz;
    p.ex = p.ey = p.ez = 0.0f;
    Ray r;
    r.sx = sx; r.sy = sy; r.sz = sz;
    r.adx = adx; r.ady = ady; r.adz = adz;
    enter_region(W, cur, g, p);
    uint32_t state = 0u, i = nSteps;
    while (state != 2u && i-- != 0u) {
        dda_step_bl(r, p);
        if (left_region(g, p)) enter_region(W, cur, g, p);
        state = occlusion_state(W, g.r, p.x, p.y, p.z);
    }
    return state != 0u ? 1u : 0u;
}

// Occupancy of window voxel q.  BYTES: one byte per voxel, q is the shared-memory byte address itself, result 0 / 1.  Else one
// bit per voxel, MSB-first inside each byte, q = 8 * address + bit number: the byte, replicated to the four byte lanes by one
// multiply, is shifted left by q & 31 = 8 k + (q & 7), which leaves bit 7 - (q & 7) in the sign.
template <bool BYTES>
__device__ __forceinline__ uint32_t aow_occ(uint32_t q) {
    uint32_t byte;
    if (BYTES) {
        asm volatile("ld.shared.u8 %0, [%1];" : "=r"(byte) : "r"(q) : "memory");
        return byte;
    }
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(byte) : "r"(q >> 3) : "memory");
    return (byte * 0x01010101u) << (q & 31u);
}

// The rays of one FAST pixel (see above): from window index qSf (an fp32 DENORMAL whose bit pattern is the index:
// denormals add like integers, so a step is a predicated FADD on the FMA pipe instead of select + integer add on the
// half-rate ALU pipe), acc0 = occupancy of the voxels already crossed, three free steps -- four when FOUR (some pixel of the
// warp has a positive normal) and this lane's pixel is one of those (`fourth`; the others stand still in that step).
// No early exit on a hit: the warp runs until its last ray is done anyway, a ray that has hit stays inside the window, and
// without the branch the load of one step overlaps the compares of the next.
// move = false: the lane has no fast pixel; it walks along without going anywhere (strides 0).
// skip: bit g set = no voxel that a ray of sample group g can still enter is occupied, for any pixel of the warp
// (AO_OPT_OCTANT): those rays are not cast; they are occluded iff one of the voxels already crossed was.
template <bool BYTES, bool FOUR>
__device__ __forceinline__ uint32_t aow_cast_rays(const AoDir* __restrict__ dirs, uint32_t orient, float qSf, uint32_t acc0, float a0x, float a1x,
                                                  float a0y, float a1y, float a0z, float a1z, bool move, bool fourth, uint32_t skip) {
    uint32_t occluded = 0;
    for (uint32_t g = 0; g < 4u; g++) {
        const uint32_t kBegin = c_aoSeg[g], kEnd = c_aoSeg[g + 1u];
        if (kBegin == kEnd) continue;
        if ((skip >> g) & 1u) {
            occluded += (BYTES ? acc0 : acc0 >> 31) * (kEnd - kBegin);
            continue;
        }
        // every sample of the group steps the same way along each axis: the strides and (exact - round) are the group's
        const AoDir first = AO_DIR_FIRST(dirs, orient, kBegin);
        const float qx = move ? first.qx : 0.0f, qy = move ? first.qy : 0.0f, qz = move ? first.qz : 0.0f;
        const float q4x = fourth ? qx : 0.0f, q4y = fourth ? qy : 0.0f, q4z = fourth ? qz : 0.0f;
        const float ax = first.qx < 0 ? a1x : a0x, ay = first.qy < 0 ? a1y : a0y, az = first.qz < 0 ? a1z : a0z;
        for (uint32_t k = kBegin; k < kEnd; k++) {
            const float4 v0 = AO_DIR_V0(dirs, orient, k);
            const float ix = v0.x, iy = v0.y, iz = v0.z;  // ray.delta = 1 / dir, |.| = absDelta
            const float adx = fabsf(ix), ady = fabsf(iy), adz = fabsf(iz);
            // deltaPos = absDelta - (exact - round) * delta (low_res.frag:379); (exact - round)[face] comes in as 0
            float px = FSUB(adx, FMUL(ax, ix));
            float py = FSUB(ady, FMUL(ay, iy));
            float pz = FSUB(adz, FMUL(az, iz));
            float qf = qSf;
            uint32_t acc = acc0;
#pragma unroll
            for (uint32_t i = 0; i < 3u; i++) {
                const bool bx = px < py && px < pz;
                const bool by = !bx && py < pz;
                const bool bz = !bx && !by;
                if (bx) { qf += qx; px = FADD(px, adx); }
                if (by) { qf += qy; py = FADD(py, ady); }
                if (bz) { qf += qz; pz = FADD(pz, adz); }
                acc |= aow_occ<BYTES>(__float_as_uint(qf));
            }
            if (FOUR) {
                const bool bx = px < py && px < pz;
                const bool by = !bx && py < pz;
                const bool bz = !bx && !by;
                if (bx) qf += q4x;
                if (by) qf += q4y;
                if (bz) qf += q4z;
                acc |= aow_occ<BYTES>(__float_as_uint(qf));
            }
            occluded += BYTES ? acc : acc >> 31;
        }
    }
    return occluded;
}

// The voxels a ray can still enter with n free steps, relative to where it stands: i + j + k in 1..n steps along the two
// tangents (in the sample group's directions) and the normal.  Sorted by i + j + k: the first 19 entries are n = 3, all 34
// are n = 4.  Three bits per entry, 21 entries per 64-bit word.
struct OctantTable {
    unsigned long long i[2], j[2], k[2];
};
__host__ __device__ constexpr OctantTable make_octant_table() {
    OctantTable t{{0ull, 0ull}, {0ull, 0ull}, {0ull, 0ull}};
    int n = 0;
    for (int sum = 1; sum <= 4; sum++)
        for (int i = 0; i <= sum; i++)
            for (int j = 0; i + j <= sum; j++) {
                const int k = sum - i - j;
                t.i[n / 21] |= (unsigned long long)i << (3 * (n % 21));
                t.j[n / 21] |= (unsigned long long)j << (3 * (n % 21));
                t.k[n / 21] |= (unsigned long long)k << (3 * (n % 21));
                n++;
            }
    return t;
}
constexpr uint32_t OCTANT_VOXELS_3 = 19, OCTANT_VOXELS_4 = 34;

__device__ __forceinline__ AoSeed shuffle_seed(const AoSeed& s, int src) {
    AoSeed o;
    o.ex = __shfl_sync(0xffffffffu, s.ex, src);
    o.ey = __shfl_sync(0xffffffffu, s.ey, src);
    o.ez = __shfl_sync(0xffffffffu, s.ez, src);
    o.x = __shfl_sync(0xffffffffu, s.x, src);
    o.y = __shfl_sync(0xffffffffu, s.y, src);
    o.z = __shfl_sync(0xffffffffu, s.z, src);
    o.r = __shfl_sync(0xffffffffu, s.r, src);
    o.g = __shfl_sync(0xffffffffu, s.g, src);
    o.b = __shfl_sync(0xffffffffu, s.b, src);
    o.pixel = __shfl_sync(0xffffffffu, s.pixel, src);
    o.face_sign = __shfl_sync(0xffffffffu, s.face_sign, src);
    o.frame_pixel = __shfl_sync(0xffffffffu, s.frame_pixel, src);
    return o;
}

// AO of up to 32 hit pixels by one warp.  Called by all 32 lanes; every lane holds a seed -- lanes without a pixel of their
// own (valid = false) a copy of some valid lane's, so that they walk along without adding windows or stores.
// warpWin: AOW_WARP_WORDS words of shared memory owned by this warp, liquidWord: one more.
__device__ __forceinline__ void aow_warp(const FrameArgs& A, const AoSeed& s, bool valid, uint32_t* warpWin, uint32_t* liquidWord) {
    // one window per DISTINCT (hit voxel, face, normal sign) of the warp.  Neighbouring pixels mostly see the same voxel
    // face (1.14 distinct windows per 8x4 pixel tile of the benchmark frame), and the window depends on nothing else, so
    // the warp builds each one ONCE, cooperatively (one lane per 4^3 brick), instead of 32 times.
    const uint32_t lane = threadIdx.x & 31u;
    const rc_frame_params& P = A.p;
    const DeviceWorld& W = A.world;
    const AoDir* table = reinterpret_cast<const AoDir*>(A.ao_table);
    const int face = (int)(s.face_sign & 3u);
    const bool negNormal = (s.face_sign & 4u) != 0u;
    const bool f0 = face == 0, f1 = face == 1, f2 = face >= 2;
    const int strideY = aow_stride_y(face), strideZ = aow_stride_z(face);

    // window origin in bricks (raw coordinates): tangent axes cover h-4 .. h+4, the normal axis
    // h+1 .. h+5 or h-5 .. h-1
    const int nOff = negNormal ? -5 : 1;
    const int b0x = (s.x + (f0 ? nOff : -4)) >> 2;
    const int b0y = (s.y + (f1 ? nOff : -4)) >> 2;
    const int b0z = (s.z + ((!f0 && !f1) ? nOff : -4)) >> 2;

    // lanes with the same window: exact match on (x, y | z, face_sign)
    const unsigned long long k1 = (unsigned long long)(uint32_t)s.x | ((unsigned long long)(uint32_t)s.y << 32);
    const unsigned long long k2 = (unsigned long long)(uint32_t)s.z | ((unsigned long long)s.face_sign << 32);
    const unsigned peers = __match_any_sync(0xffffffffu, k1) & __match_any_sync(0xffffffffu, k2);
    const uint32_t leader = (uint32_t)__ffs(peers) - 1u;
    const unsigned leaders = __ballot_sync(0xffffffffu, lane == leader);
    const uint32_t slot = (uint32_t)__popc(leaders & ((1u << leader) - 1u));
    const uint32_t nSlots = (uint32_t)__popc(leaders);

    // Few distinct windows (the usual case): one BYTE per voxel, AOW_BYTE_SLOTS windows of AOW_BYTE_WORDS words fill the
    // warp's share exactly; a DDA step then needs no shift / mask to reach its voxel.  More: one bit per voxel.
    const bool byteMode = nSlots <= AOW_BYTE_SLOTS;
    const uint32_t slotWords = byteMode ? AOW_BYTE_WORDS : (uint32_t)AOW_WORDS;
    __syncwarp();  // the warp's previous chunk / tile is done with the windows
    for (uint32_t i = lane; i < nSlots * slotWords; i += 32u) warpWin[i] = 0u;
    if (lane == 0u) *liquidWord = 0u;
    __syncwarp();

    // fill: task = (window, brick); 18 bricks per window (2 along the normal, 3 x 3 across)
    {
        const uint32_t nTasks = nSlots * 18u;
        for (uint32_t base = 0; base < nTasks; base += 32u) {
            const uint32_t task = base + lane;
            const bool work = task < nTasks;
            const uint32_t sl = work ? task / 18u : 0u, b = work ? task - sl * 18u : 0u;
            const int src = (int)__fns(leaders, 0u, (int)sl + 1);  // lane of the sl-th leader
            const int lb0x = __shfl_sync(0xffffffffu, b0x, src), lb0y = __shfl_sync(0xffffffffu, b0y, src);
            const int lb0z = __shfl_sync(0xffffffffu, b0z, src);
            const int lface = __shfl_sync(0xffffffffu, face, src);
            if (work) {
                // brick b of a window whose extent is 2 bricks along its normal axis and 3 along the others
                const int nbx = lface == 0 ? 2 : 3, nby = lface == 1 ? 2 : 3;
                const int bx = (int)b % nbx, by = ((int)b / nbx) % nby, bz = (int)b / (nbx * nby);
                const int sY = aow_stride_y(lface), sZ = aow_stride_z(lface);
                Cursor cur;
                cursor_reset(W, cur);
                Region r;
                fetch_region(W, cur, wrap_coord(W, (lb0x + bx) << 2), wrap_coord(W, (lb0y + by) << 2), wrap_coord(W, (lb0z + bz) << 2), r);
                if (r.mask != 0ull) {
                    if (r.info & BR_LIQUID) atomicOr(liquidWord, 1u << sl);
                    uint32_t* win = warpWin + sl * slotWords;
                    const int pBrick = (bx << 2) + sY * (by << 2) + sZ * (bz << 2);
#ifdef RC_CHECKED
                    // bounds of the window fill (compute-sanitizer is closed on the pool: -DRC_CHECKED builds trap instead)
                    if (pBrick < 0 || pBrick + 3 + sY * 3 + sZ * 3 >= 32 * AOW_WORDS || (sl + 1u) * slotWords > AOW_WARP_WORDS)
                        __trap();  // a window is 32 * AOW_WORDS voxels (bits, or bytes in byte mode)
#endif
#pragma unroll
                    for (int z2 = 0; z2 < 4; z2++) {
                        const uint32_t slice = (uint32_t)(r.mask >> (16 * z2)) & 0xFFFFu;
                        if (slice == 0u) continue;
#pragma unroll
                        for (int y2 = 0; y2 < 4; y2++) {
                            const uint32_t row = (slice >> (4 * y2)) & 15u;   // the 4 voxels of one x-row of the brick
                            const int p0 = pBrick + sY * y2 + sZ * z2;        // multiple of 4
                            if (row == 0u) continue;
                            if (byteMode) {
                                // one word = exactly this row (bit i of the nibble -> byte i): no other task writes it
                                win[p0 >> 2] = (row * 0x00204081u) & 0x01010101u;
                            } else {
                                // bits are stored MSB-first inside each BYTE: bit p of the window = bit 7 - (p & 7) of byte
                                // p >> 3, so that the ray loop loads one byte at (index >> 3) and shifts the bit into the sign
                                atomicOr(&win[p0 >> 5], (__brev(row) >> 28) << ((p0 & 24) + 4 - (p0 & 4)));
                            }
                        }
                    }
                }
            }
        }
    }
    __syncwarp();
    const bool liquid = ((*liquidWord >> slot) & 1u) != 0u;

    const float fx = (float)s.x, fy = (float)s.y, fz = (float)s.z;
    // (pos.exact - pos.round) for a positive and for a negative step sign (exact is biased by -1)
    float a0x = FSUB(s.ex, fx), a1x = FSUB(FSUB(s.ex, 1.0f), fx);
    float a0y = FSUB(s.ey, fy), a1y = FSUB(FSUB(s.ey, 1.0f), fy);
    float a0z = FSUB(s.ez, fz), a1z = FSUB(FSUB(s.ez, 1.0f), fz);
    const int strideN = f0 ? 1 : (f1 ? strideY : strideZ);
    const int q0 = (s.x - (b0x << 2)) + strideY * (s.y - (b0y << 2)) + strideZ * (s.z - (b0z << 2));
    // index of the voxel the rays start from, counted from shared address 0: (8 *) address of this lane's window + its
    // index in the window: 1 voxel out from the hit voxel for a positive normal, 2 for a negative one
    const uint32_t winAddr = (uint32_t)__cvta_generic_to_shared(warpWin + slot * slotWords);
    const uint32_t q1 = (byteMode ? winAddr : 8u * winAddr) + (uint32_t)(q0 + (negNormal ? -strideN : strideN));
    const uint32_t qS = negNormal ? q1 - (uint32_t)strideN : q1;
#ifdef RC_CHECKED
    {
        // every window index a ray of this pixel can reach (5 steps from the hit voxel, the first through the face) lies
        // inside the lane's window
        const int tang = 4 * (f0 ? strideY + strideZ : (f1 ? 1 + strideZ : 1 + strideY));
        const int lo = q0 + (negNormal ? -5 * strideN : strideN) - tang, hi = q0 + (negNormal ? -strideN : 5 * strideN) + tang;
        if (lo < 0 || hi >= 32 * AOW_WORDS) __trap();
    }
#endif

    // FAST pixel (see the top of this section): pos.exact on the hit face, strictly inside the voxel along both tangents
    // for either step sign, no LIQUID in the window
    const float aF = f0 ? a0x : (f1 ? a0y : a0z);
    const float tlo = 9.5367431640625e-07f, thi = 1.0f - 9.5367431640625e-07f;  // 2^-20, 1 - 2^-20
    const float ta0 = f0 ? a0y : a0x, ta1 = f0 ? a1y : a1x, tb0 = f2 ? a0y : a0z, tb1 = f2 ? a1y : a1z;  // the two tangent axes
    const bool fast = !liquid && aF == 0.0f && ta0 >= tlo && ta0 <= thi && tb0 >= tlo && tb0 <= thi && -ta1 >= tlo && -ta1 <= thi &&
                      -tb1 >= tlo && -tb1 <= thi;
    uint32_t occluded = 0u;
    if (__any_sync(0xffffffffu, fast)) {
        const bool four = __any_sync(0xffffffffu, fast && !negNormal);
        uint32_t acc0 = byteMode ? aow_occ<true>(q1) : aow_occ<false>(q1);
        if (negNormal) acc0 |= byteMode ? aow_occ<true>(qS) : aow_occ<false>(qS);
        // AO_OPT_OCTANT: which sample groups (signs of the two tangent components) can still meet an occupied voxel at
        // all?  The free steps from qS each go along one axis in the group's direction: 19 (3 steps) or 34 (4 steps)
        // voxels per group.  Tested once per distinct window, the lanes sharing the 4 x 19 / 34 voxels; a group is skipped
        // when it is empty for every window of the warp (the loop over groups stays warp-uniform).
        uint32_t skip = 0u;
        if ((A.ao_opts & AO_OPT_OCTANT) != 0u && byteMode) {
            constexpr OctantTable OT = make_octant_table();
            uint32_t myOcc = 0u;
            for (uint32_t sl = 0; sl < nSlots; sl++) {
                const int src = (int)__fns(leaders, 0u, (int)sl + 1);
                const uint32_t lqS = __shfl_sync(0xffffffffu, qS, src);  // shared byte address of that window's start voxel
                const uint32_t lfs = __shfl_sync(0xffffffffu, s.face_sign, src);
                const int lface = (int)(lfs & 3u);
                const int st1 = aow_stride_y(lface), st2 = aow_stride_z(lface);
                const int sN = (lfs & 4u) ? -sel3i(lface, 1, st1, st2) : sel3i(lface, 1, st1, st2);
                const int sU = sel3i(lface, st1, st2, 1), sV = sel3i(lface, st2, 1, st1);  // axes (face + 1) % 3, (face + 2) % 3
                const uint32_t nVox = (lfs & 4u) ? OCTANT_VOXELS_3 : OCTANT_VOXELS_4;
                uint32_t h = 0u;
                for (uint32_t t = lane; t < 4u * nVox; t += 32u) {
                    const uint32_t g = t / nVox, o = t - g * nVox;
                    const uint32_t wd = o >= 21u ? 1u : 0u, sh = 3u * (o - 21u * wd);
                    const int vi = (int)(((wd ? OT.i[1] : OT.i[0]) >> sh) & 7ull), vj = (int)(((wd ? OT.j[1] : OT.j[0]) >> sh) & 7ull),
                              vk = (int)(((wd ? OT.k[1] : OT.k[0]) >> sh) & 7ull);
                    const int off = vi * ((g & 1u) ? -sU : sU) + vj * ((g & 2u) ? -sV : sV) + vk * sN;
                    h |= aow_occ<true>(lqS + (uint32_t)off) << g;
                }
                h = __reduce_or_sync(0xffffffffu, h);
                if (slot == sl) myOcc = h;
            }
            // lanes whose pixel is not fast do not cast window rays: they have no say
            skip = __reduce_and_sync(0xffffffffu, fast ? (~myOcc & 15u) : 15u);
        }
        // (exact - round) of the face axis: absDelta - 0 * delta = absDelta
        if (f0) a0x = a1x = 0.0f;
        else if (f1) a0y = a1y = 0.0f;
        else a0z = a1z = 0.0f;
        const uint32_t orient = (uint32_t)(face * 2 + (negNormal ? 1 : 0));
        const AoDir* dirs = table + (size_t)orient * RC_MAX_AO_SAMPLES;
        const float qSf = __uint_as_float(qS);
        const bool fourth = !negNormal;
        if (byteMode)
            occluded = four ? aow_cast_rays<true, true>(dirs, orient, qSf, acc0, a0x, a1x, a0y, a1y, a0z, a1z, fast, fourth, skip)
                            : aow_cast_rays<true, false>(dirs, orient, qSf, acc0, a0x, a1x, a0y, a1y, a0z, a1z, fast, fourth, skip);
        else
            occluded = four ? aow_cast_rays<false, true>(dirs, orient, qSf, acc0, a0x, a1x, a0y, a1y, a0z, a1z, fast, fourth, skip)
                            : aow_cast_rays<false, false>(dirs, orient, qSf, acc0, a0x, a1x, a0y, a1y, a0z, a1z, fast, fourth, skip);
    }
    // The other pixels, one at a time: their samples dealt to the lanes, each ray on the tree itself (castAo as it stands)
    unsigned slow = __ballot_sync(0xffffffffu, valid && !fast);
    while (slow) {
        const int src = __ffs(slow) - 1;
        slow &= slow - 1u;
        const int hx = __shfl_sync(0xffffffffu, s.x, src), hy = __shfl_sync(0xffffffffu, s.y, src), hz = __shfl_sync(0xffffffffu, s.z, src);
        const float hex = __shfl_sync(0xffffffffu, s.ex, src), hey = __shfl_sync(0xffffffffu, s.ey, src),
                    hez = __shfl_sync(0xffffffffu, s.ez, src);
        const uint32_t hfs = __shfl_sync(0xffffffffu, s.face_sign, src);
        const int hface = (int)(hfs & 3u);
        const AoDir* hd = table + (size_t)(hface * 2 + ((hfs & 4u) ? 1 : 0)) * RC_MAX_AO_SAMPLES;
        const float hfx = (float)hx, hfy = (float)hy, hfz = (float)hz;
        uint32_t cnt = 0u;
        for (uint32_t k = lane; k < P.ao_samples; k += 32u) {
            const float4 v0 = __ldg(reinterpret_cast<const float4*>(hd + k));
            const float4 v1 = __ldg(reinterpret_cast<const float4*>(hd + k) + 1);
            const float ix = v0.x, iy = v0.y, iz = v0.z;
            const float adx = fabsf(ix), ady = fabsf(iy), adz = fabsf(iz);
            const bool nx = v0.w < 0, ny = v1.x < 0, nz = v1.y < 0;  // the strides' signs are the ray's step signs
            // pos.exact re-biased for the ray's signs; deltaPos = absDelta - (exact - round) * delta; deltaPos[face] -= absDelta[face]
            const float ex = nx ? FSUB(hex, 1.0f) : hex, ey = ny ? FSUB(hey, 1.0f) : hey, ez = nz ? FSUB(hez, 1.0f) : hez;
            float px = FSUB(adx, FMUL(FSUB(ex, hfx), ix));
            float py = FSUB(ady, FMUL(FSUB(ey, hfy), iy));
            float pz = FSUB(adz, FMUL(FSUB(ez, hfz), iz));
            if (hface == 0) px = FSUB(px, adx);
            else if (hface == 1) py = FSUB(py, ady);
            else pz = FSUB(pz, adz);
            cnt += ao_ray_general(W, hx, hy, hz, px, py, pz, adx, ady, adz, nx ? -1 : 1, ny ? -1 : 1, nz ? -1 : 1, AOW_STEPS);
        }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        if ((int)lane == src) occluded = cnt;
    }
    if (valid) {
        const float ao = FSUB(1.0f, FMUL(P.ao_strength, FDIV((float)occluded, (float)P.ao_samples)));
        store_pixel(A, s.pixel, s.frame_pixel, F3{FMUL(s.r, ao), FMUL(s.g, ao), FMUL(s.b, ao)});
        if (A.hits) A.hits[s.pixel].ao_occluded = (uint8_t)occluded;
    }
}


// ---------------------------------------------------------------------------
// One 8x4-pixel tile by one warp: the pixel of low_res.frag main() for each lane, the stores, and the AO of the tile's hit
// pixels -- cast by this same warp right away (FUSED: the seeds never leave the registers, there is no seed list and no
// second launch) or appended to the seed list for an AO kernel.
// ---------------------------------------------------------------------------
template <bool FUSED>
__device__ __forceinline__ void trace_tile(const FrameArgs& A, uint32_t px, uint32_t localRow, uint32_t lane, uint32_t* warpWin,
                                           uint32_t* liquidWord, const uint4* top, uint32_t ntop) {
    // local band -> global row
    const uint32_t band = (localRow / BAND_ROWS) * A.world_size + A.rank;
    const uint32_t py = band * BAND_ROWS + (localRow % BAND_ROWS);
    const bool active = px < A.p.width && py < A.p.height && localRow < A.local_rows;

    PixelResult R;
    R.traced = false;
    R.shadowRay = false;
    R.aoSeed = false;
    R.hit.kind = RC_KIND_SKY;
    R.hit.steps = 0;
#ifdef RC_COST_MAP
    const long long costT0 = clock64();
    R.mark[0] = R.mark[1] = R.mark[2] = 0;
#endif
    shade_pixel(A, active, px, py, R, top, ntop);
    if (active) {
        const uint32_t idx = local_index(A, px, localRow);
        // AO pixels get their final colour from the AO rays
        if (!(R.aoSeed)) store_pixel(A, idx, frame_pixel_of(A, px, localRow), R.color);
        if (A.hits) A.hits[idx] = R.hit;
#ifdef RC_COST_MAP
        // developer probe (tools/cost_map.py): cycles this warp spent in shade_pixel, in the alpha channel of the float frame
        if (A.rgba32f) {
            const long long t1 = clock64();
            const long long m0 = R.mark[0] ? R.mark[0] : t1, m1 = R.mark[1] ? R.mark[1] : m0, m2 = R.mark[2] ? R.mark[2] : t1;
            // x: to the end of loop A / free flight, y: ratios + replay, z: loop B + shading, w: everything (rest = shadow ray)
            A.rgba32f[idx] = make_float4((float)(m0 - costT0), (float)(m1 - m0), (float)(m2 - m1), (float)(t1 - costT0));
        }
#endif
    }

    // AO seeds
    const bool seed = active && R.aoSeed;
    const unsigned seedMask = __ballot_sync(0xffffffffu, seed);
    if (seedMask) {
        const int leader = __ffs(seedMask) - 1;
        AoSeed s;
        s.ex = R.sx_;
        s.ey = R.sy_;
        s.ez = R.sz_;
        s.x = R.hit.x;
        s.y = R.hit.y;
        s.z = R.hit.z;
        s.r = R.color.x;
        s.g = R.color.y;
        s.b = R.color.z;
        s.pixel = local_index(A, px, localRow);
        const uint32_t sneg = R.hit.face == 0 ? (R.hit.flags & RC_HF_STEPX_NEG)
                                              : (R.hit.face == 1 ? (R.hit.flags & RC_HF_STEPY_NEG) : (R.hit.flags & RC_HF_STEPZ_NEG));
        // normal sign = -step[face]: negative normal iff the step was positive
        s.face_sign = (uint32_t)R.hit.face | (sneg ? 0u : 4u);
        s.frame_pixel = frame_pixel_of(A, px, localRow);
        if (FUSED) {
            if ((int)lane == leader) atomicAdd(&counter_slot(A).ao_seed_count, (unsigned)__popc(seedMask));
            // lanes without a hit pixel walk along with the first seeded lane's
            const AoSeed sh = shuffle_seed(s, leader);
            if (!seed) s = sh;
            aow_warp(A, s, seed, warpWin, liquidWord);
        } else {
            // warp-aggregated append
            unsigned base = 0;
            if ((int)lane == leader) base = atomicAdd(&A.counters->seed_cursor.next, (unsigned)__popc(seedMask));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (seed) {
                const uint32_t at = base + __popc(seedMask & ((1u << lane) - 1u));
#ifdef RC_CHECKED
                if (at >= A.local_rows * A.p.width) __trap();  // seed cursor beyond the list
#endif
                A.ao_seeds[at] = s;
            }
        }
    }

    // frame statistics: one atomic per warp per counter
    const unsigned tracedMask = __ballot_sync(0xffffffffu, active && R.traced);
    const unsigned shadowMask = __ballot_sync(0xffffffffu, active && R.shadowRay);
    const unsigned hitMask = __ballot_sync(0xffffffffu, active && R.hit.kind == RC_KIND_HIT);
    unsigned stepSum = active ? (unsigned)R.hit.steps : 0u;
    for (int o = 16; o > 0; o >>= 1) stepSum += __shfl_xor_sync(0xffffffffu, stepSum, o);
    if (lane == 0) {
        FrameCounters& fc = counter_slot(A);
        if (tracedMask) atomicAdd(&fc.primary_rays, (unsigned long long)__popc(tracedMask));
        if (shadowMask) atomicAdd(&fc.shadow_rays, (unsigned long long)__popc(shadowMask));
        if (hitMask) atomicAdd(&fc.hit_pixels, (unsigned long long)__popc(hitMask));
        if (stepSum) atomicAdd(&fc.primary_steps, (unsigned long long)stepSum);
    }
}

// grid.x = ceil(width/16), grid.y = local_rows/8 ; block = 128 threads = 4 warps,
// each warp an 8x4 pixel tile, the block a 16x8 tile.
#ifndef RC_PRIMARY_WARPS
#define RC_PRIMARY_WARPS 4  // warps per CTA: 4 = a 16x8 pixel tile, 2 = 16x4, 1 = 8x4
#endif
#ifndef RC_PRIMARY_MIN_BLOCKS
#define RC_PRIMARY_MIN_BLOCKS (32 / RC_PRIMARY_WARPS)  // 64 registers: 32 warps per SM
#endif
constexpr uint32_t PRIMARY_TILE_W = RC_PRIMARY_WARPS >= 2 ? 16u : 8u, PRIMARY_TILE_H = RC_PRIMARY_WARPS >= 4 ? 8u : 4u;
// dynamic shared memory of the kernels that cast AO rays: [warps][AOW_WARP_WORDS] windows, [warps] liquid words
constexpr uint32_t aow_smem_bytes(uint32_t warps) { return warps * (AOW_WARP_WORDS + 1u) * 4u; }

// FUSED = false: primary + shading + shadow rays, hit pixels append AO seeds (SCHED_TWO_KERNELS, SCHED_AO_PERSISTENT).
// FUSED = true: the AO rays too (SCHED_FUSED; needs aow_smem_bytes(RC_PRIMARY_WARPS) of dynamic shared memory).
template <bool FUSED>
__global__ void __launch_bounds__(32 * RC_PRIMARY_WARPS, RC_PRIMARY_MIN_BLOCKS) primary_kernel(const __grid_constant__ FrameArgs A) {
    extern __shared__ __align__(16) uint32_t smem[];
    // an AO kernel launched behind this one with programmatic stream serialization may take the SM slots this grid's last
    // wave leaves free; it waits (griddepcontrol.wait) for the whole grid before it reads a seed
    asm volatile("griddepcontrol.launch_dependents;");
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t px = blockIdx.x * PRIMARY_TILE_W + (warp & 1u) * 8u + (lane & 7u);
    const uint32_t localRow = blockIdx.y * PRIMARY_TILE_H + (warp >> 1) * 4u + (lane >> 3);
    trace_tile<FUSED>(A, px, localRow, lane, FUSED ? smem + warp * AOW_WARP_WORDS : nullptr,
                      FUSED ? smem + RC_PRIMARY_WARPS * AOW_WARP_WORDS + warp : nullptr, nullptr, 0u);
}

// SCHED_PERSISTENT / SCHED_PERSISTENT_STAGED: resident warps pull 8x4-pixel tiles from a global atomic queue
// (FrameCounters::tile_cursor) until the frame is done.  Tile order: Morton inside super-tiles of 8x8 tiles (64x32 pixels),
// super-tiles row by row -- consecutive queue entries are neighbours on the screen, so the warps of an SM work on rays
// that share tree nodes.  STAGED: the top tree levels (the breadth-first prefix of DeviceWorld::n_top nodes: root,
// 256^3 and 64^3 levels) are copied into shared memory once per CTA and every descent reads them from there.
constexpr uint32_t PERSISTENT_WARPS = 4;
constexpr uint32_t STAGED_MAX_NODES = 1024;  // 16 KB
template <bool STAGED>
__global__ void __launch_bounds__(32 * PERSISTENT_WARPS, RC_PRIMARY_MIN_BLOCKS) frame_persistent_kernel(const __grid_constant__ FrameArgs A) {
    extern __shared__ __align__(16) uint32_t smem[];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t* const warpWin = smem + warp * AOW_WARP_WORDS;
    uint32_t* const liquidWord = smem + PERSISTENT_WARPS * AOW_WARP_WORDS + warp;
    const uint4* top = nullptr;
    uint32_t ntop = 0u;
    if (STAGED) {
        // after the windows and liquid words (PERSISTENT_WARPS * (AOW_WARP_WORDS + 1) words: a multiple of 16 bytes)
        uint4* t = reinterpret_cast<uint4*>(smem + PERSISTENT_WARPS * (AOW_WARP_WORDS + 1u));
        ntop = min(A.world.n_top, STAGED_MAX_NODES);
        for (uint32_t i = threadIdx.x; i < ntop; i += blockDim.x) t[i] = __ldg(&A.world.nodes[i]);
        __syncthreads();
        top = t;
    }
    const uint32_t tilesX = (A.p.width + 7u) / 8u, tilesY = (A.local_rows + 3u) / 4u;
    const uint32_t supersX = (tilesX + 7u) / 8u, supersY = (tilesY + 7u) / 8u;
    const uint32_t nTiles = supersX * supersY * 64u;
    // the next tile is claimed while the current one is traced: the atomic's round trip never shows
    uint32_t next = 0u;
    if (lane == 0u) next = atomicAdd(&A.counters->tile_cursor.next, 1u);
    for (;;) {
        const uint32_t t = __shfl_sync(0xffffffffu, next, 0);
        if (t >= nTiles) break;
        if (lane == 0u) next = atomicAdd(&A.counters->tile_cursor.next, 1u);
        const uint32_t sup = t >> 6, m = t & 63u;
        // Morton: even bits of m -> x, odd bits -> y
        const uint32_t mx = (m & 1u) | ((m >> 1) & 2u) | ((m >> 2) & 4u), my = ((m >> 1) & 1u) | ((m >> 2) & 2u) | ((m >> 3) & 4u);
        const uint32_t tx = (sup % supersX) * 8u + mx, ty = (sup / supersX) * 8u + my;
        if (tx >= tilesX || ty >= tilesY) continue;
        trace_tile<true>(A, tx * 8u + (lane & 7u), ty * 4u + (lane >> 3), lane, warpWin, liquidWord, top, ntop);
    }
}

// SCHED_TWO_KERNELS: one warp per 32 consecutive seeds of the list, grid sized for the worst case (every pixel a hit);
// warps beyond the list leave at once.
#ifndef RC_AO_MIN_BLOCKS
#define RC_AO_MIN_BLOCKS 8  // 64 registers
#endif
__global__ void __launch_bounds__(AO_THREADS, RC_AO_MIN_BLOCKS) ao_window_kernel(const __grid_constant__ FrameArgs A) {
    extern __shared__ __align__(16) uint32_t smem[];
    const uint32_t nSeeds = A.counters->seed_cursor.next;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t warp = threadIdx.x >> 5;
    if ((t & ~31u) >= nSeeds) return;  // whole warp beyond the list
    const bool valid = t < nSeeds;
    const AoSeed s = A.ao_seeds[valid ? t : nSeeds - 1u];  // idle lanes of the last warp shadow a real seed
    aow_warp(A, s, valid, smem + warp * AOW_WARP_WORDS, smem + (AO_THREADS / 32) * AOW_WARP_WORDS + warp);
}

// SCHED_AO_PERSISTENT: as many CTAs as the device holds at once; every warp pulls 32-seed chunks from a global atomic cursor
// (FrameCounters::ao_cursor) until the list is empty -- no CTAs beyond the list, no tail of half-empty waves.  Launched with
// programmatic stream serialization: the grid may become resident while the primary grid drains, and waits here until
// that grid has completed and its seeds are visible.
__global__ void __launch_bounds__(AO_THREADS, RC_AO_MIN_BLOCKS) ao_window_persistent_kernel(const __grid_constant__ FrameArgs A) {
    extern __shared__ __align__(16) uint32_t smem[];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const uint32_t nSeeds = *(volatile const unsigned int*)&A.counters->seed_cursor.next;
    const uint32_t nChunks = (nSeeds + 31u) / 32u;
    uint32_t next = 0u;
    if (lane == 0u) next = atomicAdd(&A.counters->ao_cursor.next, 1u);
    for (;;) {
        const uint32_t c = __shfl_sync(0xffffffffu, next, 0);
        if (c >= nChunks) break;
        if (lane == 0u) next = atomicAdd(&A.counters->ao_cursor.next, 1u);
        const uint32_t t = c * 32u + lane;
        const bool valid = t < nSeeds;
        const AoSeed s = A.ao_seeds[valid ? t : nSeeds - 1u];
        aow_warp(A, s, valid, smem + warp * AOW_WARP_WORDS, smem + (AO_THREADS / 32) * AOW_WARP_WORDS + warp);
    }
}


// Frame composition after the NCCL gather: row y of the frame lives in rank (y/16) % world,
// local row ((y/16) / world) * 16 + y % 16. One thread per 16-byte (or 4-byte) word of the frame:
// coalesced reads of whole rows, coalesced writes.
template <typename WordT>
__global__ void __launch_bounds__(256) compose_kernel(const WordT* __restrict__ gathered, WordT* __restrict__ dst,
                                                      uint32_t wordsPerRow, uint32_t height, uint32_t worldSize,
                                                      uint32_t paddedRows) {
    const uint64_t total = (uint64_t)wordsPerRow * height;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t y = (uint32_t)(i / wordsPerRow), xw = (uint32_t)(i % wordsPerRow);
        const uint32_t band = y / BAND_ROWS;
        const uint32_t rank = band % worldSize;
        const uint32_t srcRow = (band / worldSize) * BAND_ROWS + (y % BAND_ROWS);
        dst[i] = gathered[((uint64_t)rank * paddedRows + srcRow) * wordsPerRow + xw];
    }
}

// RAY_CASTER::castRayFromCam (src/ray_caster.cpp:54-87): fp32 quotient 1/dir widened to double,
// double deltaPos, the host getBlock's plain torus (W.leak is 0 here).
struct PickArgs {
    float px, py, pz, dx, dy, dz;
    int steps;
};
__global__ void pick_kernel(DeviceWorld W, PickArgs a, int* out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const float pos[3] = {a.px, a.py, a.pz}, dir[3] = {a.dx, a.dy, a.dz};
    int step[3], rnd[3], last[3];
    double ad[3], t[3];
    for (int k = 0; k < 3; k++) {
        step[k] = dir[k] < 0.0f ? -1 : 1;
        const double delta = (double)__fdiv_rn(1.0f, dir[k]);
        ad[k] = fabs(delta);
        rnd[k] = (int)truncf(pos[k]);
        double exact = (double)pos[k];
        if (step[k] < 0) exact = __dsub_rn(exact, 1.0);
        t[k] = __dsub_rn(ad[k], __dmul_rn(__dsub_rn(exact, (double)rnd[k]), delta));
        last[k] = rnd[k];
    }
    Cursor cur;
    cursor_reset(W, cur);
    int left = 0, steps = a.steps;
    while (steps-- > 0) {
        for (int k = 0; k < 3; k++) last[k] = rnd[k];
        const int ax = (t[0] < t[1] && t[0] < t[2]) ? 0 : (t[1] < t[2] ? 1 : 2);
        rnd[ax] += step[ax];
        t[ax] = __dadd_rn(t[ax], ad[ax]);
        int shift;
        const uint32_t w = lookup(W, cur, rnd[0], rnd[1], rnd[2], shift);
        if (w != 0u && (word_flags(w) & 0x10u) == 0u) {
            left = steps;
            break;
        }
    }
    for (int k = 0; k < 3; k++) {
        out[k] = rnd[k];
        out[3 + k] = last[k];
    }
    out[6] = left;
}

// Highest occupied y per column block (edge 2^bs voxels, SKY_GRID x SKY_GRID blocks over the torus), the input of the
// sky gates (FrameArgs::sky_gate).  Thread (i, j, k) follows child i of the root, child j of that, child k of that (most
// threads find a clear mask bit on the way and leave) and walks the subtree below -- a 16^3 cell at max_depth 6 --
// depth-first with an explicit stack; a uniform cell counts for every block it covers, a brick for the block it lies in
// (bs >= 2) with its highest occupied layer.  A uniform cell or brick met on the way down belongs to the thread whose
// remaining digits are 0.  Only nodes reachable from the root are seen, so runs abandoned by rc_update_world do not count.
struct TopsFrame {
    unsigned long long mask, todo;
    uint32_t base, kind;
    int ox, oy, oz;
};
constexpr int TOPS_MAX_LEVELS = 12;
constexpr int TOPS_PREFIX = 3;  // levels resolved by the thread index: 64^3 threads
__device__ __forceinline__ void tops_add_cell(int* tops, int bs, int x0, int z0, int edge, int top) {
    const int bx0 = x0 >> bs, bx1 = (x0 + edge - 1) >> bs, bz0 = z0 >> bs, bz1 = (z0 + edge - 1) >> bs;
    for (int bz = bz0; bz <= bz1; bz++)
        for (int bx = bx0; bx <= bx1; bx++) atomicMax(&tops[(bz & (SKY_GRID - 1)) * SKY_GRID + (bx & (SKY_GRID - 1))], top);
}
__device__ __forceinline__ void tops_add_brick(int* tops, int bs, unsigned long long mask, int ox, int oy, int oz) {
    if (mask == 0ull) return;
    int ymax = 0;
    for (int k = 1; k < 4; k++)
        if (mask & (0x000F000F000F000Full << (4 * k))) ymax = k;
    tops_add_cell(tops, bs, ox, oz, 4, oy + ymax);
}
__global__ void __launch_bounds__(128) world_tops_kernel(DeviceWorld W, int bs, int* tops) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;  // 64^3 threads
    TopsFrame st[TOPS_MAX_LEVELS];
    {
        const uint4 n = __ldg(&W.nodes[0]);
        st[0].mask = (unsigned long long)n.x | ((unsigned long long)n.y << 32);
        st[0].base = n.z;
        st[0].kind = n.w;
        st[0].ox = st[0].oy = st[0].oz = 0;
    }
    // the way down: digit l of the thread index picks the child at level l
    for (int l = 0; l < TOPS_PREFIX; l++) {
        TopsFrame& f = st[0];
        const int shift = W.root_shift - 2 * l;  // log2 of the edge of this node's child cells
        const uint32_t rest = l == 0 ? t : (t & ((1u << (6 * (TOPS_PREFIX - l))) - 1u));  // this digit and the ones below
        if ((f.kind & PK_KIND_MASK) == PK_BRICK || shift <= 0) {
            if (rest == 0u) tops_add_brick(tops, bs, f.mask, f.ox, f.oy, f.oz);
            return;
        }
        const int idx = (int)(rest >> (6 * (TOPS_PREFIX - 1 - l)));
        if (!((f.mask >> idx) & 1ull)) return;
        const int cx = f.ox + ((idx & 3) << shift), cy = f.oy + (((idx >> 2) & 3) << shift), cz = f.oz + ((idx >> 4) << shift);
        const uint4 n = __ldg(&W.nodes[f.base + (uint32_t)__popcll(f.mask & ((1ull << idx) - 1ull))]);
        if ((n.w & PK_KIND_MASK) == PK_LEAF) {
            if ((rest & ((1u << (6 * (TOPS_PREFIX - 1 - l))) - 1u)) == 0u)
                tops_add_cell(tops, bs, cx, cz, 1 << shift, cy + (1 << shift) - 1);
            return;
        }
        f.mask = (unsigned long long)n.x | ((unsigned long long)n.y << 32);
        f.base = n.z;
        f.kind = n.w;
        f.ox = cx;
        f.oy = cy;
        f.oz = cz;
    }
    // the subtree below, depth-first
    st[0].todo = st[0].mask;
    int sp = 0;
    while (sp >= 0) {
        TopsFrame& f = st[sp];
        const int shift = W.root_shift - 2 * (sp + TOPS_PREFIX);
        if ((f.kind & PK_KIND_MASK) == PK_BRICK || shift <= 0) {
            tops_add_brick(tops, bs, f.mask, f.ox, f.oy, f.oz);
            sp--;
            continue;
        }
        if (f.todo == 0ull) {
            sp--;
            continue;
        }
        const int idx = __ffsll((long long)f.todo) - 1;
        f.todo &= f.todo - 1ull;
        const int cx = f.ox + ((idx & 3) << shift), cy = f.oy + (((idx >> 2) & 3) << shift), cz = f.oz + ((idx >> 4) << shift);
        const uint4 n = __ldg(&W.nodes[f.base + (uint32_t)__popcll(f.mask & ((1ull << idx) - 1ull))]);
        if ((n.w & PK_KIND_MASK) == PK_LEAF) {
            tops_add_cell(tops, bs, cx, cz, 1 << shift, cy + (1 << shift) - 1);
            continue;
        }
        if (sp + 1 >= TOPS_MAX_LEVELS) continue;
        TopsFrame& c = st[sp + 1];
        c.mask = (unsigned long long)n.x | ((unsigned long long)n.y << 32);
        c.base = n.z;
        c.kind = n.w;
        c.ox = cx;
        c.oy = cy;
        c.oz = cz;
        c.todo = c.mask;
        sp++;
    }
}

// Top table (DeviceWorld::top_table): thread = one aligned cell of shift top_shift; walks down from the root to the node of
// that cell and leaves its index (bit 31 set), or the shift of the empty cell it ran into on the way.  A uniform cell
// (PK_LEAF) met on the way is entered with its own node index: the lookup reads the material word from it.
__global__ void __launch_bounds__(128) top_table_kernel(DeviceWorld W, uint32_t* __restrict__ table) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t dimMask = (1u << W.top_bits) - 1u;
    if ((t >> (3 * W.top_bits)) != 0u) return;
    const uint32_t x = (t & dimMask) << W.top_shift, y = ((t >> W.top_bits) & dimMask) << W.top_shift,
                   z = ((t >> (2 * W.top_bits)) & dimMask) << W.top_shift;
    uint32_t ni = 0u;
    uint4 n = __ldg(&W.nodes[0]);
    for (int s = W.root_shift; s >= W.top_shift; s -= 2) {
        const unsigned long long mask = (unsigned long long)n.x | ((unsigned long long)n.y << 32);
        const uint32_t idx = (((z >> s) & 3u) << 4) | (((y >> s) & 3u) << 2) | ((x >> s) & 3u);
        if (!((mask >> idx) & 1ull)) {
            table[t] = (uint32_t)s;
            return;
        }
        ni = n.z + (uint32_t)__popcll(mask & ((1ull << idx) - 1ull));
        n = __ldg(&W.nodes[ni]);
        if ((n.w & PK_KIND_MASK) == PK_LEAF) break;
    }
    table[t] = 0x80000000u | ni;
}

// Gate table for one step budget (FrameArgs::sky_gate): thread = (quadrant, block); rings of Chebyshev block distance
// 0-1, 2-3, 4-7, 8-15, 16+ inside the window a ray can cross in maxSteps steps.
__global__ void __launch_bounds__(128) sky_gate_kernel(const int* __restrict__ tops, int bs, uint32_t maxSteps, uint4* __restrict__ gates) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 4u * SKY_GRID * SKY_GRID) return;
    const int bx = (int)(t % SKY_GRID), bz = (int)((t / SKY_GRID) % SKY_GRID), q = (int)(t / (SKY_GRID * SKY_GRID));
    const int dx = (q & 1) ? -1 : 1, dz = (q & 2) ? -1 : 1;
    const int nb = (int)(((1u << bs) - 1u + maxSteps) >> bs);  // blocks a ray can cross along one axis
    uint32_t g[5] = {0u, 0u, 0u, 0u, 0u};
    for (int j = 0; j <= nb; j++)
        for (int i = 0; i <= nb; i++) {
            const int d = max(i, j);
            const int ring = d < 2 ? 0 : (d < 4 ? 1 : (d < 8 ? 2 : (d < 16 ? 3 : 4)));
            const int x = bx + dx * i, z = bz + dz * j;
            const uint32_t v = (x < 0 || x >= SKY_GRID || z < 0 || z >= SKY_GRID) ? 65535u : (uint32_t)(tops[z * SKY_GRID + x] + 1);
            g[ring] = max(g[ring], v);
        }
    gates[t] = make_uint4(g[0] | (g[1] << 16), g[2] | (g[3] << 16), g[4], 0u);
}

// rc_frame_sync_*: see kernels.h
__global__ void sync_signal_kernel(uint32_t* flag, uint32_t value) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        __threadfence_system();  // this device's earlier stores (its pixels in the owner's frame) before the flag
        atomicMax_system(flag, value);  // flags only grow: signals from two streams may arrive in either order
    }
}
__global__ void sync_wait_kernel(const volatile uint32_t* flags, uint32_t strideWords, uint32_t count, uint32_t value, uint32_t* timedOut) {
    if (blockIdx.x == 0 && threadIdx.x < count) {
        const volatile uint32_t* f = flags + (size_t)threadIdx.x * strideWords;
        const long long t0 = clock64();
        while ((int)(*f - value) < 0) {
            __nanosleep(200);
            if (clock64() - t0 > 4000000000ll) {  // ~2 s at 2 GHz: a peer died or the protocol broke -- do not hang the device
                atomicAdd(timedOut, 1u);
                break;
            }
        }
    }
    __threadfence_system();
}

__global__ void lookup_kernel(DeviceWorld W, const int* xyz, uint32_t n, uint32_t* outWords, int* outShift) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    Cursor cur;
    cursor_reset(W, cur);
    int shift = -1;
    outWords[t] = lookup(W, cur, xyz[3 * t], xyz[3 * t + 1], xyz[3 * t + 2], shift);
    outShift[t] = shift;
}

__global__ void detmath_kernel(const float* in, uint32_t n, float* outSin, float* outExp) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    outSin[t] = det_sin(in[t]);
    outExp[t] = det_exp(in[t]);
}

// *table: AoDir[6][RC_MAX_AO_SAMPLES] on the current device (allocated here on first use; one per device context).
cudaError_t upload_hemisphere(const float* dirs, uint32_t n, cudaStream_t stream, void** table) {
    if (n > RC_MAX_AO_SAMPLES) return cudaErrorInvalidValue;
    float inv[3 * RC_MAX_AO_SAMPLES];
    for (uint32_t i = 0; i < 3 * n; i++) inv[i] = 1.0f / dirs[i];  // IEEE division, as the oracle's buildRay does per ray
    cudaError_t e = cudaMemcpyToSymbolAsync(c_hemi, dirs, 3 * n * sizeof(float), 0, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return e;
    e = cudaMemcpyToSymbolAsync(c_hemi_inv, inv, 3 * n * sizeof(float), 0, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return e;

    // oriented table of the window kernel: d[face] = sg*up, d[(face+1)%3] = x, d[(face+2)%3] = y; samples grouped by
    // the signs of (x, y), see c_aoSeg
    static AoDir host[6 * RC_MAX_AO_SAMPLES];
    uint32_t order[RC_MAX_AO_SAMPLES], seg[5] = {0, 0, 0, 0, 0};
    {
        uint32_t m = 0;
        for (uint32_t g = 0; g < 4; g++) {
            seg[g] = m;
            for (uint32_t k = 0; k < n; k++)
                if ((uint32_t)(dirs[3 * k + 0] < 0.0f) + 2u * (uint32_t)(dirs[3 * k + 1] < 0.0f) == g) order[m++] = k;
        }
        seg[4] = m;  // == n
    }
    e = cudaMemcpyToSymbolAsync(c_aoSeg, seg, sizeof(seg), 0, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return e;
    for (int face = 0; face < 3; face++) {
        const int strides[3] = {1, face == 0 ? 8 : 16, face == 0 ? 96 : (face == 1 ? 128 : 192)};
        for (int neg = 0; neg < 2; neg++)
            for (uint32_t slot = 0; slot < n; slot++) {
                const uint32_t k = order[slot];
                float d[3], iv[3];
                const float up = dirs[3 * k + 2], iup = inv[3 * k + 2];
                d[face] = neg ? -up : up;
                iv[face] = neg ? -iup : iup;  // 1/(-up) == -(1/up) exactly
                d[(face + 1) % 3] = dirs[3 * k + 0];
                iv[(face + 1) % 3] = inv[3 * k + 0];
                d[(face + 2) % 3] = dirs[3 * k + 1];
                iv[(face + 2) % 3] = inv[3 * k + 1];
                AoDir& o = host[(size_t)(face * 2 + neg) * RC_MAX_AO_SAMPLES + slot];
                o.ix = iv[0];
                o.iy = iv[1];
                o.iz = iv[2];
                auto denorm = [](bool neg, int stride) {
                    const uint32_t bits = (neg ? 0x80000000u : 0u) | (uint32_t)stride;
                    float f;
                    memcpy(&f, &bits, 4);
                    return f;
                };
                o.qx = denorm(d[0] < 0.0f, strides[0]);
                o.qy = denorm(d[1] < 0.0f, strides[1]);
                o.qz = denorm(d[2] < 0.0f, strides[2]);
                o.lx = face == 0 ? fabsf(iv[0]) : 0.0f;
                o.ly = face == 1 ? fabsf(iv[1]) : 0.0f;
            }
    }
    static_assert(sizeof(host) == AO_TABLE_BYTES, "kernels.h states the table's size");
    if (!*table) {
        e = cudaMalloc(table, sizeof(host));
        if (e != cudaSuccess) return e;
    }
    e = cudaMemcpyAsync(*table, host, sizeof(host), cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return e;
#ifdef RC_AO_TABLE_CONSTANT
    e = cudaMemcpyToSymbolAsync(c_aoTable, host, sizeof(host), 0, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return e;
#endif
    return cudaStreamSynchronize(stream);  // inv[] lives on this stack frame
}

void release_hemisphere(void* table) {
    if (table) cudaFree(table);
}

// Kernels with dynamic shared memory above 48 KB need the opt-in once per device; cheap enough to repeat per launch.
template <typename K>
static cudaError_t allow_smem(K kernel, uint32_t bytes) {
    if (bytes <= 48u * 1024u) return cudaSuccess;
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

cudaError_t launch_frame(const FrameArgs& args, uint32_t schedule, int numSms, cudaEvent_t midEvent, cudaStream_t stream,
                         uint32_t* launches, uint32_t* scheduleUsed) {
    *launches = 0;
    const bool ao = args.p.ao_samples > 0 && args.local_rows > 0;
    // the window code serves ao_steps == 5 (the reference's count, light_scattering.frag:231); other counts walk the tree
    // with a brick cache (ao_cached_kernel)
    const bool window = ao && args.ao_table && args.p.ao_steps == AOW_STEPS;
    const bool fusable = !ao || window;
    if (schedule >= SCHED_COUNT)  // SCHED_AUTO
        schedule = (uint64_t)args.local_rows * args.p.width >= (4u << 20) ? SCHED_AO_PERSISTENT : SCHED_TWO_KERNELS;
    if (schedule >= SCHED_FUSED && !fusable) schedule = SCHED_TWO_KERNELS;
    if (schedule == SCHED_AO_PERSISTENT && !window) schedule = SCHED_TWO_KERNELS;
    *scheduleUsed = schedule;
    if (args.local_rows == 0 || args.p.width == 0) return cudaSuccess;
    cudaError_t e;

    if (schedule == SCHED_PERSISTENT || schedule == SCHED_PERSISTENT_STAGED) {
        const bool staged = schedule == SCHED_PERSISTENT_STAGED;
        const uint32_t ntop = staged ? (args.world.n_top < STAGED_MAX_NODES ? args.world.n_top : STAGED_MAX_NODES) : 0u;
        const uint32_t bytes = aow_smem_bytes(PERSISTENT_WARPS) + ntop * 16u;
        const unsigned blocks = (unsigned)numSms * (unsigned)RC_PRIMARY_MIN_BLOCKS;
        if (staged) {
            if ((e = allow_smem(frame_persistent_kernel<true>, bytes)) != cudaSuccess) return e;
            frame_persistent_kernel<true><<<blocks, 32 * PERSISTENT_WARPS, bytes, stream>>>(args);
        } else {
            if ((e = allow_smem(frame_persistent_kernel<false>, bytes)) != cudaSuccess) return e;
            frame_persistent_kernel<false><<<blocks, 32 * PERSISTENT_WARPS, bytes, stream>>>(args);
        }
        ++*launches;
        return cudaGetLastError();
    }

    dim3 grid((args.p.width + PRIMARY_TILE_W - 1u) / PRIMARY_TILE_W, (args.local_rows + PRIMARY_TILE_H - 1u) / PRIMARY_TILE_H);
    if (schedule == SCHED_FUSED) {
        const uint32_t bytes = aow_smem_bytes(RC_PRIMARY_WARPS);
        if ((e = allow_smem(primary_kernel<true>, bytes)) != cudaSuccess) return e;
        primary_kernel<true><<<grid, 32 * RC_PRIMARY_WARPS, bytes, stream>>>(args);
        ++*launches;
        return cudaGetLastError();
    }
    if ((e = allow_smem(primary_kernel<false>, args.debug_primary_smem)) != cudaSuccess) return e;
    primary_kernel<false><<<grid, 32 * RC_PRIMARY_WARPS, args.debug_primary_smem, stream>>>(args);
    ++*launches;
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (midEvent && (e = cudaEventRecord(midEvent, stream)) != cudaSuccess) return e;
    if (!ao) return cudaSuccess;

    const uint32_t maxSeeds = args.local_rows * args.p.width;
    const uint32_t bytes = aow_smem_bytes(AO_THREADS / 32);
    if (schedule == SCHED_AO_PERSISTENT) {
        // as many CTAs as fit at once (64 registers x 128 threads: 8 per SM), never more than the list could need
        unsigned blocks = (unsigned)numSms * (unsigned)RC_AO_MIN_BLOCKS;
        const unsigned most = (maxSeeds + AO_THREADS - 1u) / AO_THREADS;
        if (blocks > most) blocks = most;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(blocks);
        cfg.blockDim = dim3(AO_THREADS);
        cfg.dynamicSmemBytes = bytes;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = midEvent ? 0 : 1;  // an event between the two launches: plain stream order
        e = cudaLaunchKernelEx(&cfg, ao_window_persistent_kernel, args);
        ++*launches;
        return e;
    }
    const unsigned blocks = (maxSeeds + AO_THREADS - 1u) / AO_THREADS;
    if (window)
        ao_window_kernel<<<blocks, AO_THREADS, bytes, stream>>>(args);
    else
        ao_cached_kernel<<<blocks, AO_THREADS, 0, stream>>>(args);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_compose(const void* gathered, void* dst, uint32_t width, uint32_t height, uint32_t worldSize,
                           uint32_t paddedRows, uint32_t bytesPerElement, cudaStream_t stream) {
    const uint64_t rowBytes = (uint64_t)width * bytesPerElement;
    if (rowBytes == 0 || height == 0) return cudaSuccess;
    const bool wide = rowBytes % 16 == 0 && ((uintptr_t)gathered % 16 == 0) && ((uintptr_t)dst % 16 == 0);
    const uint64_t words = rowBytes / (wide ? 16 : 4) * height;
    uint64_t blocks = (words + 255) / 256;
    if (blocks > 148ull * 32) blocks = 148ull * 32;  // grid-stride: 32 CTAs of 256 threads per SM
    if (wide)
        compose_kernel<uint4><<<(unsigned)blocks, 256, 0, stream>>>((const uint4*)gathered, (uint4*)dst,
                                                                    (uint32_t)(rowBytes / 16), height, worldSize, paddedRows);
    else
        compose_kernel<uint32_t><<<(unsigned)blocks, 256, 0, stream>>>((const uint32_t*)gathered, (uint32_t*)dst,
                                                                       (uint32_t)(rowBytes / 4), height, worldSize, paddedRows);
    return cudaGetLastError();
}

cudaError_t launch_pick(const DeviceWorld& world, const float* p, const float* d, int steps, int* out7, cudaStream_t stream) {
    DeviceWorld w = world;
    w.leak = 0u;  // the host's getBlock shifts signed ints: plain torus
    PickArgs a{p[0], p[1], p[2], d[0], d[1], d[2], steps};
    pick_kernel<<<1, 32, 0, stream>>>(w, a, out7);
    return cudaGetLastError();
}

cudaError_t launch_world_tops(const DeviceWorld& world, int skyShift, int* tops, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(tops, 0xFF, SKY_GRID * SKY_GRID * sizeof(int), stream);  // -1: empty
    if (e != cudaSuccess) return e;
    world_tops_kernel<<<64 * 64 * 64 / 128, 128, 0, stream>>>(world, skyShift, tops);
    return cudaGetLastError();
}

cudaError_t launch_top_table(const DeviceWorld& world, uint32_t* table, cudaStream_t stream) {
    const uint32_t cells = 1u << (3 * world.top_bits);
    top_table_kernel<<<(cells + 127u) / 128u, 128, 0, stream>>>(world, table);
    return cudaGetLastError();
}

cudaError_t launch_sky_gates(const int* tops, int skyShift, uint32_t maxSteps, uint4* gates, cudaStream_t stream) {
    sky_gate_kernel<<<4 * SKY_GRID * SKY_GRID / 128, 128, 0, stream>>>(tops, skyShift, maxSteps, gates);
    return cudaGetLastError();
}

cudaError_t launch_sync_signal(uint32_t* flag, uint32_t value, cudaStream_t stream) {
    sync_signal_kernel<<<1, 32, 0, stream>>>(flag, value);
    return cudaGetLastError();
}

cudaError_t launch_sync_wait(const uint32_t* flags, uint32_t strideWords, uint32_t count, uint32_t value, uint32_t* timedOut,
                             cudaStream_t stream) {
    if (count == 0) return cudaSuccess;
    if (count > 64) return cudaErrorInvalidValue;
    sync_wait_kernel<<<1, 64, 0, stream>>>(flags, strideWords, count, value, timedOut);
    return cudaGetLastError();
}

cudaError_t launch_lookup(const DeviceWorld& world, const int* xyz, uint32_t n, uint32_t* outWords, int* outShift,
                          cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    lookup_kernel<<<(n + 127u) / 128u, 128, 0, stream>>>(world, xyz, n, outWords, outShift);
    return cudaGetLastError();
}

cudaError_t launch_detmath(const float* in, uint32_t n, float* outSin, float* outExp, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    detmath_kernel<<<(n + 127u) / 128u, 128, 0, stream>>>(in, n, outSin, outExp);
    return cudaGetLastError();
}

}  // namespace rc
