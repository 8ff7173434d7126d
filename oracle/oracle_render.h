/* oracle/oracle_render.h -- TEST INFRASTRUCTURE ONLY.
 *
 * C interface of the CPU oracle: a plain IEEE-fp32 restatement of the
 * reference's full-frame renderer src/shaders/low_res.frag (the only
 * implementation of the traversal the reference has; GLSL cannot run on this
 * box).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  The product (librc.so) never does.
 *
 * PARITY STATUS: pinned against the reference's shader SOURCE, unpinned at the GL
 * driver.  oracle/ref_shader.cpp compiles src/shaders/low_res.frag itself as C++
 * (oracle/_ref/libref_shader.so) and tests/test_ref_shader_cpu.py holds this
 * restatement to it float for float (FragColor, the pos.exact / pos.round debug
 * outputs, the globals' final state) over the baseline frames, a camera sweep and
 * a material-zoo world.  The world it traverses is built by the reference's own
 * unmodified CPU code (oracle/_ref/libref_world.so); its getBlock state machine
 * is checked voxel for voxel against the reference's getBlock, its DDA against
 * the reference's test.cpp scenario and RAY_CASTER::castRayFromCam.  What a GLSL
 * compiler would do to the shader's float arithmetic (contraction, sin / exp /
 * normalize) is "unavailable on this box" and stays unpinned; the reference holds
 * no golden frames.  The AO pass has no reference implementation: its semantics
 * are defined by this build.
 */
#ifndef ORACLE_RENDER_H
#define ORACLE_RENDER_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* The two SSBOs of low_res.frag:33-39, as flat arrays. */
typedef struct OracleWorld {
    const uint64_t* nodes;   /* 2 x u64 per node (u64vec2 nodes[]) */
    uint64_t n_nodes;
    const uint32_t* arrays;  /* 64 x u32 per array (uint[64] children_array[]) */
    uint64_t n_arrays;
    int32_t max_depth;       /* MAX_DEPTH (low_res.frag:110) = 6; 7 = declared extension */
    int32_t _pad;
} OracleWorld;

/* Uniforms of low_res.frag:16-27 + the shader's compile-time constants. */
typedef struct OracleFrame {
    float origin[3];        /* trueOrigin */
    float camera_dir[3];    /* cameraDir (normalised by the host) */
    float sun_dir[3];       /* sunDir */
    float mouse_pos[2];     /* mousePos */
    float proj_plane[2];    /* projPlaneSize */
    float delta_time;       /* deltaTime */
    float ao_strength;      /* A10 (defined by this build) */
    int32_t looking_at[3];  /* lookingAtBlock */
    int32_t render_pos_data;/* renderPosData */
    uint32_t width, height; /* resolution */
    uint32_t max_steps;     /* numSteps = 300 (low_res.frag:310) */
    uint32_t shadow_steps;  /* 75 (low_res.frag:382) */
    uint32_t ao_samples;    /* 0 = AO off */
    uint32_t ao_steps;      /* 5 (light_scattering.frag:231) */
    uint32_t flags;         /* ORACLE_F_* */
    uint32_t row_begin, row_end; /* rows [row_begin,row_end) are rendered (bounded samples) */
    int32_t threads;        /* OpenMP threads, <=0: all */
    uint32_t band_rows;     /* with band_stride > 1: only rows with (row / band_rows) % band_stride == band_offset */
    uint32_t band_stride;   /* 0 or 1: every row (multi-GPU band sharding / bounded CPU samples) */
    uint32_t band_offset;
} OracleFrame;

#define ORACLE_F_CROSSHAIR 1u  /* evaluate the crosshair test (low_res.frag:259-262) */
#define ORACLE_F_SHADOWS   2u  /* cast the sun shadow ray (low_res.frag:373-391) */

/* Per-pixel parity record. Byte layout shared with rc_hit (include/rc.h). */
typedef struct OracleHit {
    int32_t x, y, z;        /* pos.round when the primary loops exit */
    uint16_t steps;         /* nextIntersectDDA calls in the primary phase */
    uint16_t shadow_steps;  /* nextIntersectDDA calls of the shadow ray */
    uint8_t face;           /* lastHit at primary exit: 0 x, 1 y, 2 z */
    uint8_t kind;           /* ORACLE_KIND_* */
    uint8_t flags;          /* ORACLE_HF_* */
    uint8_t ao_occluded;    /* occluded AO rays */
} OracleHit;

enum {
    ORACLE_KIND_SKY = 0,
    ORACLE_KIND_HIT = 1,
    ORACLE_KIND_CAMERA_INSIDE = 2,
    ORACLE_KIND_CROSSHAIR = 3,
    ORACLE_KIND_HIGHLIGHT = 4,
    ORACLE_KIND_DEBUG = 5
};
#define ORACLE_HF_SHADOWED   0x01u
#define ORACLE_HF_REFLECTED  0x02u
#define ORACLE_HF_REFRACTED  0x04u  /* currentRefractiveIndex != 1.0 at exit */
#define ORACLE_HF_FACING_SUN 0x08u
#define ORACLE_HF_STEPX_NEG  0x10u
#define ORACLE_HF_STEPY_NEG  0x20u
#define ORACLE_HF_STEPZ_NEG  0x40u

/* Reference-algorithm memory operations, for the roofline's ALGORITHMIC bytes
 * (SURVEY 8d): 16 B per nodes[] read + 4 B per children_array[][] read. */
typedef struct OracleCounters {
    uint64_t node_reads[3];   /* [0] primary, [1] shadow, [2] AO */
    uint64_t array_reads[3];
    uint64_t dda_steps[3];
    uint64_t rays[3];
    uint64_t kind_pixels[6];
} OracleCounters;

/* Renders rows [row_begin,row_end). Any output pointer may be NULL.
 * hits/rgba8/rgba32f are full-frame buffers (width*height entries), row 0 = bottom
 * (gl_FragCoord convention). Returns 0, or <0 on malformed input. */
int oracle_render(const OracleWorld* world, const OracleFrame* frame, OracleHit* hits, uint8_t* rgba8,
                  float* rgba32f, OracleCounters* counters);

/* getBlock of the shader (stateless form: fresh stack per call). out = {block.x, block.y}. */
void oracle_get_block(const OracleWorld* world, int32_t x, int32_t y, int32_t z, uint64_t* out);

void oracle_get_block_batch(const OracleWorld* world, const int32_t* xyz, uint64_t n, uint64_t* out);

/* getMortonPos (src/main.cpp:45-66) generalised to max_depth-1 two-bit groups per axis. */
uint64_t oracle_morton(int32_t x, int32_t y, int32_t z, int32_t max_depth);

/* The literal 10-bit spread of the reference (masks 0x030000FF, 0x0300F00F, 0x030C30C3). */
uint32_t oracle_morton_ref10(int32_t x, int32_t y, int32_t z);

/* Bare DDA (no tree): n steps from origin along dir; out = {round.xyz, lastHit}.
 * use_double != 0 follows src/ray_caster.cpp / test.cpp (double deltas, trunc);
 * 0 follows low_res.frag (fp32 deltas, floor). */
void oracle_dda_walk(const float* origin, const float* dir, uint32_t n, int use_double, int32_t* out);

/* Deterministic elementary functions shared BY SPECIFICATION with the CUDA kernels
 * (DESIGN.md "deterministic transcendentals"); exposed so tests can compare them
 * with libm and with the device versions. */
float oracle_sinf(float x);
float oracle_expf(float x);

/* 64-direction (or n) hemisphere table, gen_hemisphare_distrib.py:4-12 restated in
 * double and rounded to fp32; out holds 3*n floats in tangent space (x, y, up). */
void oracle_hemisphere(uint32_t n, float* out);

/* OpenMP threads the last oracle_render really ran on (OMP_NUM_THREADS or a launcher such as torchrun can give fewer
 * than asked for); bench.py reports this as cpu_baseline.cores. */
int oracle_last_threads(void);

/* Lets tests feed the exact table the product uses instead of the libm-computed one. */
void oracle_set_hemisphere(uint32_t n, const float* dirs);

#ifdef __cplusplus
}
#endif
#endif
