/* include/voxel_host.h -- C-ABI of libvoxelhost.so: a stand-in for the HOST side
 * the reference keeps (world generation, the 64-tree block store, the CPU picking
 * ray), for machines where the reference's own host cannot be built (it needs
 * GLFW/GLEW/glm and <windows.h>: build.bat:4-17).
 *
 * It is NOT part of the drop-in boundary (that is rc.h).  bench.py, the headless
 * host (host/main_headless.cpp) and the tests use it to produce synthetic worlds
 * in the reference's pool wire format, which rc_upload_world then ingests exactly
 * as it would ingest nodeBlocks[]/arrayBlocks[] of the real host.
 *
 * Every function restates (does not copy) the behaviour of the cited reference
 * routine; tests/test_voxel_host.py holds it to the reference's own code compiled
 * unmodified (oracle/_ref): identical noise values, identical getBlock for every
 * voxel, identical flattened world.
 */
#ifndef VOXEL_HOST_H
#define VOXEL_HOST_H

#include <stdint.h>

#include "rc.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vh_world vh_world;

/* Block (src/globals.hpp:76-80). */
typedef struct vh_block {
    uint64_t color;   /* 3 x 21-bit packed colour, src/types.hpp:6-9 */
    uint32_t flags;   /* properties, src/globals.hpp:68-74 (leaf bit added by the store) */
    float metadata;
} vh_block;

#define VH_REFLECTIVE 0x2u
#define VH_REFRACTIVE 0x4u
#define VH_LUMINESCENT 0x8u
#define VH_LIQUID 0x10u

/* Empty store with a branch root (src/voxel_data/tetrahexa_tree.cpp:13-18).
 * max_depth = 6 is the reference (1024^3 torus); 7 (4096^3) and 8 are extensions. */
vh_world* vh_create(uint32_t max_depth);
void vh_destroy(vh_world* w);

/* putBlock (tetrahexa_tree.cpp:176-291). level = max_depth is one voxel, each level
 * less is a 4x larger cube. Returns 0, or -1 on bad arguments. */
int vh_put_block(vh_world* w, int32_t x, int32_t y, int32_t z, const vh_block* block, int level);

/* getBlock (tetrahexa_tree.cpp:113-157). Returns 1 and fills *out if solid, 0 if empty. */
int vh_get_block(const vh_world* w, int32_t x, int32_t y, int32_t z, vh_block* out);

/* n lookups as the shader sees them (low_res.frag:493-532): out[2i] = colour or ~0 when empty,
 * out[2i+1] = flags | metadata bits << 32 (0 when empty). */
void vh_get_block_batch(const vh_world* w, const int32_t* xyz, uint64_t n, uint64_t* out);

/* Flat copies of the two pools (what the SSBOs would hold): n_node_blocks * 4 MiB and
 * n_array_blocks * 4 MiB bytes, see vh_describe. */
void vh_copy_pools(const vh_world* w, void* nodes_out, void* arrays_out);

/* The start-up test blocks of initTetraHexaTree (tetrahexa_tree.cpp:20-27), without the
 * one at (1000,1000,1000) that the reference's root-array aliasing makes unreachable
 * (SURVEY F11), so that this world and the reference's agree voxel for voxel. */
void vh_add_startup_blocks(vh_world* w);

/* genWorld (src/world_gen.cpp:13-42) for a width x length footprint: three octaves of
 * 2-D OpenSimplex noise (seeds[3]; reference 42, 64, 100), water / grass / dirt / stone. */
int vh_generate_terrain(vh_world* w, int32_t width, int32_t length, const int64_t* seeds);

/* Fills an rc_world_desc that points into the store's blocks (valid until the next
 * mutation of w). This is the struct the real host would fill from nodeBlocks[] /
 * arrayBlocks[] / nodeNextAllocIndex / arrayNextAllocIndex. */
int vh_describe(const vh_world* w, rc_world_desc* out);

/* RGB_TO_U64 (src/types.hpp:6-9). */
uint64_t vh_rgb(int r, int g, int b);

/* cameraPos, normalize(cameraDir), sun (src/globals.cpp:20-23, src/main.cpp:195). */
void vh_default_camera(float* pos, float* dir, float* sun);

/* RAY_CASTER::castRayFromCam (src/ray_caster.cpp:54-87): out = pos.xyz, lastPos.xyz, steps left. */
void vh_pick(const vh_world* w, const float* cam_pos, const float* cam_dir, int steps, int32_t* out);

/* OpenSimplexNoise::Noise(seed).eval(x, y) (include/OpenSimplexNoise.cpp:52-208,2518-2523). */
double vh_noise2(int64_t seed, double x, double y);
void vh_noise2_grid(int64_t seed, double scale, int32_t w, int32_t l, double* out);

#ifdef __cplusplus
}
#endif
#endif /* VOXEL_HOST_H */
