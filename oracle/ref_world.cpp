// oracle/ref_world.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Thin extern "C" harness around the reference's OWN, UNMODIFIED CPU sources,
// compiled where they lie under /root/reference by oracle/Makefile into
// oracle/_ref/libref_world.so:
//   src/globals.cpp, src/world_gen.cpp, src/ray_caster.cpp,
//   src/voxel_data/tetrahexa_tree.cpp, src/voxel_data/voxel_allocator.cpp,
//   include/OpenSimplexNoise.cpp
// It gives the tests (a) the default world exactly as the reference host builds
// it (initTetraHexaTree() + genWorld(), src/main.cpp:181-190), in the two-pool
// wire format the shader reads (src/shaders/low_res.frag:33-39), (b) the
// reference's getBlock/putBlock/deleteBlock and CPU picking ray, and (c) a
// capture of what updateSsboData() pushes through glBufferSubData.
//
// Nothing under /root/reference is copied: the reference files are compiled in
// place; this file only calls their public functions.
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <unistd.h>

#include <vector>

#include <OpenSimplexNoise.h>

#include "globals.hpp"
#include "ray_caster.hpp"
#include "types.hpp"
#include "voxel_data/tetrahexa_tree.hpp"
#include "voxel_data/voxel_allocator.hpp"
#include "world_gen.hpp"

ShimGlHooks g_shimGlHooks = {nullptr, nullptr};
GLuint g_shimBoundSsbo = 0;
GLuint g_shimNextBufferName = 0;

namespace {

bool g_built = false;

// The reference prints progress with printf (tetrahexa_tree.cpp:29,39); keep
// library users' stdout clean while it runs.
struct StdoutMute {
    int saved;
    StdoutMute() {
        fflush(stdout);
        saved = dup(1);
        FILE* devnull = fopen("/dev/null", "w");
        if (devnull) {
            dup2(fileno(devnull), 1);
            fclose(devnull);
        }
    }
    ~StdoutMute() {
        fflush(stdout);
        if (saved >= 0) {
            dup2(saved, 1);
            close(saved);
        }
    }
};

struct UploadRecord {
    uint32_t buffer;
    uint64_t offset;
    uint64_t size;
    const void* src;
};
std::vector<UploadRecord> g_uploads;
std::vector<UploadRecord> g_allocs;

void hookBufferData(GLuint buffer, GLsizeiptr size) {
    g_allocs.push_back({buffer, 0, (uint64_t)size, nullptr});
}
void hookBufferSubData(GLuint buffer, GLintptr off, GLsizeiptr size, const void* data) {
    g_uploads.push_back({buffer, (uint64_t)off, (uint64_t)size, data});
}

uint32_t countBlocks(const BlockPtr* blocks) {
    uint32_t n = 0;
    while (n < (uint32_t)MAX_BLOCKS && blocks[n].ptr) n++;
    return n;
}

}  // namespace

extern "C" {

// Build the reference's start-up world. gen_world = 0 stops after
// initTetraHexaTree() (the 8 start-up test blocks only).
int refworld_build(int gen_world) {
    if (g_built) return 1;
    StdoutMute mute;
    DEBUG_LEVEL = 0;
    initTetraHexaTree();
    if (gen_world) genWorld();
    cameraDir = glm::normalize(cameraDir);  // src/main.cpp:195
    g_built = true;
    return 0;
}

uint32_t refworld_block_bytes(void) { return (uint32_t)BLOCK_SIZE; }
uint32_t refworld_node_bytes_used(void) { return nodeNextAllocIndex; }
uint32_t refworld_array_bytes_used(void) { return arrayNextAllocIndex; }
uint32_t refworld_num_node_blocks(void) { return countBlocks(nodeBlocks); }
uint32_t refworld_num_array_blocks(void) { return countBlocks(arrayBlocks); }
const void* refworld_node_block(uint32_t i) { return nodeBlocks[i].ptr; }
const void* refworld_array_block(uint32_t i) { return arrayBlocks[i].ptr; }
int refworld_node_block_modified(uint32_t i) { return nodeBlocks[i].modified ? 1 : 0; }
int refworld_array_block_modified(uint32_t i) { return arrayBlocks[i].modified ? 1 : 0; }
uint32_t refworld_root_index(void) { return root.index; }
int refworld_max_depth(void) { return maxDepth; }

// Flat copies of the pools == the SSBO contents after updateSsboData().
// dst must hold n_blocks * block_bytes.
void refworld_copy_nodes(void* dst) {
    uint32_t n = countBlocks(nodeBlocks);
    for (uint32_t i = 0; i < n; i++) memcpy((uint8_t*)dst + (size_t)i * BLOCK_SIZE, nodeBlocks[i].ptr, BLOCK_SIZE);
}
void refworld_copy_arrays(void* dst) {
    uint32_t n = countBlocks(arrayBlocks);
    for (uint32_t i = 0; i < n; i++) memcpy((uint8_t*)dst + (size_t)i * BLOCK_SIZE, arrayBlocks[i].ptr, BLOCK_SIZE);
}

// src/voxel_data/tetrahexa_tree.cpp:113-157. Do not call inside [768,1023]^3
// (the reference exits the process there).
void refworld_get_block(int x, int y, int z, uint64_t* color, uint32_t* flags, float* metadata) {
    Block b = getBlock(Pos{x, y, z});
    *color = b.color;
    *flags = b.flags;
    *metadata = b.metadata;
}

// Batch getBlock: xyz = 3*n ints, out = 2*n u64 {color, flags | metadata bits << 32}.
void refworld_get_block_batch(const int* xyz, uint64_t n, uint64_t* out) {
    for (uint64_t i = 0; i < n; i++) {
        Block b = getBlock(Pos{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]});
        uint32_t mbits;
        memcpy(&mbits, &b.metadata, 4);
        out[2 * i] = b.color;
        out[2 * i + 1] = (uint64_t)b.flags | ((uint64_t)mbits << 32);
    }
}

void refworld_put_block(int x, int y, int z, uint32_t flags, uint64_t color, float metadata, int level) {
    StdoutMute mute;  // deleteChildren prints unconditionally (tetrahexa_tree.cpp:160)
    putBlock(Pos{x, y, z}, Block{flags, color, metadata}, level);
}

void refworld_delete_block(int x, int y, int z, int level) {
    StdoutMute mute;
    deleteBlock(Pos{x, y, z}, level);
}

// RAY_CASTER::castRayFromCam (src/ray_caster.cpp:54-87) with the camera globals
// set from the arguments. out = {pos.xyz, lastPos.xyz, steps}.
void refworld_pick(const float* cam_pos, const float* cam_dir, int steps, int* out) {
    glm::vec3 savedPos = cameraPos, savedDir = cameraDir;
    cameraPos = glm::vec3(cam_pos[0], cam_pos[1], cam_pos[2]);
    cameraDir = glm::vec3(cam_dir[0], cam_dir[1], cam_dir[2]);
    RayResult r = RAY_CASTER::castRayFromCam(steps);
    out[0] = r.pos.x; out[1] = r.pos.y; out[2] = r.pos.z;
    out[3] = r.lastPos.x; out[4] = r.lastPos.y; out[5] = r.lastPos.z;
    out[6] = r.steps;
    cameraPos = savedPos;
    cameraDir = savedDir;
}

// Batch form used by the "reference CPU ray_caster path" baseline (BASELINE.json
// configs[0]): one castRayFromCam(steps) per direction, from one origin.
// out: 7 ints per ray. Single-threaded by construction (camera is a global).
void refworld_pick_batch(const float* cam_pos, const float* dirs, uint32_t n, int steps, int* out) {
    glm::vec3 savedPos = cameraPos, savedDir = cameraDir;
    cameraPos = glm::vec3(cam_pos[0], cam_pos[1], cam_pos[2]);
    for (uint32_t i = 0; i < n; i++) {
        cameraDir = glm::vec3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
        RayResult r = RAY_CASTER::castRayFromCam(steps);
        int* o = out + 7 * (size_t)i;
        o[0] = r.pos.x; o[1] = r.pos.y; o[2] = r.pos.z;
        o[3] = r.lastPos.x; o[4] = r.lastPos.y; o[5] = r.lastPos.z;
        o[6] = r.steps;
    }
    cameraPos = savedPos;
    cameraDir = savedDir;
}

// Defaults of src/globals.cpp:20-23 (+ normalisation of src/main.cpp:195).
void refworld_defaults(float* cam_pos, float* cam_dir, float* sun_dir) {
    cam_pos[0] = cameraPos.x; cam_pos[1] = cameraPos.y; cam_pos[2] = cameraPos.z;
    glm::vec3 d = g_built ? cameraDir : glm::normalize(cameraDir);
    cam_dir[0] = d.x; cam_dir[1] = d.y; cam_dir[2] = d.z;
    sun_dir[0] = sun.x; sun_dir[1] = sun.y; sun_dir[2] = sun.z;
}

// include/OpenSimplexNoise.cpp 2-D eval, as world_gen.cpp:22 calls it.
double refworld_noise2(int64_t seed, double x, double y) {
    OpenSimplexNoise::Noise n(seed);
    return n.eval(x, y);
}
void refworld_noise2_grid(int64_t seed, double scale, int w, int l, double* out) {
    OpenSimplexNoise::Noise n(seed);
    for (int x = 0; x < w; x++)
        for (int z = 0; z < l; z++) out[(size_t)x * l + z] = n.eval(x * scale, z * scale);
}

// src/types.hpp:6-9
uint64_t refworld_rgb_to_u64(int r, int g, int b) { return RGB_TO_U64(r, g, b); }

// Runs the reference's updateSsboData() (voxel_allocator.hpp:38-78) against the
// no-op GL shim and records every glBufferData / glBufferSubData it issues.
// Returns the number of sub-data uploads; records are fetched with
// refworld_upload_record(i, out[4]) = {buffer, offset, size, kind(0 arrays,1 nodes)}.
uint32_t refworld_run_update_ssbo(void) {
    static bool inited = false;
    StdoutMute mute;
    g_shimGlHooks.buffer_data = hookBufferData;
    g_shimGlHooks.buffer_sub_data = hookBufferSubData;
    if (!inited) {
        initVoxelDataAllocator();
        inited = true;
    }
    g_uploads.clear();
    g_allocs.clear();
    updateSsboData();
    return (uint32_t)g_uploads.size();
}
void refworld_upload_record(uint32_t i, uint64_t* out) {
    const UploadRecord& r = g_uploads[i];
    out[0] = r.buffer;
    out[1] = r.offset;
    out[2] = r.size;
    out[3] = (r.buffer == nodeSsbo) ? 1 : 0;
}
const void* refworld_upload_src(uint32_t i) { return g_uploads[i].src; }

}  // extern "C"
