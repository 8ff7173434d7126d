// rc_api.cu -- the C-ABI of include/rc.h: process-global context, world upload
// (flatten + copy), frame launch, read-back.  Host logic only; the kernels are in
// kernels.cu.  There is deliberately no CPU rendering path in this library.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <string>
#include <vector>

#include "../../include/rc.h"
#include "../../include/rc_debug.h"
#include "kernels.h"
#include "packed_world.h"
#include "worldgen.h"

namespace {

using namespace rc;

struct Context {
    bool initialized = false;
    int device = -1;
    cudaStream_t ownStream = nullptr;
    cudaStream_t stream = nullptr;
    uint32_t rank = 0, worldSize = 1;

    // world
    bool haveWorld = false;
    uint4* dNodes = nullptr;
    uint32_t* dVoxels = nullptr;
    uint4* dPalette = nullptr;
    DeviceWorld world{};
    rc_world_info info{};
    PackedWorld packed;       // host copy of the uploaded world (debug lookups, incremental updates)
    FlattenTracking tracking; // maps + pool mirror for rc_update_world (built by its first call)
    size_t capNodes = 0, capVoxels = 0, capPalette = 0;  // device array capacities (elements)
    bool hostCopyStale = false;  // rc_generate_world: g.packed.nodes / .voxels are fetched from the device on first use
    PackedWorld debugPacked;  // result of rc_debug_flatten_host (CPU-only tests)
    FlattenTracking debugTracking;

    // frame buffers (grown on demand)
    size_t pixelCapacity = 0;
    uchar4* dRgba8 = nullptr;  // the frame buffer being rendered into (= slot[cur])
    // rc_read_image_async: two RGBA8 frame buffers, so that frame N+1 renders while frame N travels to the host
    struct FrameSlot {
        uchar4* buf = nullptr;
        size_t capacity = 0;
        bool copyPending = false;
        cudaEvent_t evCopied = nullptr;
    } slot[2];
    int cur = 0;
    // rc_frame_target_*: the whole frame on the composing GPU, written directly by every rank's kernels
    struct FrameTarget {
        uint32_t width = 0, height = 0, slots = 0;
        uchar4* ptr[2] = {nullptr, nullptr};
        bool owner = false;
        uint32_t seq = 0;
        int lastSlot = -1;
    } target;
    // rc_frame_sync_*: completion flags in the frame owner's memory (one process per GPU, no collective per frame).
    // Layout (uint32 words, 32 words = 128 bytes per flag): [0] go = frames the peers may start, [32 * (1 + r)] done[r] = frames
    // rank r has finished storing, [32 * 65] timed-out waits
    struct FrameSync {
        uint32_t* block = nullptr;
        bool owner = false;
        uint32_t frames = 0;  // frames begun through rc_frame_sync_begin / end
    } fsync;
    cudaStream_t copyStream = nullptr;
    cudaEvent_t evRendered = nullptr;
    float4* dRgba32f = nullptr;
    rc_hit* dHits = nullptr;
    AoSeed* dSeeds = nullptr;
    size_t f32Capacity = 0, hitCapacity = 0, seedCapacity = 0;
    FrameCounterBlock* dCounters = nullptr;
    // sky gates (kernels.h, FrameArgs::sky_gate): column-block tops of the current world, the gate table built from them
    int* dSkyTops = nullptr;
    uint32_t* dTopTable = nullptr;  // DeviceWorld::top_table (kernels.h), TOP_TABLE_MAX_CELLS entries, rebuilt with the tops
    bool useTopTable = true;        // rc_debug_set_option "top_table"
    uint4* dSkyGate = nullptr;
    int skyShift = 0;
    uint32_t skyGateSteps = 0xFFFFFFFFu;  // max_steps the table on the device was built for (none)
    FrameCounterBlock* hCounters = nullptr;  // pinned
    int* dPick = nullptr;
    int* hPick = nullptr;  // pinned

    // last frame
    bool haveFrame = false;
    rc_frame_params frame{};
    uint32_t localRows = 0, localBands = 0;
    bool frameHasHits = false, frameHasF32 = false, frameHasKernelTimes = false;
    uint32_t hemiLoaded = 0;
    void* aoTable = nullptr;  // oriented direction table of the window AO code (device)
    uint32_t launches = 0;
    int numSms = 0;
    // how frames are scheduled on the device / which AO specialisations run (rc_debug_set_option; same pixels either way)
    uint32_t schedule = SCHED_DEFAULT, scheduleUsed = SCHED_DEFAULT, aoOpts = AO_OPT_DEFAULT, debugPrimarySmem = 0;
    cudaEvent_t evStart = nullptr, evPrimary = nullptr, evEnd = nullptr;

    // rc_render_frame_static: device address of rc_static_scratch on this device (cudaGetSymbolAddress at rc_init), the
    // ao_samples its AO table holds and the step budget its gate table was built for
    unsigned char* staticScratch = nullptr;
    uint32_t staticHemi = 0, staticGateSteps = 0xFFFFFFFFu;

    // rc_init_multi: this device may store into device 0's memory (peer access enabled), its frame target pointers are
    // device 0's (not to be freed / unmapped here), and the event the owner's stream waits for after each frame
    bool peerToOwner = false, targetAliased = false;
    cudaEvent_t evFrameDone = nullptr;
};

// One context per device.  rc_init drives one device (contexts[0]); rc_init_multi drives several from the one host thread
// the reference's render loop runs on (src/main.cpp:208-219): the public entry points at the end of this file walk the
// contexts, everything above them is written for "the current context" g.
constexpr int RC_MAX_DEVICES = 16;
Context contexts[RC_MAX_DEVICES];
Context* cur = &contexts[0];
int nContexts = 1;
#define g (*cur)
std::string lastError;
// device side of the last rc_update_world on context 0, replayed on the other devices
WorldPatch lastPatch;
int lastUpdateMode = 0;  // 0 patched in place, 1 whole world installed again

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    lastError = buf;
    return code;
}

#define CUDA_TRY(expr)                                                                         \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) return fail(RC_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(_e)); \
    } while (0)

void freeWorld() {
    if (g.dNodes) cudaFree(g.dNodes);
    if (g.dVoxels) cudaFree(g.dVoxels);
    if (g.dPalette) cudaFree(g.dPalette);
    g.dNodes = nullptr;
    g.dVoxels = nullptr;
    g.dPalette = nullptr;
    g.hostCopyStale = false;
    g.haveWorld = false;
}

void freeFrameBuffers() {
    if (g.copyStream) cudaStreamSynchronize(g.copyStream);
    g.slot[g.cur].buf = g.dRgba8;
    for (int i = 0; i < 2; i++) {
        if (g.slot[i].buf) cudaFree(g.slot[i].buf);
        g.slot[i].buf = nullptr;
        g.slot[i].capacity = 0;
        g.slot[i].copyPending = false;
    }
    if (g.dRgba32f) cudaFree(g.dRgba32f);
    if (g.dHits) cudaFree(g.dHits);
    if (g.dSeeds) cudaFree(g.dSeeds);
    g.dRgba8 = nullptr;
    g.dRgba32f = nullptr;
    g.dHits = nullptr;
    g.dSeeds = nullptr;
    g.pixelCapacity = g.f32Capacity = g.hitCapacity = g.seedCapacity = 0;
    g.haveFrame = false;
}

constexpr uint32_t SYNC_FLAG_WORDS = 32, SYNC_MAX_RANKS = 64, SYNC_BLOCK_WORDS = SYNC_FLAG_WORDS * (SYNC_MAX_RANKS + 2);

void closeFrameSync() {
    if (g.fsync.block) {
        if (g.fsync.owner) cudaFree(g.fsync.block);
        else cudaIpcCloseMemHandle(g.fsync.block);
    }
    g.fsync = Context::FrameSync();
}

void closeFrameTarget() {
    for (int i = 0; i < 2; i++) {
        if (!g.target.ptr[i]) continue;
        if (g.targetAliased) continue;  // rc_init_multi: device 0's allocation, freed by its own context
        if (g.target.owner) cudaFree(g.target.ptr[i]);
        else cudaIpcCloseMemHandle(g.target.ptr[i]);
        g.target.ptr[i] = nullptr;
    }
    g.target = Context::FrameTarget();
    g.targetAliased = false;
}

// gen_hemisphare_distrib.py:4-12 with num_pts = n, in double, rounded to fp32;
// stored (x, y, up): the script prints (x, z, y) and the shader reads it back
// with .xzy (light_scattering.frag:230).
void hemisphereTable(uint32_t n, float* out) {
    const double pi = 3.141592653589793;
    for (uint32_t i = 0; i < n; i++) {
        const double idx = (double)i + 0.5;
        const double phi = acos(1.0 - idx * 0.85 / (double)n);
        const double theta = pi * (1.0 + sqrt(5.0)) * idx;
        out[3 * i + 0] = (float)(cos(theta) * sin(phi));
        out[3 * i + 1] = (float)(sin(theta) * sin(phi));
        out[3 * i + 2] = (float)cos(phi);
    }
}

template <typename T>
int growBuffer(T*& ptr, size_t& capacity, size_t need) {
    if (need <= capacity) return RC_OK;
    if (ptr) cudaFree(ptr);
    ptr = nullptr;
    capacity = 0;
    CUDA_TRY(cudaMalloc((void**)&ptr, need * sizeof(T)));
    capacity = need;
    return RC_OK;
}

}  // namespace

// ---- rc_render_frame_static: scratch that is never allocated -----------------------------------------------------------
// Everything a frame needs besides the world, the caller's parameters and the caller's output frame, for frames of up to
// RC_STATIC_MAX_PIXELS pixels: the frame counters / queue cursors, the sky-gate table, the oriented AO direction table and
// the AO seed list.  One file-scope array in device memory, held by every GPU that loads the library until the process
// exits (12.9 MB).  In host code the array's NAME is only a stand-in: the device address comes from cudaGetSymbolAddress,
// once per device at rc_init (one_init), and that pointer is what the kernels get.
struct StaticScratch {
    FrameCounterBlock counters;                           // cleared before every frame (static_scratch_clear_kernel)
    alignas(256) uint4 skyGate[4 * SKY_GRID * SKY_GRID];  // rebuilt when the step budget or the world changes
    alignas(256) unsigned char aoTable[AO_TABLE_BYTES];   // rebuilt when ao_samples changes
    alignas(256) AoSeed seeds[RC_STATIC_MAX_PIXELS];      // written by the primary kernel before the AO kernel reads it
};
static_assert(offsetof(StaticScratch, skyGate) % 256 == 0 && offsetof(StaticScratch, aoTable) % 256 == 0 &&
                  offsetof(StaticScratch, seeds) % 256 == 0,
              "every part of the static scratch starts on a 256-byte boundary");
static __device__ __align__(256) unsigned char rc_static_scratch[sizeof(StaticScratch)];

// Device code names the array directly.  The counters and cursors must start at zero; the rest is overwritten before it
// is read.
__global__ void static_scratch_clear_kernel() {
    uint32_t* w = reinterpret_cast<uint32_t*>(rc_static_scratch + offsetof(StaticScratch, counters));
    for (uint32_t i = threadIdx.x; i < sizeof(FrameCounterBlock) / 4u; i += blockDim.x) w[i] = 0u;
}

extern "C" {

int rc_abi_version(void) { return RC_ABI_VERSION; }

#ifndef RC_BUILD_ID
#define RC_BUILD_ID "unknown"
#endif
// sha256 of kernels.cu + kernels.h + free_flight.h as compiled into this library (csrc/Makefile): ties a profiler capture
// (profiles/traffic.json records the same digest) to the kernels that actually ran
const char* rc_build_id(void) { return RC_BUILD_ID; }

const char* rc_last_error(void) { return lastError.c_str(); }

static int one_init(int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(RC_ERR_CUDA, "rc_init: no CUDA device (%s); this library has no CPU path",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail(RC_ERR_INVALID_ARGUMENT, "rc_init: device %d of %d", device, count);
    CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(RC_ERR_UNSUPPORTED, "rc_init: device %d is sm_%d%d; librc is built for sm_100a only", device, prop.major,
                    prop.minor);
    g.numSms = prop.multiProcessorCount;
    CUDA_TRY(cudaStreamCreateWithFlags(&g.ownStream, cudaStreamNonBlocking));
    g.stream = g.ownStream;
    CUDA_TRY(cudaMalloc((void**)&g.dCounters, sizeof(FrameCounterBlock)));
    CUDA_TRY(cudaMallocHost((void**)&g.hCounters, sizeof(FrameCounterBlock)));
    memset(g.hCounters, 0, sizeof(FrameCounterBlock));
    CUDA_TRY(cudaEventCreate(&g.evStart));
    CUDA_TRY(cudaEventCreate(&g.evPrimary));
    CUDA_TRY(cudaEventCreate(&g.evEnd));
    // the static scratch of rc_render_frame_static as THIS device sees it (the device is current: cudaSetDevice above)
    CUDA_TRY(cudaGetSymbolAddress((void**)&g.staticScratch, rc_static_scratch));
    g.staticHemi = 0;
    g.staticGateSteps = 0xFFFFFFFFu;
    g.device = device;
    g.rank = 0;
    g.worldSize = 1;
    g.hemiLoaded = 0;
    g.initialized = true;
    lastError.clear();
    return RC_OK;
}

static void one_shutdown(void) {
    if (!g.initialized) return;
    cudaSetDevice(g.device);
    cudaStreamSynchronize(g.stream);
    freeWorld();
    freeFrameBuffers();
    closeFrameTarget();
    closeFrameSync();
    release_hemisphere(g.aoTable);
    g.aoTable = nullptr;
    if (g.dCounters) cudaFree(g.dCounters);
    if (g.dSkyTops) cudaFree(g.dSkyTops);
    if (g.dTopTable) cudaFree(g.dTopTable);
    g.dTopTable = nullptr;
    if (g.dSkyGate) cudaFree(g.dSkyGate);
    g.dSkyTops = nullptr;
    g.dSkyGate = nullptr;
    g.skyGateSteps = 0xFFFFFFFFu;
    if (g.hCounters) cudaFreeHost(g.hCounters);
    g.dCounters = nullptr;
    g.hCounters = nullptr;
    if (g.dPick) cudaFree(g.dPick);
    if (g.hPick) cudaFreeHost(g.hPick);
    g.dPick = nullptr;
    g.hPick = nullptr;
    if (g.evFrameDone) cudaEventDestroy(g.evFrameDone);
    g.evFrameDone = nullptr;
    g.peerToOwner = false;
    if (g.evStart) cudaEventDestroy(g.evStart);
    if (g.evPrimary) cudaEventDestroy(g.evPrimary);
    if (g.evEnd) cudaEventDestroy(g.evEnd);
    g.evStart = g.evPrimary = g.evEnd = nullptr;
    if (g.copyStream) cudaStreamDestroy(g.copyStream);
    if (g.evRendered) cudaEventDestroy(g.evRendered);
    for (int i = 0; i < 2; i++) {
        if (g.slot[i].evCopied) cudaEventDestroy(g.slot[i].evCopied);
        g.slot[i].evCopied = nullptr;
    }
    g.copyStream = nullptr;
    g.evRendered = nullptr;
    g.cur = 0;
    if (g.ownStream) cudaStreamDestroy(g.ownStream);
    g.ownStream = g.stream = nullptr;
    g.packed = PackedWorld();
    g.tracking = FlattenTracking();
    g.initialized = false;
}

static int one_set_stream(void* cuda_stream) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_set_stream: call rc_init first");
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    g.stream = cuda_stream ? (cudaStream_t)cuda_stream : g.ownStream;
    return RC_OK;
}

static int one_set_shard(uint32_t rank, uint32_t world_size) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_set_shard: call rc_init first");
    if (world_size == 0 || rank >= world_size)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_set_shard: rank %u of %u", rank, world_size);
    g.rank = rank;
    g.worldSize = world_size;
    g.haveFrame = false;
    return RC_OK;
}

static void shardLayout(uint32_t height, uint32_t rank, uint32_t worldSize, uint32_t* localBands, uint32_t* paddedBands) {
    const uint32_t totalBands = (height + BAND_ROWS - 1) / BAND_ROWS;
    *localBands = totalBands > rank ? (totalBands - rank + worldSize - 1) / worldSize : 0;
    *paddedBands = (totalBands + worldSize - 1) / worldSize;
}

int rc_shard_layout(uint32_t height, uint32_t rank, uint32_t world_size, uint32_t* local_rows, uint32_t* padded_rows) {
    if (world_size == 0 || rank >= world_size || height == 0)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_shard_layout: height %u, rank %u of %u", height, rank, world_size);
    uint32_t lb, pb;
    shardLayout(height, rank, world_size, &lb, &pb);
    // rows really rendered: the frame's last band may be partial
    uint32_t rows = 0;
    for (uint32_t b = 0; b < lb; b++) {
        const uint32_t row0 = (b * world_size + rank) * BAND_ROWS;
        rows += row0 + BAND_ROWS <= height ? BAND_ROWS : height - row0;
    }
    if (local_rows) *local_rows = rows;
    if (padded_rows) *padded_rows = pb * BAND_ROWS;
    return RC_OK;
}

int rc_compose_frame(const void* gathered, void* dst, uint32_t width, uint32_t height, uint32_t world_size,
                     uint32_t bytes_per_element) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_compose_frame: call rc_init first");
    if (!gathered || !dst || width == 0 || height == 0 || world_size == 0 || bytes_per_element == 0 ||
        bytes_per_element % 4 != 0)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_compose_frame: bad arguments");
    uint32_t lb, pb;
    shardLayout(height, 0, world_size, &lb, &pb);
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(launch_compose(gathered, dst, width, height, world_size, pb * BAND_ROWS, bytes_per_element, g.stream));
    return RC_OK;
}

// Points the kernels' view at the device arrays and fills the size fields of g.info.
static void publishWorld(uint32_t maxDepth, uint64_t nNodes, uint64_t nVoxels, uint64_t nPalette) {
    g.world.nodes = g.dNodes;
    g.world.voxels = g.dVoxels;
    g.world.palette = g.dPalette;
    g.world.n_nodes = (uint32_t)nNodes;
    g.world.n_top = 0;
    g.world.root_shift = 2 * ((int)maxDepth - 2);
    g.world.wrap = (1u << (2 * (maxDepth - 1))) - 1u;
    g.world.leak = maxDepth == 6 ? 0x300u : 0u;
    memset(&g.info, 0, sizeof(g.info));
    g.info.packed_nodes = nNodes;
    g.info.packed_voxels = nVoxels;
    g.info.packed_materials = nPalette;
    g.info.packed_bytes = nNodes * sizeof(PackedNode) + nVoxels * 4 + nPalette * sizeof(PaletteEntry);
    g.info.max_depth = maxDepth;
}

// rc_generate_world leaves the world on the device only; the host copy (save, debug lookups) is made on demand.
static int ensureHostCopy() {
    if (!g.hostCopyStale) return RC_OK;
    CUDA_TRY(cudaSetDevice(g.device));
    g.packed.nodes.resize(g.info.packed_nodes);
    g.packed.voxels.resize(g.info.packed_voxels);
    CUDA_TRY(cudaMemcpyAsync(g.packed.nodes.data(), g.dNodes, g.packed.nodes.size() * sizeof(PackedNode), cudaMemcpyDeviceToHost,
                             g.stream));
    if (!g.packed.voxels.empty())
        CUDA_TRY(cudaMemcpyAsync(g.packed.voxels.data(), g.dVoxels, g.packed.voxels.size() * sizeof(uint32_t),
                                 cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    g.hostCopyStale = false;
    return RC_OK;
}

// Column-block tops of the packed tree as it stands on the device (after every upload, update, load or generation);
// stream-ordered, nothing comes back to the host.
static int refreshSky() {
    g.skyGateSteps = 0xFFFFFFFFu;
    g.staticGateSteps = 0xFFFFFFFFu;  // the static entry point's table was built from the old tops too
    {
        // top table: cells of shift max(root_shift - 4, 4), at most 64 per axis; trees of depth 2 have none
        DeviceWorld& w = g.world;
        w.top_table = nullptr;
        w.top_shift = w.root_shift - 4 > 4 ? w.root_shift - 4 : 4;
        w.top_bits = w.root_shift + 2 - w.top_shift;
        if (w.top_bits >= 0 && w.top_bits <= 6) {
            if (!g.dTopTable) CUDA_TRY(cudaMalloc((void**)&g.dTopTable, (size_t)TOP_TABLE_MAX_CELLS * sizeof(uint32_t)));
            w.top_table = g.dTopTable;
            CUDA_TRY(launch_top_table(w, g.dTopTable, g.stream));
            if (!g.useTopTable) w.top_table = nullptr;
        }
    }
    g.skyShift = (int)g.world.root_shift + 2 - 6;  // torus edge / SKY_GRID
    if (g.skyShift < 2) return RC_OK;  // a brick would straddle blocks: no gating for such a shallow tree
    if (!g.dSkyTops) CUDA_TRY(cudaMalloc((void**)&g.dSkyTops, (size_t)SKY_GRID * SKY_GRID * sizeof(int)));
    CUDA_TRY(launch_world_tops(g.world, g.skyShift, g.dSkyTops, g.stream));
    return RC_OK;
}

// Gate table for a step budget, built on the device from the tops.
static int ensureSkyGate(uint32_t maxSteps) {
    if (g.skyGateSteps == maxSteps) return RC_OK;
    if (!g.dSkyGate) CUDA_TRY(cudaMalloc((void**)&g.dSkyGate, (size_t)4 * SKY_GRID * SKY_GRID * sizeof(uint4)));
    CUDA_TRY(launch_sky_gates(g.dSkyTops, g.skyShift, maxSteps, g.dSkyGate, g.stream));
    g.skyGateSteps = maxSteps;
    return RC_OK;
}

// Replaces the device world of the current context with pw (arrays get slack for later appends) and fills g.info.
static int installDevice(uint64_t refBytes, const PackedWorld& pw, double flattenMs) {
    auto t1 = std::chrono::steady_clock::now();
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(cudaStreamSynchronize(g.stream));  // no frame may still read the old world
    freeWorld();
    g.capNodes = pw.nodes.size() + pw.nodes.size() / 4 + 4096;
    g.capVoxels = pw.voxels.size() + pw.voxels.size() / 4 + 65536;
    g.capPalette = pw.palette.size() * 2 + 256;
    CUDA_TRY(cudaMalloc((void**)&g.dNodes, g.capNodes * sizeof(PackedNode)));
    CUDA_TRY(cudaMalloc((void**)&g.dVoxels, g.capVoxels * sizeof(uint32_t)));
    CUDA_TRY(cudaMalloc((void**)&g.dPalette, g.capPalette * sizeof(PaletteEntry)));
    CUDA_TRY(cudaMemcpyAsync(g.dNodes, pw.nodes.data(), pw.nodes.size() * sizeof(PackedNode), cudaMemcpyHostToDevice,
                             g.stream));
    if (!pw.voxels.empty())
        CUDA_TRY(cudaMemcpyAsync(g.dVoxels, pw.voxels.data(), pw.voxels.size() * sizeof(uint32_t), cudaMemcpyHostToDevice,
                                 g.stream));
    if (!pw.palette.empty())
        CUDA_TRY(cudaMemcpyAsync(g.dPalette, pw.palette.data(), pw.palette.size() * sizeof(PaletteEntry),
                                 cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    auto t2 = std::chrono::steady_clock::now();

    publishWorld(pw.max_depth, pw.nodes.size(), pw.voxels.size(), pw.palette.size());
    g.world.n_top = pw.level_begin.size() > 3 ? pw.level_begin[3] : (pw.level_begin.empty() ? 0u : pw.level_begin.back());
    g.info.ref_nodes = pw.ref_nodes;
    g.info.ref_bytes = refBytes;
    g.info.null_children = pw.null_children;
    g.info.flatten_ms = flattenMs;
    g.info.upload_ms = std::chrono::duration<double, std::milli>(t2 - t1).count();
    g.hostCopyStale = false;
    g.haveWorld = true;
    g.haveFrame = false;
    return refreshSky();
}

// ... and keeps pw as the context's host copy (debug lookups, rc_save_world, the baseline of rc_update_world).
static int installWorld(uint64_t refBytes, PackedWorld&& pw, double flattenMs) {
    const int rcode = installDevice(refBytes, pw, flattenMs);
    if (rcode != RC_OK) return rcode;
    g.packed = std::move(pw);
    return RC_OK;
}

// The other devices of rc_init_multi get the world context 0 holds: same arrays, no host copy of their own.
static int installOnOtherDevices() {
    const Context& owner = contexts[0];
    for (int i = 1; i < nContexts; i++) {
        cur = &contexts[i];
        g.tracking = FlattenTracking();
        const int rcode = installDevice(owner.info.ref_bytes, owner.packed, owner.info.flatten_ms);
        if (rcode != RC_OK) {
            cur = &contexts[0];
            return rcode;
        }
        g.packed = PackedWorld();
        g.packed.max_depth = owner.packed.max_depth;
        g.hostCopyStale = true;
    }
    cur = &contexts[0];
    return RC_OK;
}

static int one_upload_world(const rc_world_desc* desc) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_upload_world: call rc_init first");
    if (!desc) return fail(RC_ERR_INVALID_ARGUMENT, "rc_upload_world: null desc");
    CUDA_TRY(cudaSetDevice(g.device));
    auto t0 = std::chrono::steady_clock::now();
    PackedWorld pw;
    std::string err;
    g.tracking = FlattenTracking();  // the maps of the previous world are void; rc_update_world rebuilds them
    int rcode = flatten_world(*desc, pw, err);
    if (rcode != RC_OK) return fail(rcode, "%s", err.c_str());
    auto t1 = std::chrono::steady_clock::now();
    return installWorld(desc->node_bytes_used + desc->array_bytes_used, std::move(pw), std::chrono::duration<double, std::milli>(t1 - t0).count());
}

// Contiguous runs of the sorted indices (gaps up to 256 nodes are copied along: fewer, larger copies).
static int uploadNodeRanges(const PackedWorld& pw, std::vector<uint32_t> idx, uint64_t limit, uint64_t* bytes) {
    std::sort(idx.begin(), idx.end());
    size_t i = 0;
    while (i < idx.size()) {
        if (idx[i] >= limit) break;  // appended nodes travel with the tail copy
        size_t j = i;
        while (j + 1 < idx.size() && idx[j + 1] < limit && idx[j + 1] - idx[j] <= 256) j++;
        const size_t first = idx[i], count = idx[j] - idx[i] + 1;
        CUDA_TRY(cudaMemcpyAsync(g.dNodes + first, pw.nodes.data() + first, count * sizeof(PackedNode),
                                 cudaMemcpyHostToDevice, g.stream));
        *bytes += count * sizeof(PackedNode);
        i = j + 1;
    }
    return RC_OK;
}

// Device side of an incremental update, on the current context: the patched nodes in place, the appended tails, the
// sky tops; ordered after the frames already enqueued on the stream.  pw = the patched host world (context 0's).
static int applyPatchDevice(const PackedWorld& pw, const WorldPatch& patch, uint64_t refBytes, uint64_t* uploadedBytes) {
    int rcode;
    CUDA_TRY(cudaSetDevice(g.device));
    if (pw.nodes.size() > g.capNodes || pw.voxels.size() > g.capVoxels || pw.palette.size() > g.capPalette) {
        // out of slack: new, larger arrays with everything copied again (still no flatten)
        lastUpdateMode = 1;
        if ((rcode = installDevice(refBytes, pw, 0.0)) != RC_OK) return rcode;
        *uploadedBytes += g.info.packed_bytes;
        return RC_OK;
    }
    if ((rcode = uploadNodeRanges(pw, patch.node_index, patch.nodes_before, uploadedBytes)) != RC_OK) return rcode;
    if (pw.nodes.size() > patch.nodes_before) {
        const size_t n = pw.nodes.size() - patch.nodes_before;
        CUDA_TRY(cudaMemcpyAsync(g.dNodes + patch.nodes_before, pw.nodes.data() + patch.nodes_before, n * sizeof(PackedNode),
                                 cudaMemcpyHostToDevice, g.stream));
        *uploadedBytes += n * sizeof(PackedNode);
    }
    if (pw.voxels.size() > patch.voxels_before) {
        const size_t n = pw.voxels.size() - patch.voxels_before;
        CUDA_TRY(cudaMemcpyAsync(g.dVoxels + patch.voxels_before, pw.voxels.data() + patch.voxels_before, n * 4,
                                 cudaMemcpyHostToDevice, g.stream));
        *uploadedBytes += n * 4;
    }
    if (pw.palette.size() > patch.palette_before) {
        const size_t n = pw.palette.size() - patch.palette_before;
        CUDA_TRY(cudaMemcpyAsync(g.dPalette + patch.palette_before, pw.palette.data() + patch.palette_before,
                                 n * sizeof(PaletteEntry), cudaMemcpyHostToDevice, g.stream));
        *uploadedBytes += n * sizeof(PaletteEntry);
    }
    CUDA_TRY(cudaStreamSynchronize(g.stream));  // the copies read pageable host vectors
    g.world.n_nodes = (uint32_t)pw.nodes.size();
    g.info.packed_nodes = pw.nodes.size();
    g.info.packed_voxels = pw.voxels.size();
    g.info.packed_materials = pw.palette.size();
    g.info.packed_bytes = pw.nodes.size() * sizeof(PackedNode) + pw.voxels.size() * 4 + pw.palette.size() * sizeof(PaletteEntry);
    g.info.ref_bytes = refBytes;
    return refreshSky();
}

static int one_update_world(const rc_world_desc* desc, const uint32_t* dirty_node_blocks, uint32_t n_dirty_node_blocks,
                    const uint32_t* dirty_array_blocks, uint32_t n_dirty_array_blocks, rc_update_info* info) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_update_world: call rc_init first");
    if (!desc) return fail(RC_ERR_INVALID_ARGUMENT, "rc_update_world: null desc");
    CUDA_TRY(cudaSetDevice(g.device));
    rc_update_info ui;
    memset(&ui, 0, sizeof(ui));
    auto t0 = std::chrono::steady_clock::now();
    std::string err;
    WorldPatch patch;
    UpdateStats st;
    int rcode = 1;
    if (g.haveWorld && g.tracking.enabled)
        rcode = update_world(*desc, dirty_node_blocks, n_dirty_node_blocks, dirty_array_blocks, n_dirty_array_blocks, g.packed,
                             g.tracking, patch, st, err);
    if (rcode < 0) {
        // g.packed may be half patched while the device still holds the old world: force a full flatten next time, and
        // have rc_save_world / the debug lookups fetch nodes and voxels back from the device (what is rendered) instead
        // of trusting the host copy
        g.tracking = FlattenTracking();
        g.packed.nodes.clear();
        g.packed.voxels.clear();
        g.hostCopyStale = true;
        return fail(rcode, "%s", err.c_str());
    }
    if (rcode == 1) {
        // full flatten with tracking: the baseline later updates diff against
        PackedWorld pw;
        FlattenTracking tr;
        rcode = flatten_world(*desc, pw, err, &tr);
        if (rcode != RC_OK) return fail(rcode, "%s", err.c_str());
        auto t1 = std::chrono::steady_clock::now();
        ui.host_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        rcode = installWorld(desc->node_bytes_used + desc->array_bytes_used, std::move(pw), ui.host_ms);
        if (rcode != RC_OK) return rcode;
        g.tracking = std::move(tr);
        lastUpdateMode = 1;
        ui.full_rebuild = 1;
        ui.uploaded_bytes = g.info.packed_bytes;
        ui.upload_ms = g.info.upload_ms;
        if (info) *info = ui;
        return RC_OK;
    }
    auto t1 = std::chrono::steady_clock::now();
    const PackedWorld& pw = g.packed;
    ui.changed_ref_nodes = st.changed_ref_nodes;
    ui.changed_ref_arrays = st.changed_ref_arrays;
    ui.reflattened = st.reflattened;
    ui.appended_nodes = pw.nodes.size() - patch.nodes_before;
    ui.appended_voxels = pw.voxels.size() - patch.voxels_before;
    ui.garbage_nodes = g.tracking.garbage_nodes;
    ui.garbage_voxels = g.tracking.garbage_voxels;

    lastPatch = patch;
    lastUpdateMode = 0;
    if ((rcode = applyPatchDevice(pw, patch, desc->node_bytes_used + desc->array_bytes_used, &ui.uploaded_bytes)) != RC_OK) return rcode;
    auto t2 = std::chrono::steady_clock::now();
    ui.host_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    ui.upload_ms = std::chrono::duration<double, std::milli>(t2 - t1).count();
    if (info) *info = ui;
    return RC_OK;
}

int rc_save_world(const char* path) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_save_world: call rc_init first");
    if (!g.haveWorld) return fail(RC_ERR_NO_WORLD, "rc_save_world: no world uploaded");
    std::string err;
    int rcode = ensureHostCopy();
    if (rcode != RC_OK) return rcode;
    rcode = save_packed(g.packed, path, err);
    if (rcode != RC_OK) return fail(rcode, "%s", err.c_str());
    return RC_OK;
}

static int one_load_world(const char* path) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_load_world: call rc_init first");
    CUDA_TRY(cudaSetDevice(g.device));
    auto t0 = std::chrono::steady_clock::now();
    PackedWorld pw;
    std::string err;
    int rcode = load_packed(pw, path, err);
    if (rcode != RC_OK) return fail(rcode, "%s", err.c_str());
    g.tracking = FlattenTracking();
    auto t1 = std::chrono::steady_clock::now();
    return installWorld(0, std::move(pw), std::chrono::duration<double, std::milli>(t1 - t0).count());
}

static int one_generate_world(const rc_gen_params* params, rc_gen_info* info) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_generate_world: call rc_init first");
    if (!params) return fail(RC_ERR_INVALID_ARGUMENT, "rc_generate_world: null params");
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(cudaStreamSynchronize(g.stream));  // no frame may still read the old world
    GeneratedWorld gw;
    std::string err;
    int rcode = generate_world(*params, g.stream, gw, err);
    if (rcode != RC_OK) {
        if (gw.d_nodes) cudaFree(gw.d_nodes);
        if (gw.d_voxels) cudaFree(gw.d_voxels);
        if (info) *info = gw.info;  // heights are known when a column is out of range
        return fail(rcode, "%s", err.c_str());
    }
    freeWorld();
    g.tracking = FlattenTracking();
    g.dNodes = (uint4*)gw.d_nodes;
    g.dVoxels = gw.d_voxels;
    g.capNodes = gw.cap_nodes;
    g.capVoxels = gw.cap_voxels;
    g.capPalette = gw.palette.size() * 2 + 256;
    CUDA_TRY(cudaMalloc((void**)&g.dPalette, g.capPalette * sizeof(PaletteEntry)));
    CUDA_TRY(cudaMemcpyAsync(g.dPalette, gw.palette.data(), gw.palette.size() * sizeof(PaletteEntry), cudaMemcpyHostToDevice,
                             g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    publishWorld(params->max_depth, gw.n_nodes, gw.n_voxels, gw.palette.size());
    g.world.n_top = gw.level_begin.size() > 3 ? gw.level_begin[3] : (gw.level_begin.empty() ? 0u : gw.level_begin.back());
    // what the reference's pools would hold for this world: a 16-byte node per packed node and per voxel,
    // a 256-byte child array per branch (src/voxel_data/types.hpp:29-47)
    g.info.ref_nodes = gw.n_nodes + gw.n_voxels;
    g.info.ref_bytes = g.info.ref_nodes * 16 + gw.n_nodes * 256;
    g.info.flatten_ms = gw.info.total_ms;
    g.packed = PackedWorld();
    g.packed.max_depth = params->max_depth;
    g.packed.palette = std::move(gw.palette);
    g.packed.level_begin = std::move(gw.level_begin);
    g.packed.ref_nodes = g.info.ref_nodes;
    g.hostCopyStale = true;
    g.haveWorld = true;
    g.haveFrame = false;
    if (info) *info = gw.info;
    return refreshSky();
}

int rc_write_ppm(const char* path) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_write_ppm: call rc_init first");
    if (!g.haveFrame) return fail(RC_ERR_NO_FRAME, "rc_write_ppm: no frame rendered");
    std::vector<uint8_t> img((size_t)g.frame.width * g.frame.height * 4, 0);
    int rcode = rc_read_image(img.data(), img.size(), RC_RGBA8);
    if (rcode != RC_OK) return rcode;
    std::string err;
    rcode = write_ppm(path, img.data(), g.frame.width, g.frame.height, err);
    if (rcode != RC_OK) return fail(rcode, "%s", err.c_str());
    return RC_OK;
}

int rc_world_info_get(rc_world_info* out) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_world_info_get: call rc_init first");
    if (!g.haveWorld) return fail(RC_ERR_NO_WORLD, "rc_world_info_get: no world uploaded");
    if (!out) return fail(RC_ERR_INVALID_ARGUMENT, "rc_world_info_get: null out");
    *out = g.info;
    return RC_OK;
}

static int one_render_frame(const rc_frame_params* params) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_render_frame: call rc_init first");
    if (!g.haveWorld) return fail(RC_ERR_NO_WORLD, "rc_render_frame: call rc_upload_world first");
    if (!params) return fail(RC_ERR_INVALID_ARGUMENT, "rc_render_frame: null params");
    const rc_frame_params& p = *params;
    if (p.width == 0 || p.height == 0 || p.width > 65536 || p.height > 65536)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_render_frame: resolution %ux%u", p.width, p.height);
    if (p.ao_samples > RC_MAX_AO_SAMPLES)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_render_frame: ao_samples %u > %u", p.ao_samples, RC_MAX_AO_SAMPLES);
    if (p.max_steps > 65535 || p.shadow_steps > 65535)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_render_frame: step limits above 65535 do not fit rc_hit");
    CUDA_TRY(cudaSetDevice(g.device));

    // rows owned by this rank: bands b with b % world == rank
    uint32_t localBands, paddedBands;
    shardLayout(p.height, g.rank, g.worldSize, &localBands, &paddedBands);
    const uint32_t localRows = localBands * BAND_ROWS;
    // every rank's buffers hold paddedBands bands, so the frame gather can use equal-size messages
    const size_t pixels = (size_t)paddedBands * BAND_ROWS * p.width;

    int rcode;
    // a frame whose asynchronous read-back was queued stays untouched: render into the other buffer, after the
    // copy that last left it (two frames back) has landed
    g.slot[g.cur].buf = g.dRgba8;
    g.slot[g.cur].capacity = g.pixelCapacity;
    if (g.slot[g.cur].copyPending) g.cur ^= 1;
    if (g.slot[g.cur].copyPending) {
        CUDA_TRY(cudaStreamWaitEvent(g.stream, g.slot[g.cur].evCopied, 0));
        g.slot[g.cur].copyPending = false;
    }
    g.dRgba8 = g.slot[g.cur].buf;
    g.pixelCapacity = g.slot[g.cur].capacity;
    if (pixels > g.pixelCapacity && g.copyStream) CUDA_TRY(cudaStreamSynchronize(g.copyStream));
    if ((rcode = growBuffer(g.dRgba8, g.pixelCapacity, pixels)) != RC_OK) return rcode;
    if (p.flags & RC_F_RGBA32F)
        if ((rcode = growBuffer(g.dRgba32f, g.f32Capacity, pixels)) != RC_OK) return rcode;
    if (p.flags & RC_F_HITS)
        if ((rcode = growBuffer(g.dHits, g.hitCapacity, pixels)) != RC_OK) return rcode;
    if (p.ao_samples > 0) {
        if ((rcode = growBuffer(g.dSeeds, g.seedCapacity, pixels)) != RC_OK) return rcode;
        if (g.hemiLoaded != p.ao_samples) {
            std::vector<float> tab(3 * (size_t)p.ao_samples);
            hemisphereTable(p.ao_samples, tab.data());
            CUDA_TRY(upload_hemisphere(tab.data(), p.ao_samples, g.stream, &g.aoTable));
            g.hemiLoaded = p.ao_samples;
        }
    }

    FrameArgs a;
    memset(&a, 0, sizeof(a));
    a.p = p;
    a.world = g.world;
    a.rank = g.rank;
    a.inv_width = 1.0f / (float)p.width;
    a.inv_height = 1.0f / (float)p.height;
    a.world_size = g.worldSize;
    {
        // free flight needs: step counts recoverable from fp32 sums (kernels.cu, loop A: max_steps <= 1024), a tree deep
        // enough for the column grid, and a camera whose y + max_steps stays on the torus (no wrap along y)
        const float oy = floorf(p.origin[1]);
        if (!(p.flags & RC_F_NO_FREE_FLIGHT) && p.max_steps <= 1024u && g.skyShift >= 2 && oy >= 0.0f &&
            oy + (float)p.max_steps <= (float)g.world.wrap) {
            if ((rcode = ensureSkyGate(p.max_steps)) != RC_OK) return rcode;
            a.sky_gate = g.dSkyGate;
            a.sky_shift = g.skyShift;
            a.shadow_gate = p.shadow_steps <= p.max_steps ? 1u : 0u;
            a.drop_gate = (p.flags & RC_F_NO_DROP_GATES) ? 0u : 1u;
            // where a free ray ends matters only to the hit records, to render_pos_data = 2 (pos.round as colour) and to
            // the highlight test (low_res.frag:347-350) -- which a ray that has taken all max_steps steps can only pass
            // if the looked-at voxel is exactly that many steps (|dx| + |dy| + |dz|) from the camera's
            long long manhattan = 0;
            for (int k = 0; k < 3; k++) manhattan += llabs((long long)p.looking_at[k] - (long long)floorf(p.origin[k]));
            a.sky_tail = ((p.flags & (RC_F_HITS | RC_F_ALWAYS_TAIL)) || p.render_pos_data != 0 || manhattan == (long long)p.max_steps) ? 1u : 0u;
        }
    }
    // column regions need the tops table (a tree deep enough for the column grid); the torus edge is checked per ray
    if (!(p.flags & RC_F_NO_COLUMN_REGIONS) && g.skyShift >= 2 && g.dSkyTops) {
        a.col_tops = g.dSkyTops;
        a.sky_shift = g.skyShift;
    }
    a.local_rows = localRows;
    a.n_local_bands = localBands;
    a.rgba8 = g.dRgba8;
    g.target.lastSlot = -1;
    if (g.target.slots && p.width == g.target.width && p.height == g.target.height) {
        g.target.lastSlot = (int)(g.target.seq++ % g.target.slots);
        a.frame_target = g.target.ptr[g.target.lastSlot];
    }
    a.rgba32f = (p.flags & RC_F_RGBA32F) ? g.dRgba32f : nullptr;
    a.hits = (p.flags & RC_F_HITS) ? g.dHits : nullptr;
    a.ao_seeds = p.ao_samples > 0 ? g.dSeeds : nullptr;
    a.counters = g.dCounters;
    a.ao_table = p.ao_samples > 0 ? g.aoTable : nullptr;
    a.ao_opts = g.aoOpts;
    a.debug_primary_smem = g.debugPrimarySmem;
    for (int k = 0; k < 3; k++) a.sun_inv[k] = 1.0f / p.sun_dir[k];  // IEEE fp32, as buildRay(sunDir) divides (low_res.frag:67)
    {
        // cross products of low_res.frag:279-283, one rounded operation at a time (this file is compiled with
        // -ffp-contract=off: no fused multiply-add), exactly as the kernel's cross3 would per pixel
        auto cross = [](const float* u, const float* v, float* r) {
            volatile float m0 = u[1] * v[2], m1 = v[1] * u[2], m2 = u[2] * v[0], m3 = v[2] * u[0], m4 = u[0] * v[1], m5 = v[0] * u[1];
            r[0] = m0 - m1;
            r[1] = m2 - m3;
            r[2] = m4 - m5;
        };
        const float yAxis[3] = {0.0f, 1.0f, 0.0f};
        cross(p.camera_dir, yAxis, a.cam_left);
        cross(p.camera_dir, a.cam_left, a.cam_up);
    }

    // per-kernel times need an event between the two launches, which also keeps the AO grid from becoming resident
    // while the primary grid drains: only on request
    const bool kernelTimes = (p.flags & RC_F_KERNEL_TIMES) != 0;
    g.launches = 0;
    CUDA_TRY(cudaMemsetAsync(g.dCounters, 0, sizeof(FrameCounterBlock), g.stream));
    CUDA_TRY(cudaEventRecord(g.evStart, g.stream));
    CUDA_TRY(launch_frame(a, g.schedule, g.numSms, kernelTimes ? g.evPrimary : nullptr, g.stream, &g.launches, &g.scheduleUsed));
    CUDA_TRY(cudaEventRecord(g.evEnd, g.stream));
    CUDA_TRY(cudaMemcpyAsync(g.hCounters, g.dCounters, sizeof(FrameCounterBlock), cudaMemcpyDeviceToHost, g.stream));

    g.frame = p;
    g.localRows = localRows;
    g.localBands = localBands;
    g.frameHasHits = (p.flags & RC_F_HITS) != 0;
    g.frameHasF32 = (p.flags & RC_F_RGBA32F) != 0;
    g.frameHasKernelTimes = kernelTimes && g.scheduleUsed < SCHED_FUSED;
    g.haveFrame = true;
    return RC_OK;
}

static int one_frame_target_create(uint32_t width, uint32_t height, uint32_t slots, void* handles_out) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_frame_target_create: call rc_init first");
    if (width == 0 || height == 0 || width > 65536 || height > 65536 || slots == 0 || slots > 2 || !handles_out)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_frame_target_create: %ux%u, %u slots", width, height, slots);
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    closeFrameTarget();
    g.target.owner = true;
    for (uint32_t i = 0; i < slots; i++) {
        CUDA_TRY(cudaMalloc((void**)&g.target.ptr[i], (size_t)width * height * 4));
        CUDA_TRY(cudaMemsetAsync(g.target.ptr[i], 0, (size_t)width * height * 4, g.stream));
        cudaIpcMemHandle_t h;
        CUDA_TRY(cudaIpcGetMemHandle(&h, g.target.ptr[i]));
        static_assert(sizeof(h) == RC_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t is 64 bytes");
        memcpy((uint8_t*)handles_out + (size_t)i * RC_IPC_HANDLE_BYTES, &h, sizeof(h));
    }
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    g.target.width = width;
    g.target.height = height;
    g.target.slots = slots;
    return RC_OK;
}

static int one_frame_target_open(uint32_t width, uint32_t height, uint32_t slots, const void* handles) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_frame_target_open: call rc_init first");
    if (width == 0 || height == 0 || slots == 0 || slots > 2 || !handles)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_frame_target_open: %ux%u, %u slots", width, height, slots);
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    closeFrameTarget();
    for (uint32_t i = 0; i < slots; i++) {
        cudaIpcMemHandle_t h;
        memcpy(&h, (const uint8_t*)handles + (size_t)i * RC_IPC_HANDLE_BYTES, sizeof(h));
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            closeFrameTarget();
            return fail(RC_ERR_CUDA, "rc_frame_target_open: cudaIpcOpenMemHandle: %s (no peer access between the two GPUs?)",
                        cudaGetErrorString(e));
        }
        g.target.ptr[i] = (uchar4*)p;
    }
    g.target.width = width;
    g.target.height = height;
    g.target.slots = slots;
    return RC_OK;
}

static int one_frame_target_close(void) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_frame_target_close: call rc_init first");
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    closeFrameTarget();
    return RC_OK;
}

void* rc_frame_target_device(uint32_t slot) {
    if (!g.initialized || !g.target.owner || slot >= g.target.slots) return nullptr;
    return g.target.ptr[slot];
}

int rc_frame_target_last_slot(void) { return g.initialized ? g.target.lastSlot : -1; }

static int one_sync(void) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_sync: call rc_init first");
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    return RC_OK;
}

// Copies this rank's rows into their place in a full bottom-up frame at dst.
// (world_size 1: one contiguous copy.)
static int copyRowsToHost(void* dst, const void* src, size_t bytesPerPixel, cudaStream_t stream = nullptr, bool wait = true) {
    const rc_frame_params& p = g.frame;
    const size_t rowBytes = (size_t)p.width * bytesPerPixel;
    if (!stream) stream = g.stream;
    if (g.worldSize == 1) {
        CUDA_TRY(cudaMemcpyAsync(dst, src, rowBytes * p.height, cudaMemcpyDeviceToHost, stream));
    } else {
        // band lb of this rank is band lb * worldSize + rank of the frame: one strided copy for the full bands,
        // one more for a partial last band
        const size_t bandBytes = (size_t)BAND_ROWS * rowBytes;
        uint32_t fullBands = 0, tailRows = 0;
        for (uint32_t lb = 0; lb < g.localBands; lb++) {
            const uint32_t row0 = (lb * g.worldSize + g.rank) * BAND_ROWS;
            if (row0 >= p.height) break;
            if (row0 + BAND_ROWS <= p.height) fullBands++;
            else tailRows = p.height - row0;
        }
        uint8_t* first = (uint8_t*)dst + (size_t)g.rank * bandBytes;
        if (fullBands)
            CUDA_TRY(cudaMemcpy2DAsync(first, bandBytes * g.worldSize, src, bandBytes, bandBytes, fullBands,
                                       cudaMemcpyDeviceToHost, stream));
        if (tailRows)
            CUDA_TRY(cudaMemcpyAsync(first + (size_t)fullBands * bandBytes * g.worldSize,
                                     (const uint8_t*)src + (size_t)fullBands * bandBytes, tailRows * rowBytes,
                                     cudaMemcpyDeviceToHost, stream));
    }
    if (wait) CUDA_TRY(cudaStreamSynchronize(stream));
    return RC_OK;
}

static int one_read_image_async(void* dst, size_t dst_bytes, int* ticket) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_read_image_async: call rc_init first");
    if (!g.haveFrame) return fail(RC_ERR_NO_FRAME, "rc_read_image_async: no frame rendered");
    if (!dst || !ticket) return fail(RC_ERR_INVALID_ARGUMENT, "rc_read_image_async: null dst / ticket");
    const size_t pixels = (size_t)g.frame.width * g.frame.height;
    if (dst_bytes < pixels * 4) return fail(RC_ERR_BUFFER_TOO_SMALL, "rc_read_image_async: need %zu bytes", pixels * 4);
    if (g.slot[g.cur].copyPending) return fail(RC_ERR_NO_FRAME, "rc_read_image_async: this frame's read-back is already queued");
    CUDA_TRY(cudaSetDevice(g.device));
    if (!g.copyStream) {
        CUDA_TRY(cudaStreamCreateWithFlags(&g.copyStream, cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&g.evRendered, cudaEventDisableTiming));
        for (int i = 0; i < 2; i++) CUDA_TRY(cudaEventCreateWithFlags(&g.slot[i].evCopied, cudaEventDisableTiming));
    }
    CUDA_TRY(cudaEventRecord(g.evRendered, g.stream));
    CUDA_TRY(cudaStreamWaitEvent(g.copyStream, g.evRendered, 0));
    int rcode = copyRowsToHost(dst, g.dRgba8, 4, g.copyStream, false);
    if (rcode != RC_OK) return rcode;
    CUDA_TRY(cudaEventRecord(g.slot[g.cur].evCopied, g.copyStream));
    g.slot[g.cur].copyPending = true;
    *ticket = g.cur;
    return RC_OK;
}

int rc_host_register(void* ptr, size_t bytes) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_host_register: call rc_init first");
    if (!ptr || bytes == 0) return fail(RC_ERR_INVALID_ARGUMENT, "rc_host_register: null / empty range");
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return RC_OK;
}

// Full memory fence for hosts that publish "frame complete" counters next to frames in shared host memory (one frame
// assembled by several processes): fence, then store the counter (release); load the counter, then fence (acquire).  A plain
// store / load is enough on x86 only.
void rc_host_fence(void) { std::atomic_thread_fence(std::memory_order_seq_cst); }

int rc_host_unregister(void* ptr) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_host_unregister: call rc_init first");
    if (!ptr) return fail(RC_ERR_INVALID_ARGUMENT, "rc_host_unregister: null");
    if (g.copyStream) CUDA_TRY(cudaStreamSynchronize(g.copyStream));
    CUDA_TRY(cudaHostUnregister(ptr));
    return RC_OK;
}

static int one_wait_image(int ticket) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_wait_image: call rc_init first");
    if (ticket < 0 || ticket > 1 || !g.slot[ticket].evCopied)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_wait_image: ticket %d was not issued by rc_read_image_async", ticket);
    CUDA_TRY(cudaEventSynchronize(g.slot[ticket].evCopied));
    return RC_OK;
}

static int one_read_image(void* dst, size_t dst_bytes, int format) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_read_image: call rc_init first");
    if (!g.haveFrame) return fail(RC_ERR_NO_FRAME, "rc_read_image: no frame rendered");
    if (!dst) return fail(RC_ERR_INVALID_ARGUMENT, "rc_read_image: null dst");
    const size_t pixels = (size_t)g.frame.width * g.frame.height;
    if (format == RC_RGBA8) {
        if (dst_bytes < pixels * 4) return fail(RC_ERR_BUFFER_TOO_SMALL, "rc_read_image: need %zu bytes", pixels * 4);
        return copyRowsToHost(dst, g.dRgba8, 4);
    } else if (format == RC_RGBA32F) {
        if (!g.frameHasF32) return fail(RC_ERR_NO_FRAME, "rc_read_image: frame was rendered without RC_F_RGBA32F");
        if (dst_bytes < pixels * 16) return fail(RC_ERR_BUFFER_TOO_SMALL, "rc_read_image: need %zu bytes", pixels * 16);
        return copyRowsToHost(dst, g.dRgba32f, 16);
    }
    return fail(RC_ERR_INVALID_ARGUMENT, "rc_read_image: unknown format %d", format);
}

static int one_read_hits(rc_hit* dst, size_t n) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_read_hits: call rc_init first");
    if (!g.haveFrame || !g.frameHasHits) return fail(RC_ERR_NO_FRAME, "rc_read_hits: no frame rendered with RC_F_HITS");
    if (!dst) return fail(RC_ERR_INVALID_ARGUMENT, "rc_read_hits: null dst");
    const size_t pixels = (size_t)g.frame.width * g.frame.height;
    if (n < pixels) return fail(RC_ERR_BUFFER_TOO_SMALL, "rc_read_hits: need %zu records", pixels);
    return copyRowsToHost(dst, g.dHits, sizeof(rc_hit));
}

static int one_frame_stats_get(rc_frame_stats* out) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_frame_stats_get: call rc_init first");
    if (!g.haveFrame) return fail(RC_ERR_NO_FRAME, "rc_frame_stats_get: no frame rendered");
    if (!out) return fail(RC_ERR_INVALID_ARGUMENT, "rc_frame_stats_get: null out");
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    memset(out, 0, sizeof(*out));
    uint64_t seeds = g.hCounters->seed_cursor.next;  // two-kernel schedules: the seed list's length
    for (uint32_t k = 0; k < COUNTER_SLOTS; k++) {
        const FrameCounters& c = g.hCounters->slot[k];
        out->primary_rays += c.primary_rays;
        out->shadow_rays += c.shadow_rays;
        out->hit_pixels += c.hit_pixels;
        out->primary_steps += c.primary_steps;
        out->free_flight_rays += c.free_flight_rays;
        seeds += c.ao_seed_count;  // fused schedules count their seeds here
    }
    out->ao_rays = seeds * g.frame.ao_samples;
    out->kernel_launches = g.launches;
    float ms = 0.f;
    if (g.frameHasKernelTimes) {
        if (cudaEventElapsedTime(&ms, g.evStart, g.evPrimary) == cudaSuccess) out->primary_ms = ms;
        if (cudaEventElapsedTime(&ms, g.evPrimary, g.evEnd) == cudaSuccess) out->ao_ms = ms;
    }
    if (cudaEventElapsedTime(&ms, g.evStart, g.evEnd) == cudaSuccess) out->total_ms = ms;
    out->schedule = g.scheduleUsed;
    return RC_OK;
}

void* rc_device_image(int format) {
    if (!g.initialized || !g.haveFrame) return nullptr;
    if (format == RC_RGBA8) return g.dRgba8;
    if (format == RC_RGBA32F && g.frameHasF32) return g.dRgba32f;
    return nullptr;
}

void* rc_device_hits(void) {
    if (!g.initialized || !g.haveFrame || !g.frameHasHits) return nullptr;
    return g.dHits;
}

int rc_pick(const float cam_pos[3], const float cam_dir[3], int steps, rc_pick_result* out) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_pick: call rc_init first");
    if (!g.haveWorld) return fail(RC_ERR_NO_WORLD, "rc_pick: call rc_upload_world first");
    if (!cam_pos || !cam_dir || !out) return fail(RC_ERR_INVALID_ARGUMENT, "rc_pick: null argument");
    CUDA_TRY(cudaSetDevice(g.device));
    if (!g.dPick) CUDA_TRY(cudaMalloc((void**)&g.dPick, 7 * sizeof(int)));
    if (!g.hPick) CUDA_TRY(cudaMallocHost((void**)&g.hPick, 7 * sizeof(int)));
    CUDA_TRY(launch_pick(g.world, cam_pos, cam_dir, steps, g.dPick, g.stream));
    CUDA_TRY(cudaMemcpyAsync(g.hPick, g.dPick, 7 * sizeof(int), cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    for (int k = 0; k < 3; k++) {
        out->pos[k] = g.hPick[k];
        out->last_pos[k] = g.hPick[3 + k];
    }
    out->steps = g.hPick[6];
    return RC_OK;
}

int rc_hemisphere_table(uint32_t n, float* out) {
    if (n == 0 || n > RC_MAX_AO_SAMPLES || !out) return fail(RC_ERR_INVALID_ARGUMENT, "rc_hemisphere_table: n=%u", n);
    hemisphereTable(n, out);
    return RC_OK;
}

// ---- debug / test entry points (declared in include/rc_debug.h) -----------------

// Packed-form lookup on a host copy: which = 0 the uploaded world, 1 the result of
// rc_debug_flatten_host. out_word = material word (0 empty), out_shift = uniform
// cell shift, out_leaf[4] = palette entry (the reference Leaf verbatim) when solid.
int rc_debug_lookup_host(int which, int32_t x, int32_t y, int32_t z, uint32_t* out_word, int32_t* out_shift,
                         uint32_t* out_leaf) {
    if (!which && ensureHostCopy() != RC_OK) return RC_ERR_CUDA;
    const PackedWorld& pw = which ? g.debugPacked : g.packed;
    if (pw.nodes.empty()) return fail(RC_ERR_NO_WORLD, "rc_debug_lookup_host: no world (which=%d)", which);
    int shift = 0;
    uint32_t w = packed_lookup(pw, x, y, z, &shift);
    if (out_word) *out_word = w;
    if (out_shift) *out_shift = shift;
    if (w && out_leaf) memcpy(out_leaf, &pw.palette[w & MATERIAL_INDEX_MASK], 16);
    return RC_OK;
}

// Batch form: xyz = 3*n ints; out_words[n]; out_leaves[4*n] (zeros when empty).
int rc_debug_lookup_host_batch(int which, const int32_t* xyz, uint64_t n, uint32_t* out_words, int32_t* out_shifts,
                               uint32_t* out_leaves) {
    if (!which && ensureHostCopy() != RC_OK) return RC_ERR_CUDA;
    const PackedWorld& pw = which ? g.debugPacked : g.packed;
    if (pw.nodes.empty()) return fail(RC_ERR_NO_WORLD, "rc_debug_lookup_host_batch: no world (which=%d)", which);
    for (uint64_t i = 0; i < n; i++) {
        int shift = 0;
        uint32_t w = packed_lookup(pw, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], &shift);
        out_words[i] = w;
        if (out_shifts) out_shifts[i] = shift;
        if (out_leaves) {
            if (w) memcpy(out_leaves + 4 * i, &pw.palette[w & MATERIAL_INDEX_MASK], 16);
            else memset(out_leaves + 4 * i, 0, 16);
        }
    }
    return RC_OK;
}

// Same lookup through the device code path, for n voxels (xyz: 3*n ints, host memory).
int rc_debug_lookup_device(const int32_t* xyz, uint32_t n, uint32_t* out_words, int32_t* out_shifts) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_debug_lookup_device: call rc_init first");
    if (!g.haveWorld) return fail(RC_ERR_NO_WORLD, "rc_debug_lookup_device: no world");
    int *dXyz = nullptr, *dShift = nullptr;
    uint32_t* dWords = nullptr;
    CUDA_TRY(cudaMalloc((void**)&dXyz, (size_t)n * 12));
    CUDA_TRY(cudaMalloc((void**)&dShift, (size_t)n * 4));
    CUDA_TRY(cudaMalloc((void**)&dWords, (size_t)n * 4));
    CUDA_TRY(cudaMemcpyAsync(dXyz, xyz, (size_t)n * 12, cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(launch_lookup(g.world, dXyz, n, dWords, dShift, g.stream));
    CUDA_TRY(cudaMemcpyAsync(out_words, dWords, (size_t)n * 4, cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaMemcpyAsync(out_shifts, dShift, (size_t)n * 4, cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    cudaFree(dXyz);
    cudaFree(dShift);
    cudaFree(dWords);
    return RC_OK;
}

// Palette entry (Leaf verbatim) of a material word of the uploaded world.
int rc_debug_palette(uint32_t word, uint32_t* out_leaf) {
    if (g.packed.nodes.empty() && !g.hostCopyStale) return fail(RC_ERR_NO_WORLD, "rc_debug_palette: no world");
    const uint32_t idx = word & MATERIAL_INDEX_MASK;
    if (idx >= g.packed.palette.size()) return fail(RC_ERR_INVALID_ARGUMENT, "rc_debug_palette: index %u", idx);
    memcpy(out_leaf, &g.packed.palette[idx], 16);
    return RC_OK;
}

// Host-only flatten (no CUDA needed): lets CPU tests exercise the flattener.
int rc_debug_flatten_host(const rc_world_desc* desc, rc_world_info* info) {
    if (!desc) return fail(RC_ERR_INVALID_ARGUMENT, "rc_debug_flatten_host: null desc");
    PackedWorld pw;
    std::string err;
    int rcode = flatten_world(*desc, pw, err);
    if (rcode != RC_OK) return fail(rcode, "%s", err.c_str());
    if (info) {
        memset(info, 0, sizeof(*info));
        info->ref_nodes = pw.ref_nodes;
        info->ref_bytes = desc->node_bytes_used + desc->array_bytes_used;
        info->packed_nodes = pw.nodes.size();
        info->packed_voxels = pw.voxels.size();
        info->packed_materials = pw.palette.size();
        info->packed_bytes = pw.nodes.size() * 16 + pw.voxels.size() * 4 + pw.palette.size() * 16;
        info->max_depth = pw.max_depth;
        info->null_children = pw.null_children;
    }
    g.debugPacked = std::move(pw);
    return RC_OK;
}

static void fillInfo(const rc_world_desc* desc, const PackedWorld& pw, rc_world_info* info) {
    memset(info, 0, sizeof(*info));
    info->ref_nodes = pw.ref_nodes;
    info->ref_bytes = desc->node_bytes_used + desc->array_bytes_used;
    info->packed_nodes = pw.nodes.size();
    info->packed_voxels = pw.voxels.size();
    info->packed_materials = pw.palette.size();
    info->packed_bytes = pw.nodes.size() * 16 + pw.voxels.size() * 4 + pw.palette.size() * 16;
    info->max_depth = pw.max_depth;
    info->null_children = pw.null_children;
}

int rc_debug_flatten_host_tracked(const rc_world_desc* desc, rc_world_info* info) {
    if (!desc) return fail(RC_ERR_INVALID_ARGUMENT, "rc_debug_flatten_host_tracked: null desc");
    PackedWorld pw;
    FlattenTracking tr;
    std::string err;
    int rcode = flatten_world(*desc, pw, err, &tr);
    if (rcode != RC_OK) return fail(rcode, "%s", err.c_str());
    if (info) fillInfo(desc, pw, info);
    g.debugPacked = std::move(pw);
    g.debugTracking = std::move(tr);
    return RC_OK;
}

int rc_debug_update_host(const rc_world_desc* desc, const uint32_t* dirty_node_blocks, uint32_t n_dirty_node_blocks,
                         const uint32_t* dirty_array_blocks, uint32_t n_dirty_array_blocks, rc_update_info* info) {
    if (!desc) return fail(RC_ERR_INVALID_ARGUMENT, "rc_debug_update_host: null desc");
    if (g.debugPacked.nodes.empty()) return fail(RC_ERR_NO_WORLD, "rc_debug_update_host: no tracked debug world");
    std::string err;
    WorldPatch patch;
    UpdateStats st;
    auto t0 = std::chrono::steady_clock::now();
    int rcode = update_world(*desc, dirty_node_blocks, n_dirty_node_blocks, dirty_array_blocks, n_dirty_array_blocks,
                             g.debugPacked, g.debugTracking, patch, st, err);
    if (rcode < 0) return fail(rcode, "%s", err.c_str());
    rc_update_info ui;
    memset(&ui, 0, sizeof(ui));
    if (rcode == 1) {
        PackedWorld pw;
        FlattenTracking tr;
        rcode = flatten_world(*desc, pw, err, &tr);
        if (rcode != RC_OK) return fail(rcode, "%s", err.c_str());
        g.debugPacked = std::move(pw);
        g.debugTracking = std::move(tr);
        ui.full_rebuild = 1;
    } else {
        ui.changed_ref_nodes = st.changed_ref_nodes;
        ui.changed_ref_arrays = st.changed_ref_arrays;
        ui.reflattened = st.reflattened;
        ui.appended_nodes = g.debugPacked.nodes.size() - patch.nodes_before;
        ui.appended_voxels = g.debugPacked.voxels.size() - patch.voxels_before;
        ui.garbage_nodes = g.debugTracking.garbage_nodes;
        ui.garbage_voxels = g.debugTracking.garbage_voxels;
    }
    ui.host_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (info) *info = ui;
    return RC_OK;
}

int rc_debug_save_host(const char* path) {
    std::string err;
    int rcode = save_packed(g.debugPacked, path, err);
    if (rcode != RC_OK) return fail(rcode, "%s", err.c_str());
    return RC_OK;
}

int rc_debug_load_host(const char* path, rc_world_info* info) {
    PackedWorld pw;
    std::string err;
    int rcode = load_packed(pw, path, err);
    if (rcode != RC_OK) return fail(rcode, "%s", err.c_str());
    if (info) {
        rc_world_desc none;
        memset(&none, 0, sizeof(none));
        fillInfo(&none, pw, info);
    }
    g.debugPacked = std::move(pw);
    g.debugTracking = FlattenTracking();
    return RC_OK;
}

// Scheduling / specialisation switches for A/B measurements and parity tests (every setting renders the same pixels):
//   "schedule"  0 two kernels, 1 persistent AO grid + programmatic dependent launch, 2 fused, 3 persistent fused,
//               4 persistent fused with the top tree levels staged in shared memory, 5 (default) 0 or 1 by frame size
//   "ao_opts"   bit 0: AO sample groups whose octant of reachable voxels is empty are not cast
static int one_debug_set_option(const char* name, int value) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_debug_set_option: call rc_init first");
    if (!name) return fail(RC_ERR_INVALID_ARGUMENT, "rc_debug_set_option: null name");
    if (!strcmp(name, "schedule")) {
        if (value < 0 || value > (int)SCHED_AUTO) return fail(RC_ERR_INVALID_ARGUMENT, "rc_debug_set_option: schedule %d", value);
        g.schedule = (uint32_t)value;
    } else if (!strcmp(name, "ao_opts")) {
        g.aoOpts = (uint32_t)value & AO_OPT_DEFAULT;
    } else if (!strcmp(name, "top_table")) {  // A/B: 0 = every descent that leaves its cached node starts at the root
        g.useTopTable = value != 0;
        if (g.haveWorld) {
            CUDA_TRY(cudaStreamSynchronize(g.stream));
            g.world.top_table = g.useTopTable && g.dTopTable && g.world.top_bits >= 0 && g.world.top_bits <= 6 ? g.dTopTable : nullptr;
        }
    } else if (!strcmp(name, "primary_smem")) {  // A/B probe: see FrameArgs::debug_primary_smem
        if (value < 0 || value > 200 * 1024) return fail(RC_ERR_INVALID_ARGUMENT, "rc_debug_set_option: primary_smem %d", value);
        g.debugPrimarySmem = (uint32_t)value;
    } else {
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_debug_set_option: unknown option %s", name);
    }
    return RC_OK;
}

// Copy-only ceiling of the read-back path: the last frame's rows (this rank's bands of a sharded frame) are copied to dst
// `repeats` times, back to back on the copy stream, exactly as rc_read_image / rc_read_image_async copy them, with no
// rendering in between; *ms = wall-clock time of the whole batch.  dst = NULL: into a page-locked frame the call allocates
// itself (cudaHostAlloc, portable) and frees again.
int rc_debug_reread_image(void* dst, size_t dst_bytes, uint32_t repeats, float* ms) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_debug_reread_image: call rc_init first");
    if (!g.haveFrame) return fail(RC_ERR_NO_FRAME, "rc_debug_reread_image: no frame rendered");
    const size_t bytes = (size_t)g.frame.width * g.frame.height * 4;
    if (dst && dst_bytes < bytes) return fail(RC_ERR_BUFFER_TOO_SMALL, "rc_debug_reread_image: need %zu bytes", bytes);
    CUDA_TRY(cudaSetDevice(g.device));
    void* own = nullptr;
    if (!dst) {
        CUDA_TRY(cudaHostAlloc(&own, bytes, cudaHostAllocPortable));
        dst = own;
    }
    cudaStream_t st = nullptr;
    CUDA_TRY(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    int rcode = copyRowsToHost(dst, g.dRgba8, 4, st, true);  // warm-up
    const auto t0 = std::chrono::steady_clock::now();
    for (uint32_t i = 0; i < repeats && rcode == RC_OK; i++) rcode = copyRowsToHost(dst, g.dRgba8, 4, st, false);
    if (rcode == RC_OK && cudaStreamSynchronize(st) != cudaSuccess) rcode = fail(RC_ERR_CUDA, "rc_debug_reread_image: copy failed");
    const auto t1 = std::chrono::steady_clock::now();
    cudaStreamDestroy(st);
    if (own) cudaFreeHost(own);
    if (ms) *ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
    return rcode;
}

int rc_debug_struct_sizes(uint32_t out[5]) {
    if (!out) return fail(RC_ERR_INVALID_ARGUMENT, "rc_debug_struct_sizes: null out");
    out[0] = (uint32_t)sizeof(rc_world_desc);
    out[1] = (uint32_t)sizeof(rc_world_info);
    out[2] = (uint32_t)sizeof(rc_frame_params);
    out[3] = (uint32_t)sizeof(rc_hit);
    out[4] = (uint32_t)sizeof(rc_frame_stats);
    return RC_OK;
}

// Device det_sin / det_exp over n floats (host pointers).
int rc_debug_detmath(const float* in, uint32_t n, float* out_sin, float* out_exp) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_debug_detmath: call rc_init first");
    float *dIn = nullptr, *dS = nullptr, *dE = nullptr;
    CUDA_TRY(cudaMalloc((void**)&dIn, (size_t)n * 4));
    CUDA_TRY(cudaMalloc((void**)&dS, (size_t)n * 4));
    CUDA_TRY(cudaMalloc((void**)&dE, (size_t)n * 4));
    CUDA_TRY(cudaMemcpyAsync(dIn, in, (size_t)n * 4, cudaMemcpyHostToDevice, g.stream));
    CUDA_TRY(launch_detmath(dIn, n, dS, dE, g.stream));
    CUDA_TRY(cudaMemcpyAsync(out_sin, dS, (size_t)n * 4, cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaMemcpyAsync(out_exp, dE, (size_t)n * 4, cudaMemcpyDeviceToHost, g.stream));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    cudaFree(dIn);
    cudaFree(dS);
    cudaFree(dE);
    return RC_OK;
}


// =====================================================================================================================
// Public entry points: rc_init drives one device; rc_init_multi several, from the caller's one thread (the reference's host
// is one process with one render loop, src/main.cpp:173-232).  With n devices the library owns the sharding: device i
// renders the 16-row bands with band % n == i of every frame (its kernels run on its own stream, all launched before any
// is waited for), every device holds the whole world, and rc_read_image / rc_read_image_async / rc_read_hits copy each
// device's bands straight to their place in the caller's frame -- one PCIe link per device, no device-side gather.  Pass
// page-locked memory (rc_host_register) for those copies to run concurrently.
// =====================================================================================================================
struct UseContext {  // selects a context for the duration of a call; the current one is contexts[0] otherwise
    explicit UseContext(int i) { cur = &contexts[i]; }
    ~UseContext() { cur = &contexts[0]; }
};

int rc_init_multi(int n_devices, const int* devices) {
    if (n_devices < 1 || n_devices > RC_MAX_DEVICES || !devices)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_init_multi: %d devices (1..%d)", n_devices, RC_MAX_DEVICES);
    // tests on a one-GPU box: RC_ALLOW_DUPLICATE_DEVICES=1 lets one device play several (every code path of the multi-device
    // entry points runs; nothing gets faster)
    const char* dup = getenv("RC_ALLOW_DUPLICATE_DEVICES");
    const bool allowDup = dup && dup[0] == '1';
    for (int i = 0; i < n_devices && !allowDup; i++)
        for (int j = 0; j < i; j++)
            if (devices[i] == devices[j]) return fail(RC_ERR_INVALID_ARGUMENT, "rc_init_multi: device %d listed twice", devices[i]);
    rc_shutdown();
    for (int i = 0; i < n_devices; i++) {
        cur = &contexts[i];
        const int rcode = one_init(devices[i]);
        if (rcode != RC_OK) {
            nContexts = i;
            cur = &contexts[0];
            const std::string why = lastError;
            rc_shutdown();
            lastError = why;
            return rcode;
        }
        g.rank = (uint32_t)i;
        g.worldSize = (uint32_t)n_devices;
    }
    nContexts = n_devices;
    // peer access to device 0 (rc_frame_target_create: every device's kernels may store into ONE frame there)
    for (int i = 1; i < n_devices; i++) {
        cur = &contexts[i];
        int can = 0;
        if (g.device == contexts[0].device) {
            g.peerToOwner = true;  // the same device (RC_ALLOW_DUPLICATE_DEVICES)
        } else if (cudaSetDevice(g.device) == cudaSuccess && cudaDeviceCanAccessPeer(&can, g.device, contexts[0].device) == cudaSuccess && can) {
            const cudaError_t e = cudaDeviceEnablePeerAccess(contexts[0].device, 0);
            if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) g.peerToOwner = true;
            (void)cudaGetLastError();
        }
        if (cudaEventCreateWithFlags(&g.evFrameDone, cudaEventDisableTiming) != cudaSuccess) g.evFrameDone = nullptr;
    }
    cur = &contexts[0];
    return RC_OK;
}

int rc_init(int device) { return rc_init_multi(1, &device); }

int rc_device_count(void) { return contexts[0].initialized ? nContexts : 0; }

void rc_shutdown(void) {
    for (int i = nContexts - 1; i >= 0; i--) {
        cur = &contexts[i];
        one_shutdown();
    }
    nContexts = 1;
    cur = &contexts[0];
}

int rc_set_stream(void* cuda_stream) {
    if (nContexts > 1) return fail(RC_ERR_UNSUPPORTED, "rc_set_stream: with rc_init_multi the library owns one stream per device");
    return one_set_stream(cuda_stream);
}

int rc_set_shard(uint32_t rank, uint32_t world_size) {
    if (nContexts > 1) return fail(RC_ERR_UNSUPPORTED, "rc_set_shard: with rc_init_multi the library shards frames over its own devices");
    return one_set_shard(rank, world_size);
}

int rc_upload_world(const rc_world_desc* desc) {
    const int rcode = one_upload_world(desc);  // flatten once, on the host
    return rcode != RC_OK ? rcode : installOnOtherDevices();
}

int rc_update_world(const rc_world_desc* desc, const uint32_t* dirty_node_blocks, uint32_t n_dirty_node_blocks,
                    const uint32_t* dirty_array_blocks, uint32_t n_dirty_array_blocks, rc_update_info* info) {
    int rcode = one_update_world(desc, dirty_node_blocks, n_dirty_node_blocks, dirty_array_blocks, n_dirty_array_blocks, info);
    if (rcode != RC_OK || nContexts == 1) return rcode;
    // the diff and the re-flatten happened once, on context 0's mirror: the other devices get the same bytes
    if (lastUpdateMode == 1) return installOnOtherDevices();
    const Context& owner = contexts[0];
    for (int i = 1; i < nContexts && rcode == RC_OK; i++) {
        UseContext use(i);
        uint64_t bytes = 0;
        rcode = applyPatchDevice(owner.packed, lastPatch, owner.info.ref_bytes, &bytes);
        if (info) info->uploaded_bytes += bytes;
    }
    return rcode;
}

int rc_load_world(const char* path) {
    const int rcode = one_load_world(path);
    return rcode != RC_OK ? rcode : installOnOtherDevices();
}

int rc_generate_world(const rc_gen_params* params, rc_gen_info* info) {
    // every device builds the world itself (10 ms for the 2048^2 world): no host copy, no transfer between devices
    int rcode = one_generate_world(params, info);
    for (int i = 1; i < nContexts && rcode == RC_OK; i++) {
        UseContext use(i);
        rcode = one_generate_world(params, nullptr);
    }
    return rcode;
}

int rc_debug_set_option(const char* name, int value) {
    int rcode = RC_OK;
    for (int i = 0; i < nContexts && rcode == RC_OK; i++) {
        UseContext use(i);
        rcode = one_debug_set_option(name, value);
    }
    return rcode;
}

int rc_render_frame(const rc_frame_params* params) {
    if (nContexts == 1) return one_render_frame(params);
    // all devices' kernels are queued before anything is waited for: they run side by side
    for (int i = 0; i < nContexts; i++) {
        UseContext use(i);
        const int rcode = one_render_frame(params);
        if (rcode != RC_OK) return rcode;
    }
    // one frame on device 0 written by every device (rc_frame_target_create): complete, in device 0's stream order, once
    // every device's kernels have finished
    if (contexts[0].target.lastSlot >= 0) {
        for (int i = 1; i < nContexts; i++) {
            Context& c = contexts[i];
            if (!c.evFrameDone || c.target.lastSlot < 0) continue;
            CUDA_TRY(cudaSetDevice(c.device));
            CUDA_TRY(cudaEventRecord(c.evFrameDone, c.stream));
            CUDA_TRY(cudaSetDevice(contexts[0].device));
            CUDA_TRY(cudaStreamWaitEvent(contexts[0].stream, c.evFrameDone, 0));
        }
    }
    return RC_OK;
}

// rc_render_frame without a single allocation (include/rc.h): the frame goes into the caller's device buffer, the scratch
// is rc_static_scratch (its device address: g.staticScratch, from cudaGetSymbolAddress at rc_init).  No counterpart in the
// reference.  Leaves the state of the last rc_render_frame (frame buffers, statistics, events) alone.
int rc_render_frame_static(const rc_frame_params* params, void* rgba8_device, size_t rgba8_bytes) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_render_frame_static: call rc_init first");
    if (nContexts > 1) return fail(RC_ERR_UNSUPPORTED, "rc_render_frame_static: one device only (rc_init, not rc_init_multi)");
    if (!g.haveWorld) return fail(RC_ERR_NO_WORLD, "rc_render_frame_static: call rc_upload_world first");
    if (!params || !rgba8_device) return fail(RC_ERR_INVALID_ARGUMENT, "rc_render_frame_static: null params / frame");
    const rc_frame_params& p = *params;
    if (p.width == 0 || p.height == 0 || p.width > 65536 || p.height > 65536)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_render_frame_static: resolution %ux%u", p.width, p.height);
    const uint64_t pixels = (uint64_t)p.width * p.height;
    if (pixels > RC_STATIC_MAX_PIXELS)
        return fail(RC_ERR_INVALID_ARGUMENT,
                    "rc_render_frame_static: %ux%u = %llu pixels, more than RC_STATIC_MAX_PIXELS = %u (the scratch of this entry "
                    "point is a fixed array; use rc_render_frame for larger frames)",
                    p.width, p.height, (unsigned long long)pixels, RC_STATIC_MAX_PIXELS);
    if (rgba8_bytes < pixels * 4)
        return fail(RC_ERR_BUFFER_TOO_SMALL, "rc_render_frame_static: the frame needs %llu bytes", (unsigned long long)(pixels * 4));
    if (p.ao_samples > RC_MAX_AO_SAMPLES)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_render_frame_static: ao_samples %u > %u", p.ao_samples, RC_MAX_AO_SAMPLES);
    if (p.max_steps > 65535 || p.shadow_steps > 65535)
        return fail(RC_ERR_INVALID_ARGUMENT, "rc_render_frame_static: step limits above 65535");
    if (p.flags & (RC_F_HITS | RC_F_RGBA32F | RC_F_KERNEL_TIMES))
        return fail(RC_ERR_UNSUPPORTED, "rc_render_frame_static: RGBA8 frames only (no RC_F_HITS / RC_F_RGBA32F / RC_F_KERNEL_TIMES)");
    if (g.worldSize != 1) return fail(RC_ERR_UNSUPPORTED, "rc_render_frame_static: unsharded frames only (rc_set_shard(0, 1))");
    if (!g.staticScratch) return fail(RC_ERR_CUDA, "rc_render_frame_static: the scratch array has no device address");
    CUDA_TRY(cudaSetDevice(g.device));
    // a DEVICE address: only offsets are taken from it on the host
    StaticScratch* const scratch = reinterpret_cast<StaticScratch*>(g.staticScratch);

    if (p.ao_samples > 0) {
        // the oriented table goes into the scratch; the __constant__ tables (c_hemi, c_aoSeg) are shared with rc_render_frame,
        // whose own oriented table then no longer matches them: it reloads on its next AO frame (hemiLoaded = 0)
        const bool constantsMatch = g.hemiLoaded == 0 || g.hemiLoaded == p.ao_samples;
        if (g.staticHemi != p.ao_samples || !constantsMatch) {
            float tab[3 * RC_MAX_AO_SAMPLES];
            hemisphereTable(p.ao_samples, tab);
            void* table = scratch->aoTable;  // non-null: upload_hemisphere fills it and allocates nothing
            CUDA_TRY(upload_hemisphere(tab, p.ao_samples, g.stream, &table));
            g.staticHemi = p.ao_samples;
            g.hemiLoaded = 0;
        }
    }

    FrameArgs a;
    memset(&a, 0, sizeof(a));
    a.p = p;
    a.world = g.world;
    a.rank = 0;
    a.world_size = 1;
    a.inv_width = 1.0f / (float)p.width;
    a.inv_height = 1.0f / (float)p.height;
    {
        // free flight: the conditions of rc_render_frame, the gate table in the scratch
        const float oy = floorf(p.origin[1]);
        if (!(p.flags & RC_F_NO_FREE_FLIGHT) && p.max_steps <= 1024u && g.skyShift >= 2 && g.dSkyTops && oy >= 0.0f &&
            oy + (float)p.max_steps <= (float)g.world.wrap) {
            if (g.staticGateSteps != p.max_steps) {
                CUDA_TRY(launch_sky_gates(g.dSkyTops, g.skyShift, p.max_steps, scratch->skyGate, g.stream));
                g.staticGateSteps = p.max_steps;
            }
            a.sky_gate = scratch->skyGate;
            a.sky_shift = g.skyShift;
            a.shadow_gate = p.shadow_steps <= p.max_steps ? 1u : 0u;
            a.drop_gate = (p.flags & RC_F_NO_DROP_GATES) ? 0u : 1u;
            long long manhattan = 0;
            for (int k = 0; k < 3; k++) manhattan += llabs((long long)p.looking_at[k] - (long long)floorf(p.origin[k]));
            a.sky_tail = ((p.flags & RC_F_ALWAYS_TAIL) || p.render_pos_data != 0 || manhattan == (long long)p.max_steps) ? 1u : 0u;
        }
    }
    if (!(p.flags & RC_F_NO_COLUMN_REGIONS) && g.skyShift >= 2 && g.dSkyTops) {
        a.col_tops = g.dSkyTops;
        a.sky_shift = g.skyShift;
    }
    a.n_local_bands = (p.height + BAND_ROWS - 1) / BAND_ROWS;
    a.local_rows = a.n_local_bands * BAND_ROWS;  // rows beyond the frame's height are never touched (py < height)
    a.rgba8 = (uchar4*)rgba8_device;
    a.ao_seeds = p.ao_samples > 0 ? scratch->seeds : nullptr;
    a.counters = &scratch->counters;
    a.ao_table = p.ao_samples > 0 ? (const void*)scratch->aoTable : nullptr;
    a.ao_opts = g.aoOpts;
    for (int k = 0; k < 3; k++) a.sun_inv[k] = 1.0f / p.sun_dir[k];
    {
        // low_res.frag:279-283, one rounded operation at a time (as in rc_render_frame)
        auto cross = [](const float* u, const float* v, float* r) {
            volatile float m0 = u[1] * v[2], m1 = v[1] * u[2], m2 = u[2] * v[0], m3 = v[2] * u[0], m4 = u[0] * v[1], m5 = v[0] * u[1];
            r[0] = m0 - m1;
            r[1] = m2 - m3;
            r[2] = m4 - m5;
        };
        const float yAxis[3] = {0.0f, 1.0f, 0.0f};
        cross(p.camera_dir, yAxis, a.cam_left);
        cross(p.camera_dir, a.cam_left, a.cam_up);
    }

    // the counters and queue cursors must start at zero; everything else in the scratch is written before it is read
    static_scratch_clear_kernel<<<1, 256, 0, g.stream>>>();
    CUDA_TRY(cudaGetLastError());
    uint32_t launches = 0, used = 0;
    CUDA_TRY(launch_frame(a, g.schedule, g.numSms, nullptr, g.stream, &launches, &used));
    return RC_OK;
}

int rc_frame_target_create(uint32_t width, uint32_t height, uint32_t slots, void* handles_out) {
    if (nContexts == 1) return one_frame_target_create(width, height, slots, handles_out);
    // device 0 owns the frame; the others reach it through peer access (no IPC handles inside one process)
    uint8_t scratch[2 * RC_IPC_HANDLE_BYTES];
    for (int i = 1; i < nContexts; i++)
        if (!contexts[i].peerToOwner)
            return fail(RC_ERR_UNSUPPORTED, "rc_frame_target_create: device %d has no peer access to device %d", contexts[i].device,
                        contexts[0].device);
    for (int i = 1; i < nContexts; i++) {
        UseContext use(i);
        CUDA_TRY(cudaSetDevice(g.device));
        CUDA_TRY(cudaStreamSynchronize(g.stream));
        closeFrameTarget();
    }
    const int rcode = one_frame_target_create(width, height, slots, handles_out ? handles_out : scratch);
    if (rcode != RC_OK) return rcode;
    for (int i = 1; i < nContexts; i++) {
        Context& c = contexts[i];
        c.target = contexts[0].target;
        c.target.owner = false;
        c.targetAliased = true;
    }
    return RC_OK;
}

int rc_frame_target_open(uint32_t width, uint32_t height, uint32_t slots, const void* handles) {
    if (nContexts > 1)
        return fail(RC_ERR_UNSUPPORTED, "rc_frame_target_open: with rc_init_multi the frame target lives on the library's device 0");
    return one_frame_target_open(width, height, slots, handles);
}

// ---- rc_frame_sync_*: frame completion across processes without a collective ------------------------------------------
int rc_frame_sync_create(void* handle_out) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_frame_sync_create: call rc_init first");
    if (nContexts > 1) return fail(RC_ERR_UNSUPPORTED, "rc_frame_sync_create: rc_init_multi orders its devices with events");
    if (!handle_out) return fail(RC_ERR_INVALID_ARGUMENT, "rc_frame_sync_create: null handle");
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    closeFrameSync();
    CUDA_TRY(cudaMalloc((void**)&g.fsync.block, SYNC_BLOCK_WORDS * 4));
    g.fsync.owner = true;
    CUDA_TRY(cudaMemset(g.fsync.block, 0, SYNC_BLOCK_WORDS * 4));
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, g.fsync.block));
    memcpy(handle_out, &h, sizeof(h));
    return RC_OK;
}

int rc_frame_sync_open(const void* handle) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_frame_sync_open: call rc_init first");
    if (nContexts > 1) return fail(RC_ERR_UNSUPPORTED, "rc_frame_sync_open: rc_init_multi orders its devices with events");
    if (!handle) return fail(RC_ERR_INVALID_ARGUMENT, "rc_frame_sync_open: null handle");
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    closeFrameSync();
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return fail(RC_ERR_CUDA, "rc_frame_sync_open: cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
    g.fsync.block = (uint32_t*)p;
    g.fsync.owner = false;
    return RC_OK;
}

int rc_frame_sync_close(void) {
    if (!g.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_frame_sync_close: call rc_init first");
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    closeFrameSync();
    return RC_OK;
}

// Frame f (0, 1, 2 ... counted per sync block) of a peer may start storing into the owner's frame target once the owner
// has allowed it: its slot is free again (whatever read the frame that was there before has finished).
//   peer : rc_frame_sync_begin()  -- its stream waits for go >= f          -> rc_render_frame -> rc_frame_sync_end()
//   owner: rc_frame_sync_allow(f + 1, stream) whenever slot (f + 1) % slots is known to be free, on the stream that frees it
//          (the copy stream behind a read-back, or its own stream when nothing reads that slot); rc_render_frame ->
//          rc_frame_sync_end()  -- its stream waits for done[r] >= f + 1 of every peer: the frame is complete from there on.
int rc_frame_sync_begin(void) {
    if (!g.initialized || !g.fsync.block) return fail(RC_ERR_NOT_INITIALIZED, "rc_frame_sync_begin: no sync block (rc_frame_sync_create / open)");
    CUDA_TRY(cudaSetDevice(g.device));
    if (!g.fsync.owner)
        CUDA_TRY(launch_sync_wait(g.fsync.block, SYNC_FLAG_WORDS, 1, g.fsync.frames, g.fsync.block + SYNC_FLAG_WORDS * (SYNC_MAX_RANKS + 1), g.stream));
    return RC_OK;
}

int rc_frame_sync_allow(uint32_t frame, void* cuda_stream) {
    if (!g.initialized || !g.fsync.block || !g.fsync.owner) return fail(RC_ERR_NOT_INITIALIZED, "rc_frame_sync_allow: not the owner of a sync block");
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(launch_sync_signal(g.fsync.block, frame, cuda_stream ? (cudaStream_t)cuda_stream : g.stream));
    return RC_OK;
}

int rc_frame_sync_end(void) {
    if (!g.initialized || !g.fsync.block) return fail(RC_ERR_NOT_INITIALIZED, "rc_frame_sync_end: no sync block (rc_frame_sync_create / open)");
    if (g.worldSize > SYNC_MAX_RANKS) return fail(RC_ERR_UNSUPPORTED, "rc_frame_sync_end: more than %u ranks", SYNC_MAX_RANKS);
    CUDA_TRY(cudaSetDevice(g.device));
    const uint32_t done = g.fsync.frames + 1u;
    if (g.fsync.owner) {
        if (g.worldSize > 1)
            CUDA_TRY(launch_sync_wait(g.fsync.block + SYNC_FLAG_WORDS * 2, SYNC_FLAG_WORDS, g.worldSize - 1, done,
                                      g.fsync.block + SYNC_FLAG_WORDS * (SYNC_MAX_RANKS + 1), g.stream));
    } else {
        CUDA_TRY(launch_sync_signal(g.fsync.block + SYNC_FLAG_WORDS * (1 + g.rank), done, g.stream));
    }
    g.fsync.frames = done;
    return RC_OK;
}

uint32_t rc_frame_sync_frames(void) { return g.initialized ? g.fsync.frames : 0u; }

// Owner: waits that gave up (a peer that never signalled) since rc_frame_sync_create; synchronises the stream.
int rc_frame_sync_timeouts(uint32_t* out) {
    if (!g.initialized || !g.fsync.block || !out) return fail(RC_ERR_NOT_INITIALIZED, "rc_frame_sync_timeouts: no sync block");
    CUDA_TRY(cudaSetDevice(g.device));
    CUDA_TRY(cudaStreamSynchronize(g.stream));
    CUDA_TRY(cudaMemcpy(out, g.fsync.block + SYNC_FLAG_WORDS * (SYNC_MAX_RANKS + 1), 4, cudaMemcpyDeviceToHost));
    return RC_OK;
}

int rc_frame_target_close(void) {
    int rcode = RC_OK;
    for (int i = nContexts - 1; i >= 0; i--) {  // the aliases first, the owner last
        UseContext use(i);
        const int r = one_frame_target_close();
        if (r != RC_OK) rcode = r;
    }
    return rcode;
}

int rc_sync(void) {
    int rcode = RC_OK;
    for (int i = 0; i < nContexts; i++) {
        UseContext use(i);
        const int r = one_sync();
        if (r != RC_OK) rcode = r;
    }
    return rcode;
}

int rc_read_image(void* dst, size_t dst_bytes, int format) {
    if (nContexts == 1) return one_read_image(dst, dst_bytes, format);
    // every device queues the copy of its bands, then all are waited for: the links work in parallel
    const Context& c0 = contexts[0];
    if (!c0.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_read_image: call rc_init first");
    if (!c0.haveFrame) return fail(RC_ERR_NO_FRAME, "rc_read_image: no frame rendered");
    if (!dst) return fail(RC_ERR_INVALID_ARGUMENT, "rc_read_image: null dst");
    const size_t pixels = (size_t)c0.frame.width * c0.frame.height;
    const size_t bpp = format == RC_RGBA8 ? 4 : (format == RC_RGBA32F ? 16 : 0);
    if (!bpp) return fail(RC_ERR_INVALID_ARGUMENT, "rc_read_image: unknown format %d", format);
    if (format == RC_RGBA32F && !c0.frameHasF32) return fail(RC_ERR_NO_FRAME, "rc_read_image: frame was rendered without RC_F_RGBA32F");
    if (dst_bytes < pixels * bpp) return fail(RC_ERR_BUFFER_TOO_SMALL, "rc_read_image: need %zu bytes", pixels * bpp);
    for (int i = 0; i < nContexts; i++) {
        UseContext use(i);
        CUDA_TRY(cudaSetDevice(g.device));
        const int rcode = copyRowsToHost(dst, format == RC_RGBA8 ? (const void*)g.dRgba8 : (const void*)g.dRgba32f, bpp, nullptr, false);
        if (rcode != RC_OK) return rcode;
    }
    return rc_sync();
}

int rc_read_hits(rc_hit* dst, size_t n) {
    if (nContexts == 1) return one_read_hits(dst, n);
    const Context& c0 = contexts[0];
    if (!c0.initialized) return fail(RC_ERR_NOT_INITIALIZED, "rc_read_hits: call rc_init first");
    if (!c0.haveFrame || !c0.frameHasHits) return fail(RC_ERR_NO_FRAME, "rc_read_hits: no frame rendered with RC_F_HITS");
    if (!dst) return fail(RC_ERR_INVALID_ARGUMENT, "rc_read_hits: null dst");
    if (n < (size_t)c0.frame.width * c0.frame.height) return fail(RC_ERR_BUFFER_TOO_SMALL, "rc_read_hits: need %zu records", (size_t)c0.frame.width * c0.frame.height);
    for (int i = 0; i < nContexts; i++) {
        UseContext use(i);
        CUDA_TRY(cudaSetDevice(g.device));
        const int rcode = copyRowsToHost(dst, g.dHits, sizeof(rc_hit), nullptr, false);
        if (rcode != RC_OK) return rcode;
    }
    return rc_sync();
}

int rc_read_image_async(void* dst, size_t dst_bytes, int* ticket) {
    int first = -1;
    for (int i = 0; i < nContexts; i++) {
        UseContext use(i);
        int t = -1;
        const int rcode = one_read_image_async(dst, dst_bytes, &t);
        if (rcode != RC_OK) return rcode;
        if (i == 0) first = t;
        else if (t != first) return fail(RC_ERR_CUDA, "rc_read_image_async: the devices' frame slots ran out of step (%d vs %d)", t, first);
    }
    if (ticket) *ticket = first;
    return RC_OK;
}

int rc_wait_image(int ticket) {
    int rcode = RC_OK;
    for (int i = 0; i < nContexts; i++) {
        UseContext use(i);
        const int r = one_wait_image(ticket);
        if (r != RC_OK) rcode = r;
    }
    return rcode;
}

int rc_frame_stats_get(rc_frame_stats* out) {
    if (nContexts == 1) return one_frame_stats_get(out);
    if (!out) return fail(RC_ERR_INVALID_ARGUMENT, "rc_frame_stats_get: null out");
    rc_frame_stats sum;
    memset(&sum, 0, sizeof(sum));
    for (int i = 0; i < nContexts; i++) {
        UseContext use(i);
        rc_frame_stats s;
        const int rcode = one_frame_stats_get(&s);
        if (rcode != RC_OK) return rcode;
        sum.primary_rays += s.primary_rays;
        sum.shadow_rays += s.shadow_rays;
        sum.ao_rays += s.ao_rays;
        sum.hit_pixels += s.hit_pixels;
        sum.primary_steps += s.primary_steps;
        sum.kernel_launches += s.kernel_launches;
        sum.free_flight_rays += s.free_flight_rays;
        // the devices run side by side: the frame takes as long as the slowest of them
        sum.primary_ms = s.primary_ms > sum.primary_ms ? s.primary_ms : sum.primary_ms;
        sum.ao_ms = s.ao_ms > sum.ao_ms ? s.ao_ms : sum.ao_ms;
        sum.total_ms = s.total_ms > sum.total_ms ? s.total_ms : sum.total_ms;
        sum.schedule = s.schedule;
    }
    *out = sum;
    return RC_OK;
}

}  // extern "C"
