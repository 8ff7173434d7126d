// worldgen.cu -- genWorld() on the device, straight into the packed format.
//
// The reference builds its world with one putBlock() per voxel on one host core
// (src/world_gen.cpp:13-42: three octaves of 2-D OpenSimplex noise per column, then water / grass /
// dirt / stone down the column) after initTetraHexaTree()'s start-up test blocks
// (src/voxel_data/tetrahexa_tree.cpp:20-27).  Here the same world is produced without the host tree:
//   1. one thread per column evaluates the three noise octaves in double precision (simplex2.h, the
//      source the host stand-in uses: identical doubles) -> height map;
//   2. the world is a pure function voxel(x,y,z | height map, start-up blocks): one thread per 4^3
//      brick of the terrain's bounding box computes its 64-bit occupancy mask;
//   3. non-empty bricks get a hierarchical key (6 bits per tree level, child index z<<4|y<<2|x like
//      tetrahexa_tree.cpp:127-129), are radix-sorted, and the tree levels are built bottom-up by
//      run-length grouping of key prefixes: siblings end up adjacent and in child order, which is what
//      popcount addressing needs;
//   4. one thread per brick writes its voxel words.
// The result answers every lookup like the host-built tree (tests: voxel for voxel against the
// reference's own getBlock; frames hash-identical to the oracle's goldens).
#include <cub/cub.cuh>
#include <stdio.h>
#include <string.h>

#include <chrono>

#include <vector>

#include "kernels.h"
#include "packed_world.h"
#include "simplex2.h"
#include "worldgen.h"

namespace rc {
namespace {

using rc_noise::Simplex2;

struct GenExtra {
    int x0, y0, z0, size;
    uint32_t word;
};
struct GenConsts {
    Simplex2 noise[3];
    int width, length;
    int nbx, nby, nbz;  // bricks of the terrain box
    int levels;         // key groups = max_depth - 2
    uint32_t wWater, wGrass, wDirt, wStone;
    int nExtras;
    GenExtra extras[8];
};
__constant__ GenConsts c_gen;

// world_gen.cpp:22
__global__ void heights_kernel(int* __restrict__ h) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c_gen.width * c_gen.length) return;
    const int x = i / c_gen.length, z = i % c_gen.length;
    const double a = round(c_gen.noise[0].eval(x * 0.005, z * 0.005) * 30);
    const double b = round(c_gen.noise[1].eval(x * 0.05, z * 0.05) * 5);
    const double c = round(c_gen.noise[2].eval(x * 0.1, z * 0.1) * 3);
    h[i] = (int)(a + b + c + 32);
}

// Material word of voxel (x,y,z), 0 = empty: world_gen.cpp:24-39 over the start-up blocks
// (terrain is put later, so it wins where both write).  *src = start-up block that wrote it, -1 = terrain.
__device__ __forceinline__ uint32_t gen_voxel(const int* __restrict__ h, int x, int y, int z, int* src) {
    *src = -1;
    if (x < c_gen.width && z < c_gen.length) {
        const int top = h[x * c_gen.length + z];
        if (top < 20 && y > top && y <= 20) return c_gen.wWater;
        if (y == top) return top < 20 ? c_gen.wDirt : c_gen.wGrass;
        if (y > 0 && y < top) return y >= top - 3 ? c_gen.wDirt : c_gen.wStone;
    }
    uint32_t w = 0u;
    for (int e = 0; e < c_gen.nExtras; e++) {
        const GenExtra& b = c_gen.extras[e];
        if (x >= b.x0 && x < b.x0 + b.size && y >= b.y0 && y < b.y0 + b.size && z >= b.z0 && z < b.z0 + b.size) {
            w = b.word;
            *src = e;
        }
    }
    return w;
}

// A brick whose 64 voxels all come from one start-up block of edge >= 4 was never split by a later
// putBlock: the host tree holds it as ONE leaf a level up (tetrahexa_tree.cpp:197-217), the packed form
// as PK_LEAF.  Returns the occupancy mask; *uniform = that case.
__device__ __forceinline__ unsigned long long brick_mask(const int* __restrict__ h, int bx, int by, int bz, bool* uniform) {
    unsigned long long m = 0;
    int first = -2;
    bool same = true;
    for (int i = 0; i < 64; i++) {
        int src;
        const uint32_t w = gen_voxel(h, 4 * bx + (i & 3), 4 * by + ((i >> 2) & 3), 4 * bz + (i >> 4), &src);
        if (w != 0u) m |= 1ull << i;
        if (i == 0) first = src;
        same = same && w != 0u && src == first;
    }
    *uniform = same && first >= 0 && c_gen.extras[first].size >= 4;
    return m;
}

__device__ __forceinline__ unsigned long long brick_key(int bx, int by, int bz, int levels) {
    unsigned long long key = 0;
    for (int l = levels - 1; l >= 0; l--) {
        const unsigned child = (((bz >> (2 * l)) & 3) << 4) | (((by >> (2 * l)) & 3) << 2) | ((bx >> (2 * l)) & 3);
        key = (key << 6) | child;
    }
    return key;
}
__device__ __forceinline__ void key_to_brick(unsigned long long key, int levels, int& bx, int& by, int& bz) {
    bx = by = bz = 0;
    for (int l = 0; l < levels; l++) {
        const unsigned child = (unsigned)(key >> (6 * l)) & 63u;
        bx |= (int)(child & 3u) << (2 * l);
        by |= (int)((child >> 2) & 3u) << (2 * l);
        bz |= (int)((child >> 4) & 3u) << (2 * l);
    }
}

// candidate c: a brick of the terrain box, or (c >= nBox) a brick of a start-up block outside it.
// masks[c] = 0 with flags[c] = 1 marks a uniform cell.
__global__ void brick_mask_kernel(const int* __restrict__ h, const int4* __restrict__ extraBricks, int nBox, int nCand,
                                  unsigned long long* __restrict__ keys, unsigned long long* __restrict__ masks,
                                  uint32_t* __restrict__ flags) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nCand) return;
    int bx, by, bz;
    if (c < nBox) {
        bz = c % c_gen.nbz;
        by = (c / c_gen.nbz) % c_gen.nby;
        bx = c / (c_gen.nbz * c_gen.nby);
    } else {
        const int4 b = extraBricks[c - nBox];
        bx = b.x, by = b.y, bz = b.z;
    }
    bool uniform;
    const unsigned long long m = brick_mask(h, bx, by, bz, &uniform);
    keys[c] = brick_key(bx, by, bz, c_gen.levels);
    masks[c] = uniform ? 0ull : m;
    flags[c] = m != 0ull ? 1u : 0u;
}

__global__ void compact_kernel(const unsigned long long* __restrict__ keys, const unsigned long long* __restrict__ masks,
                               const uint32_t* __restrict__ flags, const uint32_t* __restrict__ pos, int n,
                               unsigned long long* __restrict__ outKeys, unsigned long long* __restrict__ outMasks) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && flags[i]) {
        outKeys[pos[i]] = keys[i];
        outMasks[pos[i]] = masks[i];
    }
}
// heads of runs of equal (key >> shift)
__global__ void flag_heads_kernel(const unsigned long long* __restrict__ keys, int n, int shift, uint32_t* __restrict__ flags) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = (i == 0 || (keys[i] >> shift) != (keys[i - 1] >> shift)) ? 1u : 0u;
}
__global__ void popc_kernel(const unsigned long long* __restrict__ masks, int n, uint32_t* __restrict__ counts) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) counts[i] = (uint32_t)__popcll(masks[i]);
}
// pass 1, per level: keys of the parents (one per run of children with equal prefix)
__global__ void parent_keys_kernel(const unsigned long long* __restrict__ childKeys, const uint32_t* __restrict__ heads,
                                   const uint32_t* __restrict__ parentIdx, int n, unsigned long long* __restrict__ parentKeys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && heads[i]) parentKeys[parentIdx[i]] = childKeys[i] >> 6;
}
// pass 2, per level, once every level's offset is known: children (sorted keys at one depth) -> parent nodes
__global__ void link_level_kernel(const unsigned long long* __restrict__ childKeys, const uint32_t* __restrict__ heads,
                                  const uint32_t* __restrict__ parentIdx, int n, uint32_t childOffset, uint32_t parentOffset,
                                  PackedNode* __restrict__ nodes) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // parentIdx is the exclusive scan of heads: a head's own entry is its parent's index, a follower's is one more
    const uint32_t p = parentIdx[i] - (heads[i] ? 0u : 1u);
    PackedNode* node = nodes + parentOffset + p;
    atomicOr((unsigned long long*)&node->mask, 1ull << (childKeys[i] & 63ull));
    if (heads[i]) {
        node->base = childOffset + (uint32_t)i;
        node->kind = PK_INNER;
    }
}
__global__ void brick_nodes_kernel(const int* __restrict__ h, const unsigned long long* __restrict__ keys,
                                   const unsigned long long* __restrict__ masks, const uint32_t* __restrict__ voxBase, int n,
                                   uint32_t nodeOffset, PackedNode* __restrict__ nodes, uint32_t* __restrict__ voxels) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int bx, by, bz, src;
    key_to_brick(keys[i], c_gen.levels, bx, by, bz);
    const unsigned long long m = masks[i];
    if (m == 0ull) {  // uniform cell
        const uint32_t w = gen_voxel(h, 4 * bx, 4 * by, 4 * bz, &src);
        nodes[nodeOffset + i] = PackedNode{0ull, w, PK_LEAF | (((w >> MATERIAL_FLAG_SHIFT) & 0x10u) ? PK_HAS_LIQUID : 0u)};
        return;
    }
    uint32_t out = voxBase[i], liquid = 0u;
    for (int b = 0; b < 64; b++) {
        if (!((m >> b) & 1ull)) continue;
        const uint32_t w = gen_voxel(h, 4 * bx + (b & 3), 4 * by + ((b >> 2) & 3), 4 * bz + (b >> 4), &src);
        if ((w >> MATERIAL_FLAG_SHIFT) & 0x10u) liquid = PK_HAS_LIQUID;
        voxels[out++] = w;
    }
    nodes[nodeOffset + i] = PackedNode{m, voxBase[i], PK_BRICK | liquid};
}

#define GEN_TRY(expr)                                                                  \
    do {                                                                               \
        cudaError_t _e = (expr);                                                       \
        if (_e != cudaSuccess) {                                                       \
            err = std::string(#expr) + ": " + cudaGetErrorString(_e);                  \
            return RC_ERR_CUDA;                                                        \
        }                                                                              \
    } while (0)

struct DeviceBuf {
    void* p = nullptr;
    ~DeviceBuf() {
        if (p) cudaFree(p);
    }
    cudaError_t alloc(size_t bytes) {
        if (p) cudaFree(p);
        p = nullptr;
        return cudaMalloc(&p, bytes ? bytes : 16);
    }
    template <typename T>
    T* as() const {
        return (T*)p;
    }
};

inline int blocksFor(size_t n) { return (int)((n + 255) / 256); }

}  // namespace

namespace {

// convertScale (src/types.hpp:8): (u64)((float)x / 255.0 * RGB_RANGE) & RGB_MASK, in double
uint64_t pack_rgb(int r, int g, int b) {
    auto scale = [](int c) -> uint64_t { return (uint64_t)((double)(float)c / 255.0 * 2097151) & 2097151ull; };
    return (scale(r) << 42) | (scale(g) << 21) | scale(b);
}

// palette word of a material, adding it on first use (the flattener's Palette::word, flatten.cpp)
uint32_t material_word(std::vector<PaletteEntry>& palette, uint64_t color, uint32_t blockFlags) {
    const uint32_t flags = blockFlags | 1u;  // leaf bit, tetrahexa_tree.cpp:203
    uint32_t idx = 0;
    for (; idx < palette.size(); idx++)
        if (palette[idx].color == color && palette[idx].flags == flags && palette[idx].metadata_bits == 0u) break;
    if (idx == palette.size()) palette.push_back(PaletteEntry{color, flags, 0u});
    return idx | ((flags & 0xFFu) << MATERIAL_FLAG_SHIFT);
}

struct Scratch {
    DeviceBuf buf;
    size_t bytes = 0;
    cudaError_t need(size_t n) {
        if (n <= bytes) return cudaSuccess;
        cudaError_t e = buf.alloc(n);
        bytes = e == cudaSuccess ? n : 0;
        return e;
    }
};

// exclusive scan of n flags / counts into pos; *total = sum (synchronises the stream)
cudaError_t scan_total(Scratch& tmp, const uint32_t* in, uint32_t* pos, size_t n, cudaStream_t stream, uint64_t* total) {
    size_t need = 0;
    cudaError_t e = cub::DeviceScan::ExclusiveSum(nullptr, need, in, pos, (int)n, stream);
    if (e != cudaSuccess) return e;
    if ((e = tmp.need(need)) != cudaSuccess) return e;
    if ((e = cub::DeviceScan::ExclusiveSum(tmp.buf.p, need, in, pos, (int)n, stream)) != cudaSuccess) return e;
    uint32_t last[2] = {0, 0};
    if ((e = cudaMemcpyAsync(&last[0], pos + n - 1, 4, cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return e;
    if ((e = cudaMemcpyAsync(&last[1], in + n - 1, 4, cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return e;
    *total = (uint64_t)last[0] + last[1];
    return cudaSuccess;
}

}  // namespace

int generate_world(const rc_gen_params& p, cudaStream_t stream, GeneratedWorld& out, std::string& err) {
    out = GeneratedWorld();
    if (p.max_depth < 3 || p.max_depth > 8) {
        err = "rc_generate_world: max_depth must be in [3,8] (reference: 6)";
        return RC_ERR_INVALID_ARGUMENT;
    }
    const int64_t limit = 1ll << (2 * (p.max_depth - 1));
    if (p.width == 0 || p.length == 0 || (int64_t)p.width > limit || (int64_t)p.length > limit) {
        err = "rc_generate_world: width / length must be in [1, world edge]";
        return RC_ERR_INVALID_ARGUMENT;
    }
    const size_t columns = (size_t)p.width * p.length;

    GenConsts hc;
    memset(&hc, 0, sizeof(hc));
    for (int i = 0; i < 3; i++) hc.noise[i] = Simplex2(p.seeds[i]);
    hc.width = (int)p.width;
    hc.length = (int)p.length;
    hc.levels = (int)p.max_depth - 2;
    // world_gen.cpp:26-38
    hc.wGrass = material_word(out.palette, pack_rgb(0, 150, 10), 0u);
    hc.wDirt = material_word(out.palette, pack_rgb(45, 18, 0), 0u);
    hc.wStone = material_word(out.palette, pack_rgb(33, 33, 33), 0u);
    hc.wWater = material_word(out.palette, pack_rgb(0, 150, 10), 0x4u | 0x10u);  // REFRACTIVE | LIQUID
    if (p.startup_blocks) {
        // tetrahexa_tree.cpp:21-27: Block{flags, color 0, metadata 0}; level 6 = one voxel, 5 = a 4^3 cell,
        // kept relative to the voxel level.  The first one (:20, at 1000^3) lands in root cell 63, whose
        // child-array slot is 0 = "no child" (SURVEY F11): it is in no world the shader can see.
        struct Item { int x, y, z; uint32_t flags; int size; };
        const Item items[] = {{10, 100, 10, 2, 1}, {100, 10, 100, 3, 1}, {20, 10, 200, 4, 4}, {1, 10, 10, 5, 1},
                              {2, 10, 10, 6, 1},   {3, 10, 10, 7, 1},    {4, 10, 10, 8, 1}};
        for (const Item& it : items) {
            GenExtra& e = hc.extras[hc.nExtras++];
            const int wrap = (int)(limit - 1);  // putBlock keeps the low bits only (tetrahexa_tree.cpp:191)
            e.size = it.size;
            e.x0 = (it.x & wrap) & ~(it.size - 1);
            e.y0 = (it.y & wrap) & ~(it.size - 1);
            e.z0 = (it.z & wrap) & ~(it.size - 1);
            e.word = material_word(out.palette, 0ull, it.flags);
        }
    }

    const auto wall0 = std::chrono::steady_clock::now();
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
    struct EventGuard {
        cudaEvent_t* e;
        ~EventGuard() {
            for (int i = 0; i < 3; i++)
                if (e[i]) cudaEventDestroy(e[i]);
        }
    } guard{ev};
    for (int i = 0; i < 3; i++) GEN_TRY(cudaEventCreate(&ev[i]));

    // 1. height map
    DeviceBuf heights;
    GEN_TRY(heights.alloc(columns * sizeof(int)));
    GEN_TRY(cudaMemcpyToSymbolAsync(c_gen, &hc, sizeof(hc), 0, cudaMemcpyHostToDevice, stream));
    GEN_TRY(cudaEventRecord(ev[0], stream));
    heights_kernel<<<blocksFor(columns), 256, 0, stream>>>(heights.as<int>());
    GEN_TRY(cudaGetLastError());
    Scratch tmp;
    DeviceBuf minmax;
    GEN_TRY(minmax.alloc(2 * sizeof(int)));
    {
        size_t need = 0, need2 = 0;
        GEN_TRY(cub::DeviceReduce::Min(nullptr, need, heights.as<int>(), minmax.as<int>(), (int)columns, stream));
        GEN_TRY(cub::DeviceReduce::Max(nullptr, need2, heights.as<int>(), minmax.as<int>() + 1, (int)columns, stream));
        GEN_TRY(tmp.need(need > need2 ? need : need2));
        GEN_TRY(cub::DeviceReduce::Min(tmp.buf.p, need, heights.as<int>(), minmax.as<int>(), (int)columns, stream));
        GEN_TRY(cub::DeviceReduce::Max(tmp.buf.p, need2, heights.as<int>(), minmax.as<int>() + 1, (int)columns, stream));
    }
    int hmm[2] = {0, 0};
    GEN_TRY(cudaMemcpyAsync(hmm, minmax.p, sizeof(hmm), cudaMemcpyDeviceToHost, stream));
    GEN_TRY(cudaEventRecord(ev[1], stream));
    GEN_TRY(cudaStreamSynchronize(stream));
    out.info.columns = columns;
    out.info.min_height = hmm[0];
    out.info.max_height = hmm[1];
    if (hmm[0] < 0) {
        err = "rc_generate_world: a column's height is below y = 0 (the reference wraps it around the torus)";
        return RC_ERR_UNSUPPORTED;
    }
    const int top = hmm[1] > 20 ? hmm[1] : 20;  // water reaches y = 20
    if (top >= limit) {
        err = "rc_generate_world: terrain is taller than the world (the reference wraps it); raise max_depth";
        return RC_ERR_UNSUPPORTED;
    }

    // 2. candidate bricks: the terrain's bounding box + the start-up blocks outside it
    hc.nbx = ((int)p.width + 3) / 4;
    hc.nby = top / 4 + 1;
    hc.nbz = ((int)p.length + 3) / 4;
    const uint64_t nBox64 = (uint64_t)hc.nbx * hc.nby * hc.nbz;
    std::vector<int4> extraBricks;
    for (int i = 0; i < hc.nExtras; i++) {
        const int4 b = make_int4(hc.extras[i].x0 >> 2, hc.extras[i].y0 >> 2, hc.extras[i].z0 >> 2, 0);
        if (b.x < hc.nbx && b.y < hc.nby && b.z < hc.nbz) continue;
        bool seen = false;
        for (const int4& o : extraBricks) seen = seen || (o.x == b.x && o.y == b.y && o.z == b.z);
        if (!seen) extraBricks.push_back(b);
    }
    const uint64_t nCand64 = nBox64 + extraBricks.size();
    if (nCand64 > 0x7FFFFFF0ull) {
        err = "rc_generate_world: more than 2^31 candidate bricks";
        return RC_ERR_UNSUPPORTED;
    }
    const int nBox = (int)nBox64, nCand = (int)nCand64;
    out.info.candidate_bricks = nCand64;
    GEN_TRY(cudaMemcpyToSymbolAsync(c_gen, &hc, sizeof(hc), 0, cudaMemcpyHostToDevice, stream));
    DeviceBuf dExtra, candKeys, candMasks, flags, pos;
    GEN_TRY(dExtra.alloc(extraBricks.size() * sizeof(int4)));
    if (!extraBricks.empty())
        GEN_TRY(cudaMemcpyAsync(dExtra.p, extraBricks.data(), extraBricks.size() * sizeof(int4), cudaMemcpyHostToDevice, stream));
    GEN_TRY(candKeys.alloc((size_t)nCand * 8));
    GEN_TRY(candMasks.alloc((size_t)nCand * 8));
    GEN_TRY(flags.alloc((size_t)nCand * 4));
    GEN_TRY(pos.alloc((size_t)nCand * 4));
    brick_mask_kernel<<<blocksFor(nCand), 256, 0, stream>>>(heights.as<int>(), dExtra.as<int4>(), nBox, nCand,
                                                            candKeys.as<unsigned long long>(), candMasks.as<unsigned long long>(),
                                                            flags.as<uint32_t>());
    GEN_TRY(cudaGetLastError());
    uint64_t nBricks = 0;
    GEN_TRY(scan_total(tmp, flags.as<uint32_t>(), pos.as<uint32_t>(), nCand, stream, &nBricks));
    out.info.bricks = nBricks;

    const int levels = hc.levels;
    std::vector<uint64_t> count(levels + 1, 0);  // nodes per depth; depth `levels` = bricks / uniform cells
    std::vector<uint32_t> offset(levels + 2, 0);
    if (nBricks == 0) {  // an empty world is a childless root
        out.n_nodes = 1;
        out.cap_nodes = 4096;
        GEN_TRY(cudaMalloc((void**)&out.d_nodes, out.cap_nodes * sizeof(PackedNode)));
        out.cap_voxels = 65536;
        GEN_TRY(cudaMalloc((void**)&out.d_voxels, out.cap_voxels * sizeof(uint32_t)));
        GEN_TRY(cudaMemsetAsync(out.d_nodes, 0, sizeof(PackedNode), stream));
        out.level_begin = {0, 1};
    } else {
        // 3. compact + sort the non-empty bricks by hierarchical key
        std::vector<DeviceBuf> keys(levels + 1), heads(levels + 1), pidx(levels + 1);
        DeviceBuf unsortedKeys, unsortedMasks, masks;
        GEN_TRY(unsortedKeys.alloc(nBricks * 8));
        GEN_TRY(unsortedMasks.alloc(nBricks * 8));
        GEN_TRY(keys[levels].alloc(nBricks * 8));
        GEN_TRY(masks.alloc(nBricks * 8));
        compact_kernel<<<blocksFor(nCand), 256, 0, stream>>>(candKeys.as<unsigned long long>(), candMasks.as<unsigned long long>(),
                                                             flags.as<uint32_t>(), pos.as<uint32_t>(), nCand,
                                                             unsortedKeys.as<unsigned long long>(),
                                                             unsortedMasks.as<unsigned long long>());
        GEN_TRY(cudaGetLastError());
        {
            size_t need = 0;
            GEN_TRY(cub::DeviceRadixSort::SortPairs(nullptr, need, unsortedKeys.as<unsigned long long>(),
                                                    keys[levels].as<unsigned long long>(), unsortedMasks.as<unsigned long long>(),
                                                    masks.as<unsigned long long>(), (int)nBricks, 0, 6 * levels, stream));
            GEN_TRY(tmp.need(need));
            GEN_TRY(cub::DeviceRadixSort::SortPairs(tmp.buf.p, need, unsortedKeys.as<unsigned long long>(),
                                                    keys[levels].as<unsigned long long>(), unsortedMasks.as<unsigned long long>(),
                                                    masks.as<unsigned long long>(), (int)nBricks, 0, 6 * levels, stream));
        }
        // 4. tree levels, bottom-up: parents = runs of children that share a key prefix
        count[levels] = nBricks;
        for (int d = levels; d >= 1; d--) {
            const int n = (int)count[d];
            GEN_TRY(heads[d].alloc((size_t)n * 4));
            GEN_TRY(pidx[d].alloc((size_t)n * 4));
            flag_heads_kernel<<<blocksFor(n), 256, 0, stream>>>(keys[d].as<unsigned long long>(), n, 6, heads[d].as<uint32_t>());
            GEN_TRY(cudaGetLastError());
            GEN_TRY(scan_total(tmp, heads[d].as<uint32_t>(), pidx[d].as<uint32_t>(), n, stream, &count[d - 1]));
            GEN_TRY(keys[d - 1].alloc(count[d - 1] * 8));
            parent_keys_kernel<<<blocksFor(n), 256, 0, stream>>>(keys[d].as<unsigned long long>(), heads[d].as<uint32_t>(),
                                                                 pidx[d].as<uint32_t>(), n, keys[d - 1].as<unsigned long long>());
            GEN_TRY(cudaGetLastError());
        }
        if (count[0] != 1) {
            err = "rc_generate_world: internal error, tree has no single root";
            return RC_ERR_CUDA;
        }
        uint64_t total = 0;
        for (int d = 0; d <= levels; d++) {
            offset[d] = (uint32_t)total;
            total += count[d];
        }
        offset[levels + 1] = (uint32_t)total;
        // voxel runs
        DeviceBuf counts, voxBase;
        GEN_TRY(counts.alloc(nBricks * 4));
        GEN_TRY(voxBase.alloc(nBricks * 4));
        popc_kernel<<<blocksFor(nBricks), 256, 0, stream>>>(masks.as<unsigned long long>(), (int)nBricks, counts.as<uint32_t>());
        GEN_TRY(cudaGetLastError());
        // 32-bit scan: refuse worlds whose voxel count cannot be indexed (the flattener's limit too)
        if (columns * (uint64_t)(top + 1) > 0xFFFFFFF0ull) {
            err = "rc_generate_world: more than 2^32 voxels";
            return RC_ERR_UNSUPPORTED;
        }
        uint64_t nVoxels = 0;
        GEN_TRY(scan_total(tmp, counts.as<uint32_t>(), voxBase.as<uint32_t>(), nBricks, stream, &nVoxels));

        out.n_nodes = total;
        out.n_voxels = nVoxels;
        out.cap_nodes = total + total / 4 + 4096;
        out.cap_voxels = nVoxels + nVoxels / 4 + 65536;
        GEN_TRY(cudaMalloc((void**)&out.d_nodes, out.cap_nodes * sizeof(PackedNode)));
        GEN_TRY(cudaMalloc((void**)&out.d_voxels, out.cap_voxels * sizeof(uint32_t)));
        GEN_TRY(cudaMemsetAsync(out.d_nodes, 0, offset[levels] * sizeof(PackedNode), stream));
        for (int d = levels; d >= 1; d--) {
            const int n = (int)count[d];
            link_level_kernel<<<blocksFor(n), 256, 0, stream>>>(keys[d].as<unsigned long long>(), heads[d].as<uint32_t>(),
                                                                pidx[d].as<uint32_t>(), n, offset[d], offset[d - 1], out.d_nodes);
            GEN_TRY(cudaGetLastError());
        }
        brick_nodes_kernel<<<blocksFor(nBricks), 256, 0, stream>>>(heights.as<int>(), keys[levels].as<unsigned long long>(),
                                                                   masks.as<unsigned long long>(), voxBase.as<uint32_t>(),
                                                                   (int)nBricks, offset[levels], out.d_nodes, out.d_voxels);
        GEN_TRY(cudaGetLastError());
        out.level_begin.assign(offset.begin(), offset.end());
        GEN_TRY(cudaEventRecord(ev[2], stream));
        GEN_TRY(cudaStreamSynchronize(stream));  // the scratch buffers above die with this scope
    }
    if (nBricks == 0) {
        GEN_TRY(cudaEventRecord(ev[2], stream));
        GEN_TRY(cudaStreamSynchronize(stream));
    }
    out.leaves = out.n_voxels;
    out.info.nodes = out.n_nodes;
    out.info.voxels = out.n_voxels;
    GEN_TRY(cudaEventElapsedTime(&out.info.heights_ms, ev[0], ev[1]));
    GEN_TRY(cudaEventElapsedTime(&out.info.build_ms, ev[1], ev[2]));
    out.info.total_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - wall0).count();
    return RC_OK;
}

}  // namespace rc
