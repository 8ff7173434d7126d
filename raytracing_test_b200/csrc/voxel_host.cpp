// voxel_host.cpp -- libvoxelhost.so (include/voxel_host.h): a from-scratch host-side
// block store, terrain generator and picking ray that speak the reference's pool
// wire format (16-byte nodes + 256-byte child-index arrays in 4 MiB blocks,
// src/voxel_data/types.hpp:29-47, src/voxel_data/voxel_allocator.hpp:12-18).
//
// Built with -ffp-contract=off: the terrain heights are functions of double-precision
// noise values and must round exactly like the reference's.
#include "../../include/voxel_host.h"
#include "simplex2.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

namespace {

constexpr uint32_t kBlockBytes = 1u << 22;  // BLOCK_SIZE (voxel_allocator.hpp:14)
constexpr uint32_t kNodeBytes = 16;
constexpr uint32_t kArrayBytes = 256;
constexpr uint32_t kLeafBit = 1u;

// A bump allocator over fixed-size malloc blocks. An allocation never straddles a block.
struct BlockPool {
    std::vector<uint8_t*> blocks;
    uint64_t used = 0;

    ~BlockPool() {
        for (uint8_t* b : blocks) free(b);
    }
    uint8_t* at(uint64_t off) const { return blocks[off / kBlockBytes] + off % kBlockBytes; }
    uint64_t take(uint32_t bytes) {
        if (used % kBlockBytes + bytes > kBlockBytes) used = (used / kBlockBytes + 1) * (uint64_t)kBlockBytes;
        while (used + bytes > (uint64_t)blocks.size() * kBlockBytes) blocks.push_back((uint8_t*)calloc(1, kBlockBytes));
        const uint64_t off = used;
        used += bytes;
        return off;
    }
};

// View of one 16-byte node in the pool.
struct NodeRef {
    uint8_t* p;
    uint64_t word0() const { uint64_t v; memcpy(&v, p, 8); return v; }
    uint32_t flags() const { uint32_t v; memcpy(&v, p + 8, 4); return v; }
    uint32_t word3() const { uint32_t v; memcpy(&v, p + 12, 4); return v; }
    void set(uint64_t w0, uint32_t fl, uint32_t w3) {
        memcpy(p, &w0, 8);
        memcpy(p + 8, &fl, 4);
        memcpy(p + 12, &w3, 4);
    }
    bool isLeaf() const { return (flags() & kLeafBit) != 0; }
};

uint32_t floatBits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
float bitsFloat(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

using rc_noise::Simplex2;

}  // namespace

struct vh_world {
    uint32_t maxDepth;
    BlockPool nodes, arrays;
    std::vector<const void*> nodePtrs, arrayPtrs;  // scratch for vh_describe

    NodeRef node(uint32_t index) const { return NodeRef{nodes.at((uint64_t)index * kNodeBytes)}; }
    uint8_t* arrayAt(uint32_t index) const { return arrays.at((uint64_t)index * kArrayBytes); }
    uint32_t child(uint32_t arrayIndex, int i) const {
        uint32_t c;
        memcpy(&c, arrayAt(arrayIndex) + 4 * i, 4);
        return c;
    }
    void setChild(uint32_t arrayIndex, int i, uint32_t c) { memcpy(arrayAt(arrayIndex) + 4 * i, &c, 4); }

    uint32_t newNode() { return (uint32_t)(nodes.take(kNodeBytes) / kNodeBytes); }
    uint32_t newNodes64() { return (uint32_t)(nodes.take(64 * kNodeBytes) / kNodeBytes); }
    uint32_t newArray() {
        const uint64_t off = arrays.take(kArrayBytes);
        memset(arrays.at(off), 0, kArrayBytes);
        return (uint32_t)(off / kArrayBytes);
    }
    uint32_t newBranch() {
        const uint32_t n = newNode();
        node(n).set(0, 0, newArray());
        return n;
    }
    // child cell of (x,y,z) under a node of the given depth: index = z<<4 | y<<2 | x
    int cellIndex(int32_t x, int32_t y, int32_t z, uint32_t depth) const {
        const int sh = 2 * (int)(maxDepth - 2 - depth);
        return (((z >> sh) & 3) << 4) | (((y >> sh) & 3) << 2) | ((x >> sh) & 3);
    }
};

extern "C" {

vh_world* vh_create(uint32_t max_depth) {
    if (max_depth < 2 || max_depth > 8) return nullptr;
    vh_world* w = new vh_world();
    w->maxDepth = max_depth;
    // node 0 is the root (stack[0] = 0, low_res.frag:267); its child array comes from the array
    // pool (the reference takes it from the node pool by mistake: SURVEY F11)
    w->newBranch();
    return w;
}

void vh_destroy(vh_world* w) { delete w; }

int vh_put_block(vh_world* w, int32_t x, int32_t y, int32_t z, const vh_block* block, int level) {
    if (!w || !block || level < 1 || level > (int)w->maxDepth) return -1;
    const uint32_t target = (uint32_t)level - 1;
    const uint32_t leafFlags = block->flags | kLeafBit;
    const uint32_t meta = floatBits(block->metadata);
    uint32_t cur = 0;
    for (uint32_t depth = 0;; depth++) {
        NodeRef n = w->node(cur);
        if (depth == target) {
            // overwrite whatever is here (a branch's subtree is simply dropped)
            n.set(block->color, leafFlags, meta);
            return 0;
        }
        const int idx = w->cellIndex(x, y, z, depth);
        if (n.isLeaf()) {
            // split a uniform cell into 64 copies of itself, then keep descending
            const uint64_t c0 = n.word0();
            const uint32_t f0 = n.flags(), m0 = n.word3();
            const uint32_t arr = w->newArray();
            const uint32_t kids = w->newNodes64();
            for (int i = 0; i < 64; i++) {
                w->node(kids + i).set(c0, f0, m0);
                w->setChild(arr, i, kids + i);
            }
            w->node(cur).set(~0ull, 0, arr);
            cur = kids + idx;
            continue;
        }
        const uint64_t bitmap = n.word0();
        const uint32_t arr = n.word3();
        if (!((bitmap >> idx) & 1)) {
            uint32_t kid;
            if (depth + 1 == target) {
                kid = w->newNode();
                w->node(kid).set(block->color, leafFlags, meta);
            } else {
                kid = w->newBranch();
            }
            // newNode/newBranch may have grown the pool: re-resolve the parent
            w->node(cur).set(bitmap | (1ull << idx), 0, arr);
            w->setChild(arr, idx, kid);
            if (depth + 1 == target) return 0;
            cur = kid;
        } else {
            cur = w->child(arr, idx);
        }
    }
}

int vh_get_block(const vh_world* w, int32_t x, int32_t y, int32_t z, vh_block* out) {
    uint32_t cur = 0;
    for (uint32_t depth = 0; depth < w->maxDepth; depth++) {
        NodeRef n = w->node(cur);
        if (n.isLeaf()) {
            if (out) {
                out->color = n.word0();
                out->flags = n.flags();
                out->metadata = bitsFloat(n.word3());
            }
            return 1;
        }
        if (depth == w->maxDepth - 1) break;
        const int idx = w->cellIndex(x, y, z, depth);
        if (!((n.word0() >> idx) & 1)) return 0;
        cur = w->child(n.word3(), idx);
        if (cur == 0) return 0;
    }
    return 0;
}

void vh_get_block_batch(const vh_world* w, const int32_t* xyz, uint64_t n, uint64_t* out) {
    // the shader's view of getBlock (low_res.frag:493-532): empty = (~0, 0), solid = (color, flags | metadata << 32)
    for (uint64_t i = 0; i < n; i++) {
        vh_block b;
        if (vh_get_block(w, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], &b)) {
            out[2 * i] = b.color;
            out[2 * i + 1] = (uint64_t)b.flags | ((uint64_t)floatBits(b.metadata) << 32);
        } else {
            out[2 * i] = ~0ull;
            out[2 * i + 1] = 0;
        }
    }
}

void vh_copy_pools(const vh_world* w, void* nodes_out, void* arrays_out) {
    for (size_t i = 0; i < w->nodes.blocks.size(); i++)
        memcpy((uint8_t*)nodes_out + i * (size_t)kBlockBytes, w->nodes.blocks[i], kBlockBytes);
    for (size_t i = 0; i < w->arrays.blocks.size(); i++)
        memcpy((uint8_t*)arrays_out + i * (size_t)kBlockBytes, w->arrays.blocks[i], kBlockBytes);
}

uint64_t vh_rgb(int r, int g, int b) {
    // convertScale (src/types.hpp:8): (u64)((float)x / 255.0 * RGB_RANGE) & RGB_MASK, in double
    auto scale = [](int c) -> uint64_t { return (uint64_t)((double)(float)c / 255.0 * 2097151) & 2097151ull; };
    return (scale(r) << 42) | (scale(g) << 21) | scale(b);
}

void vh_add_startup_blocks(vh_world* w) {
    // tetrahexa_tree.cpp:20-27: Block{flags, color 0, metadata 0}; levels are relative to the
    // reference's maxDepth 6 (5 => 4^3 cube), kept relative to the voxel level here.
    const int voxel = (int)w->maxDepth;
    struct Item { int x, y, z; uint32_t flags; int level; };
    const Item items[] = {
        {10, 100, 10, 2, voxel}, {100, 10, 100, 3, voxel}, {20, 10, 200, 4, voxel - 1}, {1, 10, 10, 5, voxel},
        {2, 10, 10, 6, voxel},   {3, 10, 10, 7, voxel},    {4, 10, 10, 8, voxel},
    };
    for (const Item& it : items) {
        vh_block b{0, it.flags, 0.0f};
        vh_put_block(w, it.x, it.y, it.z, &b, it.level);
    }
}

int vh_generate_terrain(vh_world* w, int32_t width, int32_t length, const int64_t* seeds) {
    if (!w || width <= 0 || length <= 0 || !seeds) return -1;
    const int32_t limit = 1 << (2 * (w->maxDepth - 1));
    if (width > limit || length > limit) return -1;
    const Simplex2 o1(seeds[0]), o2(seeds[1]), o3(seeds[2]);
    const int voxel = (int)w->maxDepth;
    const vh_block water{vh_rgb(0, 150, 10), VH_REFRACTIVE | VH_LIQUID, 0.0f};
    const vh_block grass{vh_rgb(0, 150, 10), 0, 0.0f};
    const vh_block dirt{vh_rgb(45, 18, 0), 0, 0.0f};
    const vh_block stone{vh_rgb(33, 33, 33), 0, 0.0f};
    for (int32_t x = 0; x < width; x++) {
        for (int32_t z = 0; z < length; z++) {
            // world_gen.cpp:22 (round() is the C library's, half away from zero)
            int y = (int)(round(o1.eval(x * 0.005, z * 0.005) * 30) + round(o2.eval(x * 0.05, z * 0.05) * 5) +
                          round(o3.eval(x * 0.1, z * 0.1) * 3) + 32);
            if (y < 20) {  // below sea level: water down to the bed, bed is dirt
                for (int i = 20; i > y; i--) vh_put_block(w, x, i, z, &water, voxel);
                vh_put_block(w, x, y, z, &dirt, voxel);
            } else {
                vh_put_block(w, x, y, z, &grass, voxel);
            }
            y--;
            for (int i = 3; y > 0 && i; i--, y--) vh_put_block(w, x, y, z, &dirt, voxel);
            for (; y > 0; y--) vh_put_block(w, x, y, z, &stone, voxel);
        }
    }
    return 0;
}

int vh_describe(const vh_world* wc, rc_world_desc* out) {
    if (!wc || !out) return -1;
    vh_world* w = const_cast<vh_world*>(wc);
    w->nodePtrs.assign(w->nodes.blocks.begin(), w->nodes.blocks.end());
    w->arrayPtrs.assign(w->arrays.blocks.begin(), w->arrays.blocks.end());
    memset(out, 0, sizeof(*out));
    out->node_blocks = w->nodePtrs.data();
    out->array_blocks = w->arrayPtrs.data();
    out->n_node_blocks = (uint32_t)w->nodePtrs.size();
    out->n_array_blocks = (uint32_t)w->arrayPtrs.size();
    out->block_bytes = kBlockBytes;
    out->root_index = 0;
    out->node_bytes_used = w->nodes.used;
    out->array_bytes_used = w->arrays.used;
    out->max_depth = w->maxDepth;
    return 0;
}

void vh_default_camera(float* pos, float* dir, float* sun) {
    // src/globals.cpp:20-23; glm::normalize(v) = v * (1/sqrt(dot(v,v)))
    auto normalize = [](float x, float y, float z, float* o) {
        const float inv = 1.0f / sqrtf(x * x + y * y + z * z);
        o[0] = x * inv;
        o[1] = y * inv;
        o[2] = z * inv;
    };
    pos[0] = 35.0f; pos[1] = 50.0f; pos[2] = 35.0f;
    normalize(1.0f, 0.0f, 1.0f, dir);
    normalize(2.0f, 1.0f, 4.0f, sun);
}

void vh_pick(const vh_world* w, const float* cam_pos, const float* cam_dir, int steps, int32_t* out) {
    // src/ray_caster.cpp:19-44,54-87: double-precision deltas of fp32 quotients, trunc() start voxel
    int step[3], rnd[3], last[3];
    double ad[3], t[3];
    for (int a = 0; a < 3; a++) {
        step[a] = cam_dir[a] < 0 ? -1 : 1;
        const double delta = (double)(1 / cam_dir[a]);
        ad[a] = fabs(delta);
        rnd[a] = (int)truncf(cam_pos[a]);
        double exact = cam_pos[a];
        if (step[a] < 0) exact -= 1;
        t[a] = ad[a] - (exact - (double)rnd[a]) * delta;
    }
    for (int a = 0; a < 3; a++) last[a] = rnd[a];
    int left = 0;
    while (steps--) {
        for (int a = 0; a < 3; a++) last[a] = rnd[a];
        const int a = (t[0] < t[1] && t[0] < t[2]) ? 0 : (t[1] < t[2] ? 1 : 2);
        rnd[a] += step[a];
        t[a] += ad[a];
        vh_block b;
        if (vh_get_block(w, rnd[0], rnd[1], rnd[2], &b) && (b.flags & VH_LIQUID) == 0) {
            left = steps;
            break;
        }
    }
    for (int a = 0; a < 3; a++) {
        out[a] = rnd[a];
        out[3 + a] = last[a];
    }
    out[6] = left;
}

double vh_noise2(int64_t seed, double x, double y) { return Simplex2(seed).eval(x, y); }

void vh_noise2_grid(int64_t seed, double scale, int32_t w, int32_t l, double* out) {
    const Simplex2 n(seed);
    for (int32_t x = 0; x < w; x++)
        for (int32_t z = 0; z < l; z++) out[(size_t)x * l + z] = n.eval(x * scale, z * scale);
}

}  // extern "C"
