// flatten.cpp -- host half of rc_upload_world: walk the host's two block pools
// (the byte layout updateSsboData() uploads, src/voxel_data/voxel_allocator.hpp:38-78)
// from the root and emit the packed breadth-first form described in packed_world.h.
#include "packed_world.h"

#include <string.h>

#include <unordered_map>

namespace rc {
namespace {

// Leaf / Branch of src/voxel_data/types.hpp:29-47, read field by field (the
// reference's structs are packed; the pools are only 16-byte aligned by offset).
struct RefNode {
    uint64_t word0;  // branch.bitmap | leaf.packedColor
    uint32_t flags;  // bit 0: leaf
    uint32_t word3;  // branch.children (array index) | leaf.metadata bits
};

struct Pools {
    const rc_world_desc& d;
    uint64_t n_nodes;
    uint64_t n_arrays;

    explicit Pools(const rc_world_desc& desc) : d(desc) {
        n_nodes = desc.node_bytes_used / 16;
        n_arrays = desc.array_bytes_used / 256;
    }
    RefNode node(uint64_t index) const {
        uint64_t off = index * 16;
        const uint8_t* p = (const uint8_t*)d.node_blocks[off / d.block_bytes] + off % d.block_bytes;
        RefNode n;
        memcpy(&n.word0, p, 8);
        memcpy(&n.flags, p + 8, 4);
        memcpy(&n.word3, p + 12, 4);
        return n;
    }
    const uint8_t* array(uint64_t index) const {
        uint64_t off = index * 256;
        return (const uint8_t*)d.array_blocks[off / d.block_bytes] + off % d.block_bytes;
    }
    uint32_t child(const uint8_t* arr, int i) const {
        uint32_t c;
        memcpy(&c, arr + 4 * i, 4);
        return c;
    }
};

struct PaletteKey {
    uint64_t color;
    uint64_t rest;  // flags | metadata_bits << 32
    bool operator==(const PaletteKey& o) const { return color == o.color && rest == o.rest; }
};
struct PaletteKeyHash {
    size_t operator()(const PaletteKey& k) const {
        uint64_t h = k.color * 0x9E3779B97F4A7C15ull ^ (k.rest + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full;
        return (size_t)(h ^ (h >> 29));
    }
};

struct Palette {
    std::unordered_map<PaletteKey, uint32_t, PaletteKeyHash> index;
    std::vector<PaletteEntry>& entries;
    // Leaves arrive in long runs of the same material: remember the last one.
    PaletteKey lastKey{~0ull, ~0ull};
    uint32_t lastWord = 0;

    explicit Palette(std::vector<PaletteEntry>& e) : entries(e) {}

    // 0 on palette overflow (> 2^24 distinct materials)
    uint32_t word(const RefNode& leaf) {
        PaletteKey k{leaf.word0, (uint64_t)leaf.flags | ((uint64_t)leaf.word3 << 32)};
        if (k == lastKey) return lastWord;
        uint32_t idx;
        auto it = index.find(k);
        if (it != index.end()) {
            idx = it->second;
        } else {
            if (entries.size() > MATERIAL_INDEX_MASK) return 0;
            idx = (uint32_t)entries.size();
            entries.push_back(PaletteEntry{leaf.word0, leaf.flags, leaf.word3});
            index.emplace(k, idx);
        }
        lastKey = k;
        lastWord = idx | ((leaf.flags & 0xFFu) << MATERIAL_FLAG_SHIFT);
        return lastWord;
    }
};

}  // namespace

int flatten_world(const rc_world_desc& desc, PackedWorld& out, std::string& err) {
    out = PackedWorld();
    if (!desc.node_blocks || !desc.array_blocks || desc.block_bytes == 0 || desc.block_bytes % 256 != 0) {
        err = "rc_upload_world: null block lists or bad block_bytes";
        return RC_ERR_INVALID_ARGUMENT;
    }
    if (desc.max_depth < 2 || desc.max_depth > 8) {
        err = "rc_upload_world: max_depth must be in [2,8] (reference: 6)";
        return RC_ERR_INVALID_ARGUMENT;
    }
    if (desc.node_bytes_used > (uint64_t)desc.n_node_blocks * desc.block_bytes ||
        desc.array_bytes_used > (uint64_t)desc.n_array_blocks * desc.block_bytes) {
        err = "rc_upload_world: bytes_used exceeds the blocks given";
        return RC_ERR_INVALID_ARGUMENT;
    }
    for (uint32_t i = 0; i < desc.n_node_blocks; i++)
        if (!desc.node_blocks[i] && (uint64_t)i * desc.block_bytes < desc.node_bytes_used) {
            err = "rc_upload_world: null node block inside the used range";
            return RC_ERR_INVALID_ARGUMENT;
        }
    for (uint32_t i = 0; i < desc.n_array_blocks; i++)
        if (!desc.array_blocks[i] && (uint64_t)i * desc.block_bytes < desc.array_bytes_used) {
            err = "rc_upload_world: null array block inside the used range";
            return RC_ERR_INVALID_ARGUMENT;
        }

    Pools pools(desc);
    if (desc.root_index >= pools.n_nodes) {
        err = "rc_upload_world: root_index outside the node pool";
        return RC_ERR_MALFORMED_WORLD;
    }
    const uint32_t D = desc.max_depth;
    const uint32_t brickDepth = D - 2;  // children of this depth are single voxels
    out.max_depth = D;

    struct Pending {
        uint32_t ref;
        uint32_t depth;
    };
    std::vector<Pending> queue;
    Palette palette(out.palette);
    // generous bound against cyclic / adversarial pools: a tree cannot have more
    // reachable nodes than the pool holds
    const uint64_t nodeCap = pools.n_nodes + 65;

    RefNode rootNode = pools.node(desc.root_index);
    out.nodes.push_back(PackedNode{0, 0, PK_INNER});
    queue.push_back(Pending{desc.root_index, 0});
    if (rootNode.flags & 1u) {
        // A root that is itself a leaf (whole torus one material): express it as an
        // inner root with 64 uniform cells so the kernel's cached node is never a leaf.
        uint32_t w = palette.word(rootNode);
        out.nodes[0] = PackedNode{~0ull, 1, PK_INNER};
        for (int i = 0; i < 64; i++) out.nodes.push_back(PackedNode{0, w, PK_LEAF | ((rootNode.flags & 0x10u) ? PK_HAS_LIQUID : 0u)});
        out.level_begin = {0, 1, 65};
        out.ref_nodes = 1;
        return RC_OK;
    }

    uint32_t currentDepth = 0;
    out.level_begin.push_back(0);
    for (size_t qi = 0; qi < queue.size(); qi++) {
        const Pending cur = queue[qi];
        if (cur.depth != currentDepth) {
            currentDepth = cur.depth;
            out.level_begin.push_back((uint32_t)qi);
        }
        const RefNode n = pools.node(cur.ref);
        out.ref_nodes++;
        if (n.flags & 1u) {  // uniform leaf above voxel level (e.g. putBlock(..., level 5) => 4^3 block)
            uint32_t w = palette.word(n);
            if (!w) {
                err = "rc_upload_world: more than 2^24 distinct materials";
                return RC_ERR_UNSUPPORTED;
            }
            out.nodes[qi] = PackedNode{0, w, PK_LEAF | ((n.flags & 0x10u) ? PK_HAS_LIQUID : 0u)};
            continue;
        }
        if (n.word3 >= pools.n_arrays) {
            err = "rc_upload_world: branch.children outside the array pool";
            return RC_ERR_MALFORMED_WORLD;
        }
        const uint8_t* arr = pools.array(n.word3);
        uint64_t mask = 0;
        if (cur.depth == brickDepth) {
            const uint64_t voxBase = out.voxels.size();
            uint32_t liquid = 0;
            for (int i = 0; i < 64; i++) {
                if (!((n.word0 >> i) & 1)) continue;
                uint32_t c = pools.child(arr, i);
                if (c == 0 || c >= pools.n_nodes) {
                    out.null_children++;
                    continue;
                }
                RefNode leaf = pools.node(c);
                if (!(leaf.flags & 1u)) {
                    out.null_children++;  // a branch at voxel level is malformed: empty (oracle does the same)
                    continue;
                }
                uint32_t w = palette.word(leaf);
                if (!w) {
                    err = "rc_upload_world: more than 2^24 distinct materials";
                    return RC_ERR_UNSUPPORTED;
                }
                out.ref_nodes++;
                out.voxels.push_back(w);
                if (leaf.flags & 0x10u) liquid = PK_HAS_LIQUID;
                mask |= 1ull << i;
            }
            if (out.voxels.size() > 0xFFFFFFF0ull) {
                err = "rc_upload_world: more than 2^32 voxels";
                return RC_ERR_UNSUPPORTED;
            }
            out.nodes[qi] = PackedNode{mask, (uint32_t)voxBase, PK_BRICK | liquid};
        } else {
            const uint64_t childBase = out.nodes.size();
            for (int i = 0; i < 64; i++) {
                if (!((n.word0 >> i) & 1)) continue;
                uint32_t c = pools.child(arr, i);
                if (c == 0 || c >= pools.n_nodes) {
                    out.null_children++;  // root cell 63 of the default world lands here
                    continue;
                }
                mask |= 1ull << i;
                out.nodes.push_back(PackedNode{0, 0, PK_INNER});
                queue.push_back(Pending{c, cur.depth + 1});
            }
            if (out.nodes.size() > nodeCap || out.nodes.size() > 0x7FFFFFF0ull) {
                err = "rc_upload_world: tree has more reachable nodes than the pool holds (cycle?)";
                return RC_ERR_MALFORMED_WORLD;
            }
            out.nodes[qi] = PackedNode{mask, (uint32_t)childBase, PK_INNER};
        }
    }
    out.level_begin.push_back((uint32_t)queue.size());
    return RC_OK;
}

uint32_t packed_lookup(const PackedWorld& w, int32_t x, int32_t y, int32_t z, int* cell_shift) {
    const int D = (int)w.max_depth;
    const uint32_t wrap = (1u << (2 * (D - 1))) - 1;
    // max_depth 6: bits 24-25 of the raw coordinate leak into bits 8-9, as in the reference's bit
    // spread (low_res.frag:454-459; see kernels.cu wrap_coord)
    const uint32_t leak = D == 6 ? 0x300u : 0u;
    const uint32_t ux = ((uint32_t)x & wrap) | (((uint32_t)x >> 16) & leak);
    const uint32_t uy = ((uint32_t)y & wrap) | (((uint32_t)y >> 16) & leak);
    const uint32_t uz = ((uint32_t)z & wrap) | (((uint32_t)z >> 16) & leak);
    int shift = 2 * (D - 2);
    const PackedNode* n = &w.nodes[0];
    for (;;) {
        int idx = (int)((((uz >> shift) & 3) << 4) | (((uy >> shift) & 3) << 2) | ((ux >> shift) & 3));
        if (!((n->mask >> idx) & 1)) {
            if (cell_shift) *cell_shift = shift;
            return 0;
        }
        uint32_t rank = (uint32_t)__builtin_popcountll(n->mask & ((1ull << idx) - 1));
        if ((n->kind & PK_KIND_MASK) == PK_BRICK) {
            if (cell_shift) *cell_shift = 0;
            return w.voxels[n->base + rank];
        }
        const PackedNode* c = &w.nodes[n->base + rank];
        if ((c->kind & PK_KIND_MASK) == PK_LEAF) {
            if (cell_shift) *cell_shift = shift;
            return c->base;
        }
        n = c;
        shift -= 2;
    }
}

}  // namespace rc
