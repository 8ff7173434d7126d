// flatten.cpp -- host half of rc_upload_world: walk the host's two block pools
// (the byte layout updateSsboData() uploads, src/voxel_data/voxel_allocator.hpp:38-78)
// from the root and emit the packed breadth-first form described in packed_world.h.
#include "packed_world.h"

#include <string.h>

#include <algorithm>
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

int validate_desc(const rc_world_desc& desc, std::string& err) {
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
    return RC_OK;
}

// Everything below packed node n (a copy of its 16 bytes) becomes garbage.
void kill_children(const PackedWorld& pw, FlattenTracking& tr, const PackedNode& top) {
    std::vector<PackedNode> stack{top};
    while (!stack.empty()) {
        const PackedNode n = stack.back();
        stack.pop_back();
        const uint32_t kind = n.kind & PK_KIND_MASK;
        const uint32_t cnt = (uint32_t)__builtin_popcountll(n.mask);
        if (kind == PK_BRICK) {
            tr.garbage_voxels += cnt;
        } else if (kind == PK_INNER) {
            for (uint32_t j = 0; j < cnt; j++) {
                const uint32_t c = n.base + j;
                if (c >= pw.nodes.size() || c >= tr.packed_dead.size() || tr.packed_dead[c]) continue;
                tr.packed_dead[c] = 1;
                tr.garbage_nodes++;
                stack.push_back(pw.nodes[c]);
            }
        }
    }
}

// Emits the packed form of ref subtrees, breadth-first, APPENDING children runs to out.nodes /
// out.voxels and writing each subtree root into the packed slot it already owns.
//
// Update mode (a subtree root that HAS an old packed node): a node's children run must stay
// contiguous, so when a node changes its run is re-emitted -- but only the run.  A child whose ref
// node is unchanged is MOVED: its 16 bytes are copied into the new run and keep pointing at its old
// descendants.  Only children that are new, or that depend on something that changed (`dirty`), are
// emitted again.  A click therefore re-emits one run per tree level it touched.
struct Emitter {
    const Pools& pools;
    PackedWorld& out;
    Palette& palette;
    FlattenTracking* tr;
    std::string& err;
    uint64_t nodeCap;
    const std::vector<uint32_t>* dirty = nullptr;  // sorted packed indices whose content is stale (update mode)

    struct Pending {
        uint32_t packed;
        uint32_t ref;
        uint32_t depth;
        bool hasOld;
        PackedNode old;  // what the slot (or the slot this node moved from) held before
    };
    std::vector<Pending> queue;

    void own_node(uint32_t ref, uint32_t packed) {
        if (tr && ref < tr->node_owner.size()) tr->node_owner[ref] = packed;
    }
    void grow_maps() {
        if (tr && tr->packed_ref.size() < out.nodes.size()) {
            tr->packed_ref.resize(out.nodes.size(), NO_OWNER);
            tr->packed_depth.resize(out.nodes.size(), 0);
            tr->packed_dead.resize(out.nodes.size(), 0);
        }
    }
    void describe(uint32_t packed, uint32_t ref, uint32_t depth) {
        if (!tr) return;
        grow_maps();
        tr->packed_ref[packed] = ref;
        tr->packed_depth[packed] = (uint8_t)depth;
        tr->packed_dead[packed] = 0;
    }
    bool is_dirty(uint32_t packed) const { return dirty && std::binary_search(dirty->begin(), dirty->end(), packed); }

    // The packed node that currently represents ref child c of a node whose old children run is
    // [runBase, runBase + runCount), or NO_OWNER.
    uint32_t reusable(uint32_t c, uint32_t depth, uint32_t runBase, uint32_t runCount) const {
        if (!tr || c >= tr->node_owner.size()) return NO_OWNER;
        const uint32_t pc = tr->node_owner[c];
        if (pc == NO_OWNER || pc < runBase || pc >= runBase + runCount || pc >= tr->packed_ref.size()) return NO_OWNER;
        if (tr->packed_ref[pc] != c || tr->packed_dead[pc] || tr->packed_depth[pc] != depth) return NO_OWNER;
        return pc;
    }

    // ref node c (unchanged) now lives in packed slot s: re-point the maps that named its old slot
    void moved(uint32_t c, uint32_t s, uint32_t depth) {
        describe(s, c, depth);
        own_node(c, s);
        const RefNode n = pools.node(c);
        if (n.flags & 1u) return;
        if (n.word3 >= pools.n_arrays) return;
        if (n.word3 < tr->array_owner.size()) tr->array_owner[n.word3] = s;
        if (depth == out.max_depth - 2) {  // a brick: its voxel leaves are owned by the brick's slot
            const uint8_t* arr = pools.array(n.word3);
            for (int i = 0; i < 64; i++) {
                if (!((n.word0 >> i) & 1)) continue;
                const uint32_t v = pools.child(arr, i);
                if (v != 0 && v < pools.n_nodes) own_node(v, s);
            }
        }
    }

    // level_begin (full flatten only): first packed index of each depth
    int run(uint32_t packedRoot, uint32_t refRoot, uint32_t depthRoot, bool rootHasOld, std::vector<uint32_t>* levelBegin) {
        const uint32_t brickDepth = out.max_depth - 2;  // children of this depth are single voxels
        queue.clear();
        queue.push_back(Pending{packedRoot, refRoot, depthRoot, rootHasOld, rootHasOld ? out.nodes[packedRoot] : PackedNode{0, 0, 0}});
        uint32_t currentDepth = depthRoot;
        if (levelBegin) levelBegin->push_back(packedRoot);
        for (size_t qi = 0; qi < queue.size(); qi++) {
            const Pending cur = queue[qi];
            if (cur.depth != currentDepth) {
                currentDepth = cur.depth;
                if (levelBegin) levelBegin->push_back(cur.packed);
            }
            const RefNode n = pools.node(cur.ref);
            out.ref_nodes++;
            own_node(cur.ref, cur.packed);
            describe(cur.packed, cur.ref, cur.depth);
            const bool oldInner = cur.hasOld && (cur.old.kind & PK_KIND_MASK) == PK_INNER;
            if (n.flags & 1u) {  // uniform leaf above voxel level (e.g. putBlock(..., level 5) => 4^3 block)
                uint32_t w = palette.word(n);
                if (!w) {
                    err = "rc_upload_world: more than 2^24 distinct materials";
                    return RC_ERR_UNSUPPORTED;
                }
                if (cur.hasOld) kill_children(out, *tr, cur.old);
                out.nodes[cur.packed] = PackedNode{0, w, PK_LEAF | ((n.flags & 0x10u) ? PK_HAS_LIQUID : 0u)};
                continue;
            }
            if (n.word3 >= pools.n_arrays) {
                err = "rc_upload_world: branch.children outside the array pool";
                return RC_ERR_MALFORMED_WORLD;
            }
            if (tr && n.word3 < tr->array_owner.size()) tr->array_owner[n.word3] = cur.packed;
            const uint8_t* arr = pools.array(n.word3);
            uint64_t mask = 0;
            if (cur.depth == brickDepth) {
                if (cur.hasOld) kill_children(out, *tr, cur.old);  // the old voxel run (or an old subtree)
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
                    own_node(c, cur.packed);
                    out.voxels.push_back(w);
                    if (leaf.flags & 0x10u) liquid = PK_HAS_LIQUID;
                    mask |= 1ull << i;
                }
                if (out.voxels.size() > 0xFFFFFFF0ull) {
                    err = "rc_upload_world: more than 2^32 voxels";
                    return RC_ERR_UNSUPPORTED;
                }
                out.nodes[cur.packed] = PackedNode{mask, (uint32_t)voxBase, PK_BRICK | liquid};
            } else {
                if (cur.hasOld && !oldInner) kill_children(out, *tr, cur.old);  // was a brick / leaf
                const uint32_t runBase = oldInner ? cur.old.base : 0u;
                const uint32_t runCount = oldInner ? (uint32_t)__builtin_popcountll(cur.old.mask) : 0u;
                const uint64_t childBase = out.nodes.size();
                for (int i = 0; i < 64; i++) {
                    if (!((n.word0 >> i) & 1)) continue;
                    uint32_t c = pools.child(arr, i);
                    if (c == 0 || c >= pools.n_nodes) {
                        out.null_children++;  // root cell 63 of the default world lands here
                        continue;
                    }
                    mask |= 1ull << i;
                    const uint32_t s = (uint32_t)out.nodes.size();
                    const uint32_t pc = oldInner ? reusable(c, cur.depth + 1, runBase, runCount) : NO_OWNER;
                    if (pc == NO_OWNER) {
                        out.nodes.push_back(PackedNode{0, 0, PK_INNER});
                        queue.push_back(Pending{s, c, cur.depth + 1, false, PackedNode{0, 0, 0}});
                        continue;
                    }
                    const PackedNode oldChild = out.nodes[pc];
                    out.nodes.push_back(oldChild);
                    tr->packed_dead[pc] = 1;  // the slot only: its descendants live on under the copy
                    tr->garbage_nodes++;
                    if (is_dirty(pc)) {
                        queue.push_back(Pending{s, c, cur.depth + 1, true, oldChild});
                    } else {
                        grow_maps();
                        moved(c, s, cur.depth + 1);
                    }
                }
                if (out.nodes.size() > nodeCap || out.nodes.size() > 0x7FFFFFF0ull) {
                    err = "rc_upload_world: tree has more reachable nodes than the pool holds (cycle?)";
                    return RC_ERR_MALFORMED_WORLD;
                }
                // old children that were neither moved nor taken over: gone with everything below them
                for (uint32_t j = 0; j < runCount; j++) {
                    const uint32_t oc = runBase + j;
                    if (oc >= out.nodes.size() || oc >= tr->packed_dead.size() || tr->packed_dead[oc]) continue;
                    tr->packed_dead[oc] = 1;
                    tr->garbage_nodes++;
                    kill_children(out, *tr, out.nodes[oc]);
                }
                out.nodes[cur.packed] = PackedNode{mask, (uint32_t)childBase, PK_INNER};
            }
        }
        return RC_OK;
    }
};

void rebuild_palette_index(Palette& palette) {
    for (uint32_t i = 0; i < palette.entries.size(); i++) {
        const PaletteEntry& e = palette.entries[i];
        palette.index.emplace(PaletteKey{e.color, (uint64_t)e.flags | ((uint64_t)e.metadata_bits << 32)}, i);
    }
}

// copies the used part of a pool into one flat mirror
void mirror_pool(const void* const* blocks, uint32_t blockBytes, uint64_t used, std::vector<uint8_t>& dst) {
    dst.assign(used, 0);
    for (uint64_t off = 0; off < used; off += blockBytes) {
        const uint64_t n = used - off < blockBytes ? used - off : blockBytes;
        memcpy(dst.data() + off, blocks[off / blockBytes], n);
    }
}

}  // namespace

int flatten_world(const rc_world_desc& desc, PackedWorld& out, std::string& err, FlattenTracking* track) {
    out = PackedWorld();
    if (track) *track = FlattenTracking();
    int rcode = validate_desc(desc, err);
    if (rcode != RC_OK) return rcode;

    Pools pools(desc);
    if (desc.root_index >= pools.n_nodes) {
        err = "rc_upload_world: root_index outside the node pool";
        return RC_ERR_MALFORMED_WORLD;
    }
    out.max_depth = desc.max_depth;
    Palette palette(out.palette);

    RefNode rootNode = pools.node(desc.root_index);
    out.nodes.push_back(PackedNode{0, 0, PK_INNER});
    if (rootNode.flags & 1u) {
        // A root that is itself a leaf (whole torus one material): express it as an
        // inner root with 64 uniform cells so the kernel's cached node is never a leaf.
        // (Not tracked: an update of such a world is a full upload.)
        uint32_t w = palette.word(rootNode);
        out.nodes[0] = PackedNode{~0ull, 1, PK_INNER};
        for (int i = 0; i < 64; i++)
            out.nodes.push_back(PackedNode{0, w, PK_LEAF | ((rootNode.flags & 0x10u) ? PK_HAS_LIQUID : 0u)});
        out.level_begin = {0, 1, 65};
        out.ref_nodes = 1;
        return RC_OK;
    }
    if (track) {
        track->node_owner.assign(pools.n_nodes, NO_OWNER);
        track->array_owner.assign(pools.n_arrays, NO_OWNER);
        track->root_index = desc.root_index;
    }
    // generous bound against cyclic / adversarial pools: a tree cannot have more
    // reachable nodes than the pool holds
    Emitter em{pools, out, palette, track, err, pools.n_nodes + 65, nullptr, {}};
    rcode = em.run(0, desc.root_index, 0, false, &out.level_begin);
    if (rcode != RC_OK) return rcode;
    out.level_begin.push_back((uint32_t)out.nodes.size());
    if (track) {
        track->packed_ref.resize(out.nodes.size(), NO_OWNER);
        track->packed_depth.resize(out.nodes.size(), 0);
        track->packed_dead.resize(out.nodes.size(), 0);
        mirror_pool(desc.node_blocks, desc.block_bytes, desc.node_bytes_used, track->node_mirror);
        mirror_pool(desc.array_blocks, desc.block_bytes, desc.array_bytes_used, track->array_mirror);
        track->enabled = true;
    }
    return RC_OK;
}

namespace {

// 16-byte (nodes) or 256-byte (arrays) granules of one dirty block that differ from the mirror;
// the mirror is brought up to date on the way.
void diff_block(const uint8_t* block, uint64_t blockOff, uint64_t bytes, uint32_t granule, std::vector<uint8_t>& mirror,
                std::vector<uint32_t>& changed) {
    uint8_t* m = mirror.data() + blockOff;
    if (memcmp(block, m, bytes) == 0) return;
    for (uint64_t off = 0; off < bytes; off += granule) {
        const uint64_t n = bytes - off < granule ? bytes - off : granule;
        if (memcmp(block + off, m + off, n) != 0) {
            changed.push_back((uint32_t)((blockOff + off) / granule));
            memcpy(m + off, block + off, n);
        }
    }
}

}  // namespace

int update_world(const rc_world_desc& desc, const uint32_t* dirtyNodeBlocks, uint32_t nDirtyNodeBlocks,
                 const uint32_t* dirtyArrayBlocks, uint32_t nDirtyArrayBlocks, PackedWorld& pw, FlattenTracking& tr,
                 WorldPatch& patch, UpdateStats& stats, std::string& err) {
    patch = WorldPatch();
    stats = UpdateStats();
    int rcode = validate_desc(desc, err);
    if (rcode != RC_OK) return rcode;
    // RC_ALL_BLOCKS: every block is a candidate
    std::vector<uint32_t> allNodes, allArrays;
    if (nDirtyNodeBlocks == RC_ALL_BLOCKS) {
        for (uint32_t i = 0; i < desc.n_node_blocks; i++) allNodes.push_back(i);
        dirtyNodeBlocks = allNodes.data();
        nDirtyNodeBlocks = (uint32_t)allNodes.size();
    }
    if (nDirtyArrayBlocks == RC_ALL_BLOCKS) {
        for (uint32_t i = 0; i < desc.n_array_blocks; i++) allArrays.push_back(i);
        dirtyArrayBlocks = allArrays.data();
        nDirtyArrayBlocks = (uint32_t)allArrays.size();
    }
    if ((nDirtyNodeBlocks && !dirtyNodeBlocks) || (nDirtyArrayBlocks && !dirtyArrayBlocks)) {
        err = "rc_update_world: null dirty-block list";
        return RC_ERR_INVALID_ARGUMENT;
    }
    if (!tr.enabled || desc.max_depth != pw.max_depth || desc.root_index != tr.root_index) return 1;
    Pools pools(desc);
    if (desc.root_index >= pools.n_nodes) {
        err = "rc_upload_world: root_index outside the node pool";
        return RC_ERR_MALFORMED_WORLD;
    }

    // the pools only grow (the reference never frees a block: voxel_allocator.hpp:101-107)
    if (desc.node_bytes_used < tr.node_mirror.size() || desc.array_bytes_used < tr.array_mirror.size()) return 1;
    tr.node_mirror.resize(desc.node_bytes_used, 0);
    tr.array_mirror.resize(desc.array_bytes_used, 0);
    tr.node_owner.resize(pools.n_nodes, NO_OWNER);
    tr.array_owner.resize(pools.n_arrays, NO_OWNER);

    std::vector<uint32_t> changedNodes, changedArrays;
    for (uint32_t k = 0; k < nDirtyNodeBlocks; k++) {
        const uint32_t b = dirtyNodeBlocks[k];
        if (b >= desc.n_node_blocks) {
            err = "rc_update_world: dirty node block index outside the block list";
            return RC_ERR_INVALID_ARGUMENT;
        }
        const uint64_t off = (uint64_t)b * desc.block_bytes;
        if (off >= desc.node_bytes_used) continue;
        const uint64_t bytes = desc.node_bytes_used - off < desc.block_bytes ? desc.node_bytes_used - off : desc.block_bytes;
        diff_block((const uint8_t*)desc.node_blocks[b], off, bytes, 16, tr.node_mirror, changedNodes);
    }
    for (uint32_t k = 0; k < nDirtyArrayBlocks; k++) {
        const uint32_t b = dirtyArrayBlocks[k];
        if (b >= desc.n_array_blocks) {
            err = "rc_update_world: dirty array block index outside the block list";
            return RC_ERR_INVALID_ARGUMENT;
        }
        const uint64_t off = (uint64_t)b * desc.block_bytes;
        if (off >= desc.array_bytes_used) continue;
        const uint64_t bytes = desc.array_bytes_used - off < desc.block_bytes ? desc.array_bytes_used - off : desc.block_bytes;
        diff_block((const uint8_t*)desc.array_blocks[b], off, bytes, 256, tr.array_mirror, changedArrays);
    }
    stats.changed_ref_nodes = (uint32_t)changedNodes.size();
    stats.changed_ref_arrays = (uint32_t)changedArrays.size();

    // packed nodes that depend on something that changed
    std::vector<uint32_t> dirty;
    for (uint32_t r : changedNodes) {
        if (r == desc.root_index && (pools.node(r).flags & 1u)) return 1;  // root turned into a leaf
        if (tr.node_owner[r] != NO_OWNER) dirty.push_back(tr.node_owner[r]);
    }
    for (uint32_t a : changedArrays)
        if (tr.array_owner[a] != NO_OWNER) dirty.push_back(tr.array_owner[a]);
    std::sort(dirty.begin(), dirty.end(), [&](uint32_t x, uint32_t y) {
        return tr.packed_depth[x] != tr.packed_depth[y] ? tr.packed_depth[x] < tr.packed_depth[y] : x < y;
    });
    dirty.erase(std::unique(dirty.begin(), dirty.end()), dirty.end());

    patch.nodes_before = pw.nodes.size();
    patch.voxels_before = pw.voxels.size();
    patch.palette_before = pw.palette.size();
    Palette palette(pw.palette);
    rebuild_palette_index(palette);
    std::vector<uint32_t> dirtyByIndex = dirty;  // Emitter::is_dirty searches by packed index
    std::sort(dirtyByIndex.begin(), dirtyByIndex.end());
    Emitter em{pools, pw, palette, &tr, err, pw.nodes.size() + pools.n_nodes + 65, &dirtyByIndex, {}};
    const uint64_t refBefore = pw.ref_nodes;
    for (uint32_t p : dirty) {
        if (tr.packed_dead[p]) continue;  // already re-emitted (or dropped) together with its parent's run
        rcode = em.run(p, tr.packed_ref[p], tr.packed_depth[p], true, nullptr);
        if (rcode != RC_OK) return rcode;
        patch.node_index.push_back(p);
        stats.reflattened++;
    }
    pw.ref_nodes = refBefore;  // (an estimate after edits; exact again after the next full upload)
    tr.packed_ref.resize(pw.nodes.size(), NO_OWNER);
    tr.packed_depth.resize(pw.nodes.size(), 0);
    tr.packed_dead.resize(pw.nodes.size(), 0);
    // compaction is the next full upload: ask for it when more than half of the arrays is garbage
    if (tr.garbage_nodes * 2 > pw.nodes.size() + 1024 || tr.garbage_voxels * 2 > pw.voxels.size() + 65536) return 1;
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
