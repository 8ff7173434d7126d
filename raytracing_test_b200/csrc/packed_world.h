// packed_world.h -- the flattened, GPU-resident form of the host's 64-tree.
//
// The reference keeps one 16-byte node per voxel plus a separate 256-byte array
// of child indices per branch (src/voxel_data/types.hpp:29-47,
// src/shaders/low_res.frag:33-39): two dependent loads per tree level, 16 B per
// voxel.  The packed form is built for the traversal kernel instead:
//
//   PackedNode (16 B, one 128-bit load per level)
//     mask : 64-bit child occupancy, bit i = child cell i  (i = z<<4 | y<<2 | x,
//            the reference's index order, tetrahexa_tree.cpp:127-129)
//     base : index of child 0; child i lives at base + popcount(mask & ((1<<i)-1))
//     kind : PK_INNER  children are PackedNodes            (nodes[base + rank])
//            PK_BRICK  children are voxels: 4x4x4 brick    (voxels[base + rank])
//            PK_LEAF   no children: the whole cell is one material, base = material word
//
//   voxel / material word (4 B):  bits 0..23 palette index, bits 24..31 = low 8
//     bits of the reference's leaf flags (bit 0 "leaf" is always set, so a
//     material word is never 0; 0 means empty).  The traversal decisions
//     (reflective 0x2, refractive 0x4, liquid 0x10; src/globals.hpp:68-74) need no
//     second load; colour comes from the palette only when a pixel is shaded.
//
//   palette entry (16 B) = the reference's Leaf verbatim {u64 packedColor, u32 flags, f32 metadata}.
//
// Nodes are emitted in breadth-first order: node 0 is the root, the top levels
// are one contiguous prefix (what the kernel stages in shared memory), siblings
// are adjacent (what popcount addressing needs).
//
// Equivalence contract (tests/test_flatten.py): for every voxel of the torus,
// packed lookup == the reference's getBlock, except that a child index of 0 is
// "no child" (SURVEY F11).
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/rc.h"

namespace rc {

enum : uint32_t { PK_INNER = 0, PK_BRICK = 1, PK_LEAF = 2 };
// kind = PK_* in the low byte, flags above it
constexpr uint32_t PK_KIND_MASK = 0xFFu;
// PK_BRICK: at least one voxel of the brick is LIQUID (0x10, src/globals.hpp:72); PK_LEAF: the material is.
// Shadow / AO rays pass through liquid (low_res.frag:384): where the flag is clear an occupancy bit
// alone says "opaque" and no material word has to be fetched.
constexpr uint32_t PK_HAS_LIQUID = 0x100u;

struct PackedNode {
    uint64_t mask;
    uint32_t base;
    uint32_t kind;
};
static_assert(sizeof(PackedNode) == 16, "PackedNode must be one 128-bit load");

struct PaletteEntry {
    uint64_t color;
    uint32_t flags;
    uint32_t metadata_bits;
};
static_assert(sizeof(PaletteEntry) == 16, "PaletteEntry mirrors the reference Leaf");

constexpr uint32_t MATERIAL_INDEX_MASK = 0x00FFFFFFu;
constexpr uint32_t MATERIAL_FLAG_SHIFT = 24;

struct PackedWorld {
    std::vector<PackedNode> nodes;
    std::vector<uint32_t> voxels;
    std::vector<PaletteEntry> palette;
    std::vector<uint32_t> level_begin;  // nodes of depth d are [level_begin[d], level_begin[d+1])
    uint32_t max_depth = 0;
    uint64_t ref_nodes = 0;
    uint32_t null_children = 0;
};

// What rc_update_world needs to patch a packed world in place after the host edited its tree
// (putBlock / deleteBlock mark whole 4 MiB blocks modified: src/voxel_data/voxel_allocator.hpp:50-56,
// 69-75).  Kept next to the PackedWorld when tracking is on.
struct FlattenTracking {
    bool enabled = false;
    std::vector<uint8_t> node_mirror, array_mirror;  // the pools as of the last upload / update
    std::vector<uint32_t> node_owner;    // ref node index  -> packed node whose 16 bytes / voxel run depend on it
    std::vector<uint32_t> array_owner;   // ref array index -> packed node built from it
    std::vector<uint32_t> packed_ref;    // packed node -> ref node index
    std::vector<uint8_t> packed_depth;   // packed node -> depth of that ref node
    std::vector<uint8_t> packed_dead;    // packed node abandoned by an update (garbage until the next full upload)
    uint32_t root_index = 0;
    uint64_t garbage_nodes = 0, garbage_voxels = 0;
};
constexpr uint32_t NO_OWNER = 0xFFFFFFFFu;

// Device-side consequences of one update: nodes rewritten in place + the appended tails.
struct WorldPatch {
    std::vector<uint32_t> node_index;  // packed nodes whose 16 bytes changed in place
    uint64_t nodes_before = 0, voxels_before = 0, palette_before = 0;  // arrays grew from these sizes
};

struct UpdateStats {
    uint32_t changed_ref_nodes = 0, changed_ref_arrays = 0, reflattened = 0;
};

// Host-side flattener: reference pools -> PackedWorld. Returns RC_OK or a negative
// rc_status with a message in err.  With `track`, also fills the maps an update needs.
int flatten_world(const rc_world_desc& desc, PackedWorld& out, std::string& err, FlattenTracking* track = nullptr);

// Incremental re-flatten after the host modified the listed 4 MiB blocks. Finds the 16-byte nodes /
// 256-byte child arrays that really changed (diff against the mirror), re-emits the packed subtrees
// that depend on them at the END of the packed arrays (children runs stay contiguous; the old runs
// become garbage) and rewrites the subtree roots in place.  Returns RC_OK, a negative rc_status, or
// 1 = "do a full flatten instead" (no tracking yet, depth / root changed, root edited, too much garbage).
int update_world(const rc_world_desc& desc, const uint32_t* dirty_node_blocks, uint32_t n_dirty_node_blocks,
                 const uint32_t* dirty_array_blocks, uint32_t n_dirty_array_blocks, PackedWorld& pw,
                 FlattenTracking& tr, WorldPatch& patch, UpdateStats& stats, std::string& err);

// world_blob.cpp: the flattened world as a file (header + the three arrays + FNV-1a checksum) and frames as PPM.
int save_packed(const PackedWorld& w, const char* path, std::string& err);
int load_packed(PackedWorld& w, const char* path, std::string& err);
int write_ppm(const char* path, const uint8_t* rgba8_bottom_up, uint32_t width, uint32_t height, std::string& err);

// Host-side lookup in the packed form (used by the flattener's self-check and by
// tests through rc_debug_lookup). Returns the material word, 0 = empty;
// *cell_shift = log2 of the edge of the uniform cell that contains the voxel.
uint32_t packed_lookup(const PackedWorld& w, int32_t x, int32_t y, int32_t z, int* cell_shift);

}  // namespace rc
