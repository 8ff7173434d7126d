// world_blob.cpp -- on-disk form of the flattened world (the reference has no file format at all:
// its world exists only as the two malloc'd pools) and a PPM writer for frames.
//
//   header (64 bytes, little endian)
//     char     magic[8]   "RCPWORLD"
//     uint32   version    1
//     uint32   max_depth
//     uint64   n_nodes, n_voxels, n_palette
//     uint64   ref_nodes
//     uint32   null_children
//     uint32   reserved
//     uint64   checksum   FNV-1a 64 over the three payload arrays, in file order
//   PackedNode[n_nodes] | uint32 voxels[n_voxels] | PaletteEntry[n_palette]
#include <stdio.h>
#include <string.h>

#include "packed_world.h"

namespace rc {
namespace {

struct BlobHeader {
    char magic[8];
    uint32_t version;
    uint32_t max_depth;
    uint64_t n_nodes, n_voxels, n_palette;
    uint64_t ref_nodes;
    uint32_t null_children;
    uint32_t reserved;
    uint64_t checksum;
};
static_assert(sizeof(BlobHeader) == 64, "blob header is 64 bytes");

const char MAGIC[8] = {'R', 'C', 'P', 'W', 'O', 'R', 'L', 'D'};

uint64_t fnv1a(uint64_t h, const void* data, size_t n) {
    const uint8_t* p = (const uint8_t*)data;
    for (size_t i = 0; i < n; i++) {
        h ^= p[i];
        h *= 0x100000001B3ull;
    }
    return h;
}

uint64_t checksum(const PackedWorld& w) {
    uint64_t h = 0xCBF29CE484222325ull;
    h = fnv1a(h, w.nodes.data(), w.nodes.size() * sizeof(PackedNode));
    h = fnv1a(h, w.voxels.data(), w.voxels.size() * sizeof(uint32_t));
    h = fnv1a(h, w.palette.data(), w.palette.size() * sizeof(PaletteEntry));
    return h;
}

}  // namespace

int save_packed(const PackedWorld& w, const char* path, std::string& err) {
    if (!path || w.nodes.empty()) {
        err = "rc_save_world: no path or no world";
        return RC_ERR_INVALID_ARGUMENT;
    }
    FILE* f = fopen(path, "wb");
    if (!f) {
        err = std::string("rc_save_world: cannot open ") + path;
        return RC_ERR_INVALID_ARGUMENT;
    }
    BlobHeader h;
    memset(&h, 0, sizeof(h));
    memcpy(h.magic, MAGIC, 8);
    h.version = 1;
    h.max_depth = w.max_depth;
    h.n_nodes = w.nodes.size();
    h.n_voxels = w.voxels.size();
    h.n_palette = w.palette.size();
    h.ref_nodes = w.ref_nodes;
    h.null_children = w.null_children;
    h.checksum = checksum(w);
    bool ok = fwrite(&h, sizeof(h), 1, f) == 1;
    ok = ok && fwrite(w.nodes.data(), sizeof(PackedNode), w.nodes.size(), f) == w.nodes.size();
    ok = ok && fwrite(w.voxels.data(), sizeof(uint32_t), w.voxels.size(), f) == w.voxels.size();
    ok = ok && fwrite(w.palette.data(), sizeof(PaletteEntry), w.palette.size(), f) == w.palette.size();
    ok = (fclose(f) == 0) && ok;
    if (!ok) {
        err = std::string("rc_save_world: short write to ") + path;
        return RC_ERR_INVALID_ARGUMENT;
    }
    return RC_OK;
}

int load_packed(PackedWorld& w, const char* path, std::string& err) {
    w = PackedWorld();
    FILE* f = path ? fopen(path, "rb") : nullptr;
    if (!f) {
        err = std::string("rc_load_world: cannot open ") + (path ? path : "(null)");
        return RC_ERR_INVALID_ARGUMENT;
    }
    BlobHeader h;
    auto bad = [&](const char* why) {
        fclose(f);
        w = PackedWorld();
        err = std::string("rc_load_world: ") + why;
        return RC_ERR_MALFORMED_WORLD;
    };
    if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, MAGIC, 8) != 0) return bad("not a packed-world blob");
    if (h.version != 1) return bad("unknown blob version");
    if (h.max_depth < 2 || h.max_depth > 8 || h.n_nodes == 0 || h.n_nodes > 0x7FFFFFF0ull || h.n_voxels > 0xFFFFFFF0ull ||
        h.n_palette > MATERIAL_INDEX_MASK + 1ull)
        return bad("header out of range");
    // the payload must be exactly as long as the header says
    if (fseek(f, 0, SEEK_END) != 0) return bad("seek failed");
    const long end = ftell(f);
    const uint64_t want = sizeof(h) + h.n_nodes * sizeof(PackedNode) + h.n_voxels * 4 + h.n_palette * sizeof(PaletteEntry);
    if (end < 0 || (uint64_t)end != want) return bad("file length does not match the header");
    fseek(f, sizeof(h), SEEK_SET);
    w.nodes.resize(h.n_nodes);
    w.voxels.resize(h.n_voxels);
    w.palette.resize(h.n_palette);
    bool ok = fread(w.nodes.data(), sizeof(PackedNode), w.nodes.size(), f) == w.nodes.size();
    ok = ok && fread(w.voxels.data(), sizeof(uint32_t), w.voxels.size(), f) == w.voxels.size();
    ok = ok && fread(w.palette.data(), sizeof(PaletteEntry), w.palette.size(), f) == w.palette.size();
    if (!ok) return bad("short read");
    w.max_depth = h.max_depth;
    w.ref_nodes = h.ref_nodes;
    w.null_children = h.null_children;
    if (checksum(w) != h.checksum) return bad("checksum mismatch");
    // Structural check, by walking from the root the way the kernels do: fetch_region / world_tops_kernel pick the array a
    // child run is read from by the tree LEVEL (children of a node whose cells have shift 0 are voxels, all others are
    // nodes), lookup() by the node's kind -- so every reachable node must have the kind its level implies, every child run
    // must lie inside the array that level reads, and a node must not be reachable at two levels.  Nodes no walk can
    // reach (runs abandoned by rc_update_world) are never indexed and are not looked at.
    {
        const int rootShift = 2 * ((int)w.max_depth - 2);
        std::vector<uint8_t> level(w.nodes.size(), 0xFF);  // shift / 2 of the node's child cells, 0xFF = not reached
        std::vector<uint32_t> todo;
        todo.push_back(0u);
        level[0] = (uint8_t)(rootShift / 2);
        while (!todo.empty()) {
            const uint32_t i = todo.back();
            todo.pop_back();
            const PackedNode& n = w.nodes[i];
            const uint32_t kind = n.kind & PK_KIND_MASK;
            const int shift = 2 * (int)level[i];
            const uint64_t cnt = (uint64_t)__builtin_popcountll(n.mask);
            if (shift == 0) {
                if (kind != PK_BRICK) return bad("a node at the brick level is not a brick");
                if (n.base + cnt > w.voxels.size()) return bad("voxel run outside the voxel array");
                continue;
            }
            if (kind != PK_INNER) return bad(kind == PK_BRICK ? "a brick above the brick level" : kind == PK_LEAF ? "a leaf with children" : "unknown node kind");
            if (n.base + cnt > w.nodes.size()) return bad("child run outside the node array");
            for (uint64_t c = 0; c < cnt; c++) {
                const uint32_t ci = n.base + (uint32_t)c;
                const PackedNode& ch = w.nodes[ci];
                const uint32_t ck = ch.kind & PK_KIND_MASK;
                if (ck == PK_LEAF) {
                    if ((ch.base & MATERIAL_INDEX_MASK) >= w.palette.size() || ch.base == 0) return bad("leaf material outside the palette");
                    continue;
                }
                if (ck != PK_INNER && ck != PK_BRICK) return bad("unknown node kind");
                const uint8_t lv = (uint8_t)(shift / 2 - 1);
                if (level[ci] == 0xFF) {
                    level[ci] = lv;
                    todo.push_back(ci);
                } else if (level[ci] != lv) {
                    return bad("a node is reachable at two tree levels");
                }
            }
        }
    }
    for (uint32_t v : w.voxels)
        if (v == 0 || (v & MATERIAL_INDEX_MASK) >= w.palette.size()) return bad("voxel material outside the palette");
    fclose(f);
    return RC_OK;
}

// Binary PPM (P6), top row first: rgba8 is bottom-up (GL convention), alpha dropped.
int write_ppm(const char* path, const uint8_t* rgba8, uint32_t width, uint32_t height, std::string& err) {
    FILE* f = path ? fopen(path, "wb") : nullptr;
    if (!f) {
        err = std::string("rc_write_ppm: cannot open ") + (path ? path : "(null)");
        return RC_ERR_INVALID_ARGUMENT;
    }
    fprintf(f, "P6\n%u %u\n255\n", width, height);
    std::vector<uint8_t> row((size_t)width * 3);
    bool ok = true;
    for (uint32_t y = 0; y < height && ok; y++) {
        const uint8_t* src = rgba8 + (size_t)(height - 1 - y) * width * 4;
        for (uint32_t x = 0; x < width; x++) {
            row[3 * x + 0] = src[4 * x + 0];
            row[3 * x + 1] = src[4 * x + 1];
            row[3 * x + 2] = src[4 * x + 2];
        }
        ok = fwrite(row.data(), 1, row.size(), f) == row.size();
    }
    ok = (fclose(f) == 0) && ok;
    if (!ok) {
        err = std::string("rc_write_ppm: short write to ") + path;
        return RC_ERR_INVALID_ARGUMENT;
    }
    return RC_OK;
}

}  // namespace rc
