// worldgen.h -- device-side world generation (worldgen.cu), called by rc_generate_world.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/rc.h"
#include "packed_world.h"

namespace rc {

// The generated world, left on the device in arrays that already have rc_update_world's append slack.
// Ownership of d_nodes / d_voxels passes to the caller.
struct GeneratedWorld {
    PackedNode* d_nodes = nullptr;
    uint32_t* d_voxels = nullptr;
    uint64_t n_nodes = 0, n_voxels = 0;
    uint64_t cap_nodes = 0, cap_voxels = 0;
    std::vector<uint32_t> level_begin;
    std::vector<PaletteEntry> palette;
    uint64_t leaves = 0;  // voxels + uniform cells: what the reference would hold as leaf nodes
    rc_gen_info info{};
};

int generate_world(const rc_gen_params& p, cudaStream_t stream, GeneratedWorld& out, std::string& err);

}  // namespace rc
