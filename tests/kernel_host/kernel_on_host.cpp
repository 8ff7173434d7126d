// kernel_on_host.cpp -- TEST INFRASTRUCTURE (tests/test_kernel_source_cpu.py builds and drives it).
//
// The device code of raytracing_test_b200/csrc/kernels.cu, compiled by g++ through tests/kernel_host/cuda_on_host.h and run on
// the host.  What reaches g++ is kernels.cu itself up to its first host-side launch helper, with its five inline-PTX lines
// rewritten (KERNEL_SOURCE, written by the test: further down the file has kernel launches that only nvcc can parse).
//   per-pixel mode (khost_render): the table kernels (world_tops_kernel, top_table_kernel, sky_gate_kernel) exactly as rc_api.cu
//     launches them, then shade_pixel() for every pixel -- camera ray, loop A over tree regions / column regions / the top
//     table, sky gates and drop gates, free flight with the closed-form leap, replay of pos.exact, loop B (mirrors,
//     refraction), shading, the shadow ray -- and the stores of trace_tile.  Whole 1080p frames in seconds.
//   kernel mode (khost_render_kernels): the __global__ functions themselves (primary_kernel<false / true>, ao_window_kernel,
//     ao_window_persistent_kernel, ao_cached_kernel, frame_persistent_kernel<false / true>), CTA by CTA as launch_frame
//     schedules them, every warp as 32 coroutines that meet at each *_sync intrinsic, the CTA's shared memory a plain array.
// Nothing here is used by the product.  Not modelled: concurrency between warps (they run one after the other).
#include "cuda_on_host.h"

#include KERNEL_SOURCE  // namespace rc { ... device code of kernels.cu ... }
#undef __noinline__

// the kernels' `extern __shared__ uint32_t smem[]`: the shared memory of the one CTA that is being emulated
namespace rc {
alignas(16) uint32_t smem[48 * 1024];
}
unsigned char* kh_shared_base = reinterpret_cast<unsigned char*>(rc::smem);

using namespace rc;

namespace {

struct HostWorld {
    PackedWorld pw;
    DeviceWorld W{};
    std::vector<uint32_t> top;
    std::vector<int> tops;
    std::vector<uint4> gates;
    uint32_t gateSteps = 0xFFFFFFFFu;
    int skyShift = 0;
    FrameCounterBlock counters;
};

std::string lastError;

// a 1-D launch of a __global__ function, one thread after the other
template <typename F>
void launch1d(unsigned blocks, unsigned threads, F&& body) {
    gridDim.x = blocks;
    blockDim.x = threads;
    for (unsigned b = 0; b < blocks; b++)
        for (unsigned t = 0; t < threads; t++) {
            blockIdx.x = b;
            threadIdx.x = t;
            body();
        }
    blockIdx.x = threadIdx.x = 0;
}

}  // namespace

extern "C" {

const char* khost_last_error(void) { return lastError.c_str(); }

// rc_upload_world's host half (flatten.cpp) + publishWorld + refreshSky of rc_api.cu, with the kernels run on the host.
void* khost_world_create(const rc_world_desc* desc) {
    HostWorld* hw = new HostWorld();
    if (flatten_world(*desc, hw->pw, lastError) != RC_OK) {
        delete hw;
        return nullptr;
    }
    DeviceWorld& w = hw->W;
    const uint32_t maxDepth = hw->pw.max_depth;
    w.nodes = reinterpret_cast<const uint4*>(hw->pw.nodes.data());
    w.voxels = hw->pw.voxels.data();
    w.palette = reinterpret_cast<const uint4*>(hw->pw.palette.data());
    w.n_nodes = (uint32_t)hw->pw.nodes.size();
    w.n_top = 0;
    w.root_shift = 2 * ((int)maxDepth - 2);
    w.wrap = (1u << (2 * (maxDepth - 1))) - 1u;
    w.leak = maxDepth == 6 ? 0x300u : 0u;
    // top table (rc_api.cu refreshSky)
    w.top_table = nullptr;
    w.top_shift = w.root_shift - 4 > 4 ? w.root_shift - 4 : 4;
    w.top_bits = w.root_shift + 2 - w.top_shift;
    if (w.top_bits >= 0 && w.top_bits <= 6) {
        hw->top.assign((size_t)1 << (3 * w.top_bits), 0u);
        const uint32_t cells = 1u << (3 * w.top_bits);
        uint32_t* table = hw->top.data();
        launch1d((cells + 127u) / 128u, 128, [&] { top_table_kernel(w, table); });
        w.top_table = table;
    }
    // column-block tops (launch_world_tops: memset to -1, 64^3 threads)
    hw->skyShift = w.root_shift + 2 - 6;
    if (hw->skyShift >= 2) {
        hw->tops.assign((size_t)SKY_GRID * SKY_GRID, -1);
        int* tops = hw->tops.data();
        const int bs = hw->skyShift;
        launch1d(64 * 64 * 64 / 128, 128, [&] { world_tops_kernel(w, bs, tops); });
    }
    return hw;
}

void khost_world_destroy(void* world) { delete static_cast<HostWorld*>(world); }

// Lookup through the device routine (lookup(): top table, cursor, popcount addressing): word + cell shift of n voxels.
void khost_lookup(void* world, const int32_t* xyz, uint32_t n, uint32_t* words, int32_t* shifts) {
    HostWorld* hw = static_cast<HostWorld*>(world);
    for (uint32_t i = 0; i < n; i++) {
        Cursor cur;
        cursor_reset(hw->W, cur);
        int shift = -1;
        words[i] = lookup(hw->W, cur, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], shift);
        shifts[i] = shift;
    }
}

// The FrameArgs rc_render_frame builds (one_render_frame of rc_api.cu), for an unsharded frame.
static void frameArgs(HostWorld* hw, const rc_frame_params& p, FrameArgs& a) {
    memset(&a, 0, sizeof(a));
    a.p = p;
    a.world = hw->W;
    a.rank = 0;
    a.world_size = 1;
    a.inv_width = 1.0f / (float)p.width;
    a.inv_height = 1.0f / (float)p.height;
    {
        const float oy = floorf(p.origin[1]);
        if (!(p.flags & RC_F_NO_FREE_FLIGHT) && p.max_steps <= 1024u && hw->skyShift >= 2 && oy >= 0.0f &&
            oy + (float)p.max_steps <= (float)hw->W.wrap) {
            if (hw->gateSteps != p.max_steps) {
                hw->gates.assign((size_t)4 * SKY_GRID * SKY_GRID, make_uint4(0, 0, 0, 0));
                const int* tops = hw->tops.data();
                uint4* gates = hw->gates.data();
                const int bs = hw->skyShift;
                const uint32_t ms = p.max_steps;
                launch1d(4 * SKY_GRID * SKY_GRID / 128, 128, [&] { sky_gate_kernel(tops, bs, ms, gates); });
                hw->gateSteps = p.max_steps;
            }
            a.sky_gate = hw->gates.data();
            a.sky_shift = hw->skyShift;
            a.shadow_gate = p.shadow_steps <= p.max_steps ? 1u : 0u;
            a.drop_gate = (p.flags & RC_F_NO_DROP_GATES) ? 0u : 1u;
            long long manhattan = 0;
            for (int k = 0; k < 3; k++) manhattan += llabs((long long)p.looking_at[k] - (long long)floorf(p.origin[k]));
            a.sky_tail = ((p.flags & (RC_F_HITS | RC_F_ALWAYS_TAIL)) || p.render_pos_data != 0 || manhattan == (long long)p.max_steps) ? 1u : 0u;
        }
    }
    if (!(p.flags & RC_F_NO_COLUMN_REGIONS) && hw->skyShift >= 2 && !hw->tops.empty()) {
        a.col_tops = hw->tops.data();
        a.sky_shift = hw->skyShift;
    }
    a.n_local_bands = (p.height + BAND_ROWS - 1) / BAND_ROWS;
    a.local_rows = a.n_local_bands * BAND_ROWS;
    memset(&hw->counters, 0, sizeof(hw->counters));
    a.counters = &hw->counters;
    a.ao_opts = AO_OPT_DEFAULT;
    for (int k = 0; k < 3; k++) a.sun_inv[k] = 1.0f / p.sun_dir[k];
    {
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
}

// Per-pixel mode: shade_pixel for every pixel + the stores of trace_tile, no warps (fast: whole 1080p frames in seconds).
// hits: width * height records (or null), rgba8: width * height * 4 bytes bottom-up, rgba32f: width * height * 4 floats or
// null.  stats[0] = rays that ended in free flight, stats[1] = hit pixels, stats[2] = shadow rays.
int khost_render(void* world, const rc_frame_params* params, rc_hit* hits, uint8_t* rgba8, float* rgba32f, uint64_t* stats) {
    HostWorld* hw = static_cast<HostWorld*>(world);
    const rc_frame_params& p = *params;
    if (p.ao_samples != 0) {
        lastError = "khost_render: per-pixel mode runs shade_pixel only (ao_samples must be 0); khost_render_kernels runs the kernels";
        return RC_ERR_UNSUPPORTED;
    }
    FrameArgs a;
    frameArgs(hw, p, a);
    uint64_t hitPixels = 0, shadowRays = 0;
    for (uint32_t py = 0; py < p.height; py++)
        for (uint32_t px = 0; px < p.width; px++) {
            blockIdx.x = px / 16;  // counter_slot
            blockIdx.y = py / 8;
            threadIdx.x = 0;
            PixelResult R;
            shade_pixel(a, true, px, py, R);
            const size_t idx = (size_t)py * p.width + px;
            // store_pixel / trace_tile
            rgba8[4 * idx + 0] = to_unorm8(R.color.x);
            rgba8[4 * idx + 1] = to_unorm8(R.color.y);
            rgba8[4 * idx + 2] = to_unorm8(R.color.z);
            rgba8[4 * idx + 3] = 0;
            if (rgba32f) {
                rgba32f[4 * idx + 0] = R.color.x;
                rgba32f[4 * idx + 1] = R.color.y;
                rgba32f[4 * idx + 2] = R.color.z;
                rgba32f[4 * idx + 3] = 0.0f;
            }
            if (hits) hits[idx] = R.hit;
            if (R.hit.kind == RC_KIND_HIT) hitPixels++;
            if (R.shadowRay) shadowRays++;
        }
    blockIdx.x = blockIdx.y = 0;
    if (stats) {
        uint64_t free = 0;
        for (uint32_t k = 0; k < COUNTER_SLOTS; k++) free += hw->counters.slot[k].free_flight_rays;
        stats[0] = free;
        stats[1] = hitPixels;
        stats[2] = shadowRays;
    }
    return RC_OK;
}

// ---- kernel mode: the __global__ functions themselves, CTA by CTA, every warp as 32 coroutines (cuda_on_host.h, KhWarp) ------
namespace {
struct KernelCall {
    void (*kernel)(const FrameArgs);
    const FrameArgs* args;
};
void warpBody(void* p) {
    const KernelCall* c = static_cast<const KernelCall*>(p);
    c->kernel(*c->args);
}
KhWarp warp;

// <<<grid, threads, (shared memory), stream>>> on the host: CTAs one after the other, their warps one after the other (none
// of these kernels synchronises across warps)
void launch(void (*kernel)(const FrameArgs), const FrameArgs& a, unsigned gx, unsigned gy, unsigned threads) {
    KernelCall call{kernel, &a};
    gridDim.x = gx;
    gridDim.y = gy;
    blockDim.x = threads;
    for (unsigned by = 0; by < gy; by++)
        for (unsigned bx = 0; bx < gx; bx++) {
            blockIdx.x = bx;
            blockIdx.y = by;
            for (unsigned w = 0; w < threads / 32u; w++) warp.run(32u * w, warpBody, &call);
        }
    blockIdx.x = blockIdx.y = 0;
    threadIdx.x = 0;
    gridDim.x = gridDim.y = 1;
}
}  // namespace

// rc_render_frame as launch_frame (kernels.cu) schedules it, on the host.  schedule: 0 primary_kernel<false> + ao_window_kernel
// (ao_cached_kernel when ao_steps != 5), 1 + ao_window_persistent_kernel, 2 primary_kernel<true> (fused), 3 / 4
// frame_persistent_kernel<false / true>.  hemi: rc_hemisphere_table(ao_samples) (3 floats per sample) when ao_samples > 0.
// schedule | 0x100: the AO octant skip off; | 0x200 / 0x400: the lanes of a warp run in descending / shuffled order between
// two rendezvous (a race check: results must not depend on it).  stats[0] = free-flight rays, [1] = hit pixels, [2] = shadow rays, [3] = AO seeds.
int khost_render_kernels(void* world, const rc_frame_params* params, uint32_t schedule, const float* hemi, rc_hit* hits, uint8_t* rgba8,
                         float* rgba32f, uint64_t* stats) {
    HostWorld* hw = static_cast<HostWorld*>(world);
    const rc_frame_params& p = *params;
    FrameArgs a;
    frameArgs(hw, p, a);
    const size_t pixels = (size_t)a.local_rows * p.width;
    std::vector<uchar4> img(pixels);
    std::vector<float4> f32(rgba32f ? pixels : 0);
    std::vector<rc_hit> hit(hits ? pixels : 0);
    std::vector<AoSeed> seeds(p.ao_samples ? pixels : 0);
    a.rgba8 = img.data();
    a.rgba32f = rgba32f ? f32.data() : nullptr;
    a.hits = hits ? hit.data() : nullptr;
    a.ao_seeds = p.ao_samples ? seeds.data() : nullptr;
    void* table = nullptr;
    if (p.ao_samples) {
        if (!hemi || upload_hemisphere(hemi, p.ao_samples, nullptr, &table) != cudaSuccess) {
            lastError = "khost_render_kernels: upload_hemisphere failed";
            return RC_ERR_INVALID_ARGUMENT;
        }
        a.ao_table = table;
    }
    const bool ao = p.ao_samples > 0;
    const bool window = ao && p.ao_steps == AOW_STEPS;
    if (schedule & 0x100u) a.ao_opts = 0u;  // A/B: AO_OPT_OCTANT off
    warp.order = (schedule >> 9) & 3u;      // 0x200: lanes run in descending order between rendezvous, 0x400: shuffled
    schedule &= 0xFFu;
    if (schedule >= 2 && ao && !window) schedule = 0;
    if (schedule == 1 && !window) schedule = 0;
    const unsigned gx = (p.width + PRIMARY_TILE_W - 1u) / PRIMARY_TILE_W, gy = (a.local_rows + PRIMARY_TILE_H - 1u) / PRIMARY_TILE_H;
    if (schedule == 3 || schedule == 4) {
        if (schedule == 4) a.world.n_top = hw->pw.level_begin.size() > 3 ? hw->pw.level_begin[3] : 0u;
        launch(schedule == 4 ? frame_persistent_kernel<true> : frame_persistent_kernel<false>, a, 8, 1, 32 * PERSISTENT_WARPS);
    } else if (schedule == 2) {
        launch(primary_kernel<true>, a, gx, gy, 32 * RC_PRIMARY_WARPS);
    } else {
        launch(primary_kernel<false>, a, gx, gy, 32 * RC_PRIMARY_WARPS);
        if (ao) {
            const unsigned maxSeeds = a.local_rows * p.width;
            if (schedule == 1)
                launch(ao_window_persistent_kernel, a, 8, 1, AO_THREADS);
            else
                launch(window ? ao_window_kernel : ao_cached_kernel, a, (maxSeeds + AO_THREADS - 1u) / AO_THREADS, 1, AO_THREADS);
        }
    }
    if (table) release_hemisphere(table);
    for (uint32_t y = 0; y < p.height; y++) {
        memcpy(rgba8 + (size_t)y * p.width * 4, &img[(size_t)y * p.width], (size_t)p.width * 4);
        if (rgba32f) memcpy(rgba32f + (size_t)y * p.width * 4, &f32[(size_t)y * p.width], (size_t)p.width * 16);
        if (hits) memcpy(hits + (size_t)y * p.width, &hit[(size_t)y * p.width], (size_t)p.width * sizeof(rc_hit));
    }
    if (stats) {
        stats[0] = stats[1] = stats[2] = 0;
        stats[3] = hw->counters.seed_cursor.next;
        for (uint32_t k = 0; k < COUNTER_SLOTS; k++) {
            stats[0] += hw->counters.slot[k].free_flight_rays;
            stats[1] += hw->counters.slot[k].hit_pixels;
            stats[2] += hw->counters.slot[k].shadow_rays;
            stats[3] += hw->counters.slot[k].ao_seed_count;
        }
    }
    return RC_OK;
}

}  // extern "C"
