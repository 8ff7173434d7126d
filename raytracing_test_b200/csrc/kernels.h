// kernels.h -- launch interface between the C-ABI layer (rc_api.cu) and the
// sm_100a kernels (kernels.cu). Device pointers + PODs only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rc.h"

namespace rc {

// Rows of the image are dealt to ranks in bands of RC_BAND_ROWS rows:
// band b belongs to rank b % world_size and is stored at local band b / world_size
// of that rank's (compact) frame buffers. world_size 1 => plain bottom-up frame.
constexpr uint32_t BAND_ROWS = 16;

struct DeviceWorld {
    const uint4* nodes;      // PackedNode[]
    const uint32_t* voxels;  // material words
    const uint4* palette;    // PaletteEntry[]
    uint32_t n_nodes;
    uint32_t n_top;          // BFS prefix staged in shared memory by the persistent kernel
    int root_shift;          // cell shift of the root's children: 2*(max_depth-2)
    uint32_t wrap;           // (1 << 2*(max_depth-1)) - 1
    uint32_t leak;           // 0x300 for max_depth 6 (the reference's Morton-spread quirk), else 0
    // Top table: for every aligned cell of shift top_shift (2^top_bits cells per axis, at most 64: 16^3 cells for the
    // reference's 1024^3 torus) the index of the node of that cell | 0x80000000, or the shift of the empty cell around it.
    // A lookup that has left its cached node starts here instead of at the root.  Null: descents start at the root
    // (trees of depth 2; rc_debug_set_option "top_table" 0).  Rebuilt on the device after every world change.
    const uint32_t* top_table;
    int top_shift, top_bits;
};

// Per-hit-pixel record handed from the primary kernel to the AO kernel.
struct AoSeed {
    float ex, ey, ez;   // un-biased pos.exact at the hit (low_res.frag:343-345)
    int x, y, z;        // hit voxel
    float r, g, b;      // shaded colour before AO
    uint32_t pixel;     // index into the (compact) frame buffers
    uint32_t face_sign; // bits 0-1 face, bit 2: normal sign negative
    uint32_t frame_pixel; // index into the whole frame (py * width + px)
};

// Frame statistics and work-queue cursors.  Every warp adds to the statistics, so they are spread over COUNTER_SLOTS
// copies, each on a 128-byte line of its own (one L2 line taking 300 k atomics per frame serialises them; the host adds the
// copies up).  The seed-list cursor and the two queue cursors return values the warps wait for: they sit on lines nothing
// else touches (slot 0's ao_seed_count is THE seed-list cursor of the two-kernel schedules; fused frames count their
// seeds per slot).
struct alignas(128) FrameCounters {
    unsigned long long primary_rays;
    unsigned long long shadow_rays;
    unsigned long long hit_pixels;
    unsigned long long primary_steps;
    unsigned int ao_seed_count;
    unsigned int free_flight_rays;  // primary rays that ended in the free-flight loop
};
struct alignas(128) QueueCursor {
    unsigned int next;
};
constexpr uint32_t COUNTER_SLOTS = 64;
struct FrameCounterBlock {
    QueueCursor seed_cursor;   // two-kernel schedules: next free entry of the AO seed list = seeds appended
    QueueCursor tile_cursor;   // persistent frame kernel: next 8x4-pixel tile (global atomic work queue)
    QueueCursor ao_cursor;     // persistent AO kernel: next 32-seed chunk (global atomic work queue)
    FrameCounters slot[COUNTER_SLOTS];
};

struct FrameArgs {
    rc_frame_params p;
    DeviceWorld world;
    uint32_t rank, world_size;
    float inv_width, inv_height;  // 1.0f / width, 1.0f / height (low_res.frag:264: 1.0 / resolution)
    // Free flight (primary loop A).  The world's columns are summarised in a SKY_GRID x SKY_GRID grid of column blocks
    // (edge B = 2^sky_shift voxels).  sky_gate[q][bz][bx] holds, for a ray in block (bx, bz) heading the way of quadrant
    // q = (step.x < 0) | (step.z < 0) << 1, five gates: 1 + the highest occupied y over the blocks at Chebyshev distance
    // 0-1, 2-3, 4-7, 8-15 and 16+ blocks in that quadrant, as far as max_steps steps can reach (u16 x 8, one 128-bit load;
    // 0 = nothing there, 65535 = the ring runs off the grid: torus wrap, never free).  A ray that steps UP is above
    // everything it can still reach when its y clears the near gate and, for each farther ring, its y plus what it
    // must have climbed before it can get there (slope x distance, rounded down, less 3) clears that ring's gate: then
    // nothing but empty space lies ahead.  Null: no gating (camera y + max_steps leaves the torus, max_steps > 1024,
    // tree too shallow).
    const uint4* sky_gate;
    int sky_shift;
    // Column regions (primary loop A).  col_tops[bz][bx] = highest occupied y of column block (bx, bz) (-1: none), the same
    // table the gates are built from.  Everything above it is empty: a ray above its block's top walks the whole
    // 2^sky_shift x 2^sky_shift column -- down to the top, up to the edge of the torus -- as ONE empty region, found with
    // one load instead of a tree descent.  Null: off (RC_F_NO_COLUMN_REGIONS, tree too shallow).
    const int* col_tops;
    uint32_t sky_tail;       // 1: a free ray's final voxel / last face are needed (hit records, position debug output, or the
                             // looked-at voxel lies exactly max_steps steps from the camera's); 0: nobody can see them
    uint32_t shadow_gate;    // the table also covers the shadow rays' reach (shadow_steps <= max_steps, flags allow it)
    uint32_t drop_gate;      // rays that step DOWN are tested against the table too (sky_drop_distance; RC_F_NO_DROP_GATES: off)
    uint32_t local_rows;     // rows stored locally = owned bands * BAND_ROWS (last band may be partial)
    uint32_t n_local_bands;
    uchar4* rgba8;           // [local_rows][width]
    uchar4* frame_target;    // optional: the whole frame [height][width] on the composing GPU (peer memory)
    float4* rgba32f;         // optional
    rc_hit* hits;            // optional
    AoSeed* ao_seeds;        // optional (ao_samples > 0)
    FrameCounterBlock* counters;
    const void* ao_table;    // AoDir[6][RC_MAX_AO_SAMPLES] of the window AO code (device), set by launch_primary / launch_ao
    uint32_t ao_opts;        // AO_OPT_*
    // the sun's ray (low_res.frag:373: buildRay(sunDir)) is the same for every pixel: 1 / sunDir, IEEE fp32, divided once on the host
    float sun_inv[3];
    // camera basis (low_res.frag:279-283), the same for every pixel: left = cross(cameraDir, (0,1,0)) (not normalised),
    // up = cross(cameraDir, left); computed once on the host with the same single-rounded fp32 operations
    float cam_left[3], cam_up[3];
    uint32_t debug_primary_smem;  // A/B probe (rc_debug_set_option "primary_smem"): dynamic shared memory the non-fused primary
                                  // kernel is launched with although it uses none -- shows what the fused kernels' smaller L1 costs
};

// Pieces of the window AO code that can be switched off for A/B measurements and tests (same results either way).
constexpr uint32_t AO_OPT_OCTANT = 1u;   // sample groups whose octant of reachable voxels is empty are not cast
constexpr uint32_t AO_OPT_DEFAULT = AO_OPT_OCTANT;

// How a frame is scheduled on the device (rc_debug_set_option "schedule"; same pixels either way).
enum FrameSchedule : uint32_t {
    SCHED_TWO_KERNELS = 0,     // primary_kernel, then ao_window_kernel over the seed list (one CTA per 128 seeds)
    SCHED_AO_PERSISTENT = 1,   // primary_kernel, then a persistent AO grid pulling 32-seed chunks from a global atomic cursor,
                               // launched with programmatic dependent launch (its prologue overlaps the primary tail)
    SCHED_FUSED = 2,           // one kernel: the warp that traced an 8x4 tile casts the AO rays of its hit pixels itself
                               // (no seed list, no second launch)
    SCHED_PERSISTENT = 3,      // SCHED_FUSED as persistent warps over a global atomic queue of Morton-ordered 8x4 tiles
    SCHED_PERSISTENT_STAGED = 4, // SCHED_PERSISTENT with the top tree levels (DeviceWorld::n_top nodes) staged in shared memory
    SCHED_COUNT,
    // default: the two-kernel schedules, whichever is faster for the frame's size (measured, DESIGN section 4: below ~4 M
    // pixels per device the one-CTA-per-128-seeds AO grid, above it the persistent AO grid; the fused kernels lose to both --
    // their code outgrows the instruction cache once warps of one SM are in different phases)
    SCHED_AUTO = SCHED_COUNT
};
constexpr uint32_t SCHED_DEFAULT = SCHED_AUTO;

// Bytes of the oriented AO direction table (AoDir[6][RC_MAX_AO_SAMPLES], 32 bytes per entry; kernels.cu asserts it).
constexpr size_t AO_TABLE_BYTES = 6u * RC_MAX_AO_SAMPLES * 32u;

// Loads the AO direction table (3*n floats: x, y, up) into __constant__ memory of the current device and the oriented
// table of the window AO code into *table (device memory, AO_TABLE_BYTES, allocated on first use -- a non-null *table is
// used as it is, nothing is allocated; pass it on as FrameArgs::ao_table).
cudaError_t upload_hemisphere(const float* dirs, uint32_t n, cudaStream_t stream, void** table);
void release_hemisphere(void* table);

// Whole frame: primary + shading + shadow rays, then (or fused with) the AO rays, scheduled as `schedule` says.
// mid_event (optional): recorded between the primary and the AO kernel when they are separate launches (per-kernel
// times; it also keeps the AO launch from overlapping the primary tail).  *launches = kernels launched.
// A schedule that cannot serve the frame (the window AO code needs ao_steps == 5) falls back to SCHED_TWO_KERNELS.
cudaError_t launch_frame(const FrameArgs& args, uint32_t schedule, int num_sms, cudaEvent_t mid_event, cudaStream_t stream,
                         uint32_t* launches, uint32_t* schedule_used);

// De-interleave gathered per-rank band buffers into a bottom-up frame (see rc_compose_frame).
cudaError_t launch_compose(const void* gathered, void* dst, uint32_t width, uint32_t height, uint32_t world_size,
                           uint32_t padded_rows, uint32_t bytes_per_element, cudaStream_t stream);

// castRayFromCam (src/ray_caster.cpp:54-87) as a 1-thread kernel; out = device int[7]: pos, lastPos, steps left.
cudaError_t launch_pick(const DeviceWorld& world, const float* cam_pos3, const float* cam_dir3, int steps, int* out7,
                        cudaStream_t stream);

// Highest occupied y per column block of the packed tree -> tops = device int[SKY_GRID * SKY_GRID] ([bz][bx], -1 = empty);
// block edge = 2^sky_shift voxels.  One depth-first walk per 64^3 subtree.
constexpr int SKY_GRID = 64;
cudaError_t launch_world_tops(const DeviceWorld& world, int sky_shift, int* tops, cudaStream_t stream);
// Fills world.top_table (TOP_TABLE_MAX_CELLS entries at most) from the packed tree: one thread per cell.
constexpr uint32_t TOP_TABLE_MAX_CELLS = 64u * 64u * 64u;
cudaError_t launch_top_table(const DeviceWorld& world, uint32_t* table, cudaStream_t stream);
// tops -> the gate table of FrameArgs::sky_gate for a step budget (4 * SKY_GRID * SKY_GRID entries of 16 bytes).
cudaError_t launch_sky_gates(const int* tops, int sky_shift, uint32_t max_steps, uint4* gates, cudaStream_t stream);

// Cross-process frame completion without a collective (rc_frame_sync_*): flags in the frame owner's memory.
// signal: *flag = max(*flag, value), after everything queued before it on the stream (system-scope fence).
// wait: lanes 0 .. count-1 each spin on flags[lane * stride_words] until it has reached `value` (wrap-safe compare), with
// a back-off; gives up after ~2 s and raises *timed_out instead of hanging the device.
cudaError_t launch_sync_signal(uint32_t* flag, uint32_t value, cudaStream_t stream);
cudaError_t launch_sync_wait(const uint32_t* flags, uint32_t stride_words, uint32_t count, uint32_t value, uint32_t* timed_out,
                             cudaStream_t stream);

// Single-voxel lookups through the same device code path the kernels use (tests).
cudaError_t launch_lookup(const DeviceWorld& world, const int* xyz, uint32_t n, uint32_t* out_words, int* out_shift,
                          cudaStream_t stream);

// Device versions of the deterministic sin/exp (tests compare them with the oracle's).
cudaError_t launch_detmath(const float* in, uint32_t n, float* out_sin, float* out_exp, cudaStream_t stream);

}  // namespace rc
