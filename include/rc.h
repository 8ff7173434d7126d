/* include/rc.h -- C-ABI of librc.so, the B200 voxel ray caster.
 *
 * Drop-in boundary for the render path of reedthorngag/raytracing_test.  The
 * reference has no FFI: its host talks to the GPU through OpenGL calls.  Each
 * entry point below replaces one group of those calls (cited as file:line of
 * the reference) and carries exactly the information they carried.  The host
 * (main.cpp, world_gen.cpp, input.cpp, voxel_data/, ray_caster.cpp) is kept
 * as it is; INTEGRATION.md shows the lines a maintainer changes.
 *
 * Conventions
 *  - plain C: pointers, sizes, PODs; no C++/torch types cross this boundary
 *  - process-global state, like the reference's GL context and globals; not
 *    thread-safe, one caller thread.  rc_init drives one GPU; rc_init_multi
 *    drives several from that one thread (the library shards every frame)
 *  - every int-returning call returns RC_OK (0) or a negative rc_status; the
 *    text of the last failure is rc_last_error().  Nothing ever calls exit()
 *    (the reference does: src/globals.hpp:32-43)
 *  - there is no CPU fallback: without a usable CUDA device rc_init fails
 */
#ifndef RC_H
#define RC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RC_ABI_VERSION 1

typedef enum rc_status {
    RC_OK = 0,
    RC_ERR_NOT_INITIALIZED = -1,
    RC_ERR_INVALID_ARGUMENT = -2,
    RC_ERR_CUDA = -3,
    RC_ERR_MALFORMED_WORLD = -4,
    RC_ERR_NO_WORLD = -5,
    RC_ERR_NO_FRAME = -6,
    RC_ERR_BUFFER_TOO_SMALL = -7,
    RC_ERR_UNSUPPORTED = -8
} rc_status;

/* ---- world ------------------------------------------------------------------
 * The host's two block pools, in the layout the shader binds as SSBOs
 * (src/shaders/low_res.frag:33-39; src/voxel_data/types.hpp:29-47):
 *   node pool : 16-byte nodes, index = byte offset >> 4   (binding 3)
 *   array pool: 64 x u32 child-index arrays, index = byte offset >> 8 (binding 2)
 * held as lists of 4 MiB malloc blocks (src/voxel_data/voxel_allocator.hpp:12-18).
 * Exactly what updateSsboData() walks (voxel_allocator.hpp:38-78). */
typedef struct rc_world_desc {
    const void* const* node_blocks;   /* nodeBlocks[i].ptr */
    const void* const* array_blocks;  /* arrayBlocks[i].ptr */
    uint32_t n_node_blocks;
    uint32_t n_array_blocks;
    uint32_t block_bytes;             /* BLOCK_SIZE = 1 << 22 */
    uint32_t root_index;              /* root.index (0) */
    uint64_t node_bytes_used;         /* nodeNextAllocIndex */
    uint64_t array_bytes_used;        /* arrayNextAllocIndex */
    uint32_t max_depth;               /* maxDepth = 6 (tetrahexa_tree.hpp:6); 7 and 8 are extensions */
    uint32_t reserved;
} rc_world_desc;

typedef struct rc_world_info {
    uint64_t ref_nodes;       /* reachable nodes of the host tree */
    uint64_t ref_bytes;       /* node_bytes_used + array_bytes_used */
    uint64_t packed_nodes;    /* 16-byte packed nodes on the device */
    uint64_t packed_voxels;   /* 4-byte voxel material words on the device */
    uint64_t packed_materials;
    uint64_t packed_bytes;    /* total device bytes of the flattened world */
    uint32_t max_depth;
    uint32_t null_children;   /* bitmap bits whose child index was 0 */
    double flatten_ms;
    double upload_ms;
} rc_world_info;

/* ---- frame ------------------------------------------------------------------
 * One field per uniform of low_res.frag:16-27 as set by render()
 * (src/main.cpp:83-95), plus the shader's compile-time constants as knobs.
 * originMortonPos (main.cpp:86) is derived inside the library. */
typedef struct rc_frame_params {
    float origin[3];         /* trueOrigin     <- cameraPos            main.cpp:83 */
    float camera_dir[3];     /* cameraDir (normalised by the host)     main.cpp:84,195 */
    float sun_dir[3];        /* sunDir         <- sun                  main.cpp:88 */
    float mouse_pos[2];      /* mousePos (crosshair), used with RC_F_CROSSHAIR  main.cpp:85 */
    float proj_plane[2];     /* projPlaneSize = (tan45, tan45*h/w)     main.cpp:94 */
    float delta_time;        /* deltaTime      <- glfwGetTime()        main.cpp:90 */
    float ao_strength;       /* colour *= 1 - ao_strength * occluded/ao_samples */
    int32_t looking_at[3];   /* lookingAtBlock <- castRayFromCam(30).pos  main.cpp:81,89 */
    int32_t render_pos_data; /* renderPosData  <- sendDebugFrame       main.cpp:87 */
    uint32_t width, height;  /* resolution                             main.cpp:93 */
    uint32_t max_steps;      /* numSteps = 300                         low_res.frag:310; <= 65535 (rc_hit.steps is 16 bits).
                                Above 1024 free flight is off (step counts are recovered from fp32 sums, exact up to there):
                                same frames, sky rays walk every step */
    uint32_t shadow_steps;   /* 75                                     low_res.frag:382; <= 65535 */
    uint32_t ao_samples;     /* 0 = AO off; 64 for BASELINE configs 3-5 (<= RC_MAX_AO_SAMPLES) */
    uint32_t ao_steps;       /* 5                                      light_scattering.frag:231 */
    uint32_t flags;          /* RC_F_* */
} rc_frame_params;

#define RC_F_CROSSHAIR 0x1u  /* white disc of radius 3 at mouse_pos (low_res.frag:259-262) */
#define RC_F_SHADOWS   0x2u  /* sun shadow ray (low_res.frag:373-391) */
#define RC_F_HITS      0x4u  /* also write the per-pixel parity records (rc_read_hits) */
#define RC_F_RGBA32F   0x8u  /* also keep the unclamped float colour (glReadPixels GL_FLOAT, main.cpp:38) */
#define RC_F_NO_FREE_FLIGHT 0x10u /* tests / measurements: primary rays that rise above everything they can still reach
                                     keep walking cells instead of finishing their steps in the free-flight loop
                                     (same frame, same hit records either way) */
#define RC_F_ALWAYS_TAIL 0x40u /* tests / measurements: work out where a free-flight ray ends even when nothing can show it
                                  (no hit records, no position debug output, looked-at voxel out of its reach) */
#define RC_F_NO_LEAP   0x20u /* tests / measurements: the free-flight loop walks every step instead of leaping over most of
                                them in closed form (same frame, same hit records either way) */

#define RC_F_NO_COLUMN_REGIONS 0x100u /* tests / measurements: primary rays above the highest voxel of their 16 x 16 column block
                                         look the empty space up in the tree, cell by cell, instead of taking the whole column
                                         above the terrain as one empty region (same frame, same hit records either way) */
#define RC_F_NO_DROP_GATES 0x200u /* tests / measurements: only primary rays that step UP can enter free flight; rays that step down
                                     but lose less height over their remaining steps than separates them from everything they can
                                     still reach walk on through the tree (same frame, same hit records either way) */
#define RC_F_KERNEL_TIMES 0x80u /* measurements: also time the primary and the AO kernel separately (rc_frame_stats.primary_ms /
                                   ao_ms).  The event between the two launches keeps the AO grid from becoming resident while
                                   the primary grid drains, and a fused frame (one kernel) has no such split: without this flag
                                   only total_ms is filled in */

#define RC_MAX_AO_SAMPLES 256u

/* Per-pixel parity record (the "hit buffer"): precedent = the shader's debug
 * outputs pos.round / pos.exact (low_res.frag:335-341). */
typedef struct rc_hit {
    int32_t x, y, z;        /* pos.round when the primary loops exit */
    uint16_t steps;         /* nextIntersectDDA calls of the primary phase */
    uint16_t shadow_steps;  /* nextIntersectDDA calls of the shadow ray */
    uint8_t face;           /* lastHit: 0 x, 1 y, 2 z */
    uint8_t kind;           /* RC_KIND_* */
    uint8_t flags;          /* RC_HF_* */
    uint8_t ao_occluded;    /* occluded AO rays of this pixel */
} rc_hit;

enum {
    RC_KIND_SKY = 0,            /* low_res.frag:352-355 */
    RC_KIND_HIT = 1,            /* low_res.frag:357-391 */
    RC_KIND_CAMERA_INSIDE = 2,  /* low_res.frag:300-306 */
    RC_KIND_CROSSHAIR = 3,      /* low_res.frag:259-262 */
    RC_KIND_HIGHLIGHT = 4,      /* low_res.frag:347-350 */
    RC_KIND_DEBUG = 5           /* low_res.frag:335-341 */
};
#define RC_HF_SHADOWED   0x01u
#define RC_HF_REFLECTED  0x02u
#define RC_HF_REFRACTED  0x04u
#define RC_HF_FACING_SUN 0x08u
#define RC_HF_STEPX_NEG  0x10u
#define RC_HF_STEPY_NEG  0x20u
#define RC_HF_STEPZ_NEG  0x40u

typedef enum rc_image_format {
    RC_RGBA8 = 0,    /* 4 B/pixel, alpha 0, row 0 = bottom (GL), what glfwSwapBuffers shows (main.cpp:216) */
    RC_RGBA32F = 1   /* 16 B/pixel, needs RC_F_RGBA32F (glReadPixels GL_RGBA/GL_FLOAT, main.cpp:38) */
} rc_image_format;

typedef struct rc_frame_stats {
    uint64_t primary_rays;   /* pixels that traced (all but crosshair pixels) */
    uint64_t shadow_rays;
    uint64_t ao_rays;
    uint64_t hit_pixels;     /* RC_KIND_HIT */
    uint64_t primary_steps;  /* DDA steps, summed (0 unless RC_F_HITS) */
    uint32_t kernel_launches;/* kernels launched by the last rc_render_frame */
    uint32_t free_flight_rays; /* primary rays that finished their step budget in free flight (kernels.cu, loop A) */
    float primary_ms;        /* RC_F_KERNEL_TIMES and a two-kernel schedule: device time of each kernel (CUDA events on the */
    float ao_ms;             /*   library's stream), else 0 */
    float total_ms;          /* device time of the whole frame; valid after rc_sync / rc_read_* */
    uint32_t schedule;       /* how the frame was scheduled: 0 primary kernel + AO kernel, 1 the same with a persistent AO grid
                                (programmatic dependent launch), 2 one fused kernel, 3 / 4 persistent fused kernel (global
                                atomic tile queue; 4: top tree levels staged in shared memory) */
} rc_frame_stats;

/* ---- lifetime -----------------------------------------------------------------
 * rc_init replaces createWindow()+setupOpenGl() (src/setup.cpp:55-111,414-426):
 * selects the CUDA device, creates the stream, loads the constant tables.
 * rc_shutdown replaces glfwDestroyWindow/glfwTerminate (src/main.cpp:229-230). */
int rc_init(int device);
void rc_shutdown(void);

/* Several GPUs of one box behind the same calls, from the same single host thread (the reference's host is one process with
 * one render loop, src/main.cpp:173-232): rc_init_multi(n, devices) instead of rc_init.  Every device holds the
 * whole world (rc_upload_world flattens once and copies to each; rc_update_world diffs once and patches each;
 * rc_generate_world runs on each); rc_render_frame deals the 16-row bands of the frame to the devices (band % n == i) and
 * queues every device's kernels on its own stream before any is waited for; rc_read_image / rc_read_image_async /
 * rc_read_hits copy each device's bands to their place in the caller's frame, every device over its own PCIe link (pass
 * page-locked memory, rc_host_register, for the copies to overlap) -- there is no device-side gather.  rc_frame_stats_get
 * sums the rays and reports the slowest device's times.  With rc_frame_target_create the devices' kernels ALSO store every
 * pixel into one frame on devices[0] through NVLink peer access (rc_frame_target_device; complete in that device's stream
 * order after rc_render_frame).  rc_set_stream, rc_set_shard and rc_frame_target_open belong to the one-process-per-GPU
 * deployment (rc_init) and are refused.  rc_pick, rc_device_image, rc_device_hits, rc_compose_frame act on devices[0]. */
int rc_init_multi(int n_devices, const int* devices);
int rc_device_count(void); /* devices the library drives (0 before rc_init) */
const char* rc_last_error(void);
int rc_abi_version(void);
/* sha256 (hex) of the kernel sources this library was compiled from; profiles/traffic.json records the digest of the
 * sources a profiler capture was taken with, so a reader can tell whether the two belong together. */
const char* rc_build_id(void);

/* Launch on a caller-owned cudaStream_t (e.g. torch's current stream) instead of
 * the library's own; NULL restores the internal stream. */
int rc_set_stream(void* cuda_stream);

/* Image sharding for one-process-per-GPU rendering: this process renders the bands of
 * 16 rows with band % world_size == rank (band b = rows 16 b .. 16 b + 15, counted from the bottom
 * row) and leaves the others untouched.  (0,1) = whole frame. */
int rc_set_shard(uint32_t rank, uint32_t world_size);

/* Layout of a sharded frame: how many rows this rank renders (local_rows, bands of 16 rows dealt
 * round-robin, the last band of the frame may be partial) and how many rows every rank's buffers
 * hold (padded_rows = the largest per-rank count rounded up to whole bands; equal on all ranks so a
 * gather can use equal-size messages). Needs no device. */
int rc_shard_layout(uint32_t height, uint32_t rank, uint32_t world_size, uint32_t* local_rows, uint32_t* padded_rows);

/* After a gather of every rank's rc_device_image / rc_device_hits (rank-major, world_size buffers
 * of padded_rows * width elements, DEVICE memory): de-interleave the bands into the bottom-up
 * frame at dst (DEVICE memory, width * height elements). bytes_per_element: 4 (RC_RGBA8),
 * 16 (RC_RGBA32F) or 20 (rc_hit). One kernel on the library's stream. */
int rc_compose_frame(const void* gathered, void* dst, uint32_t width, uint32_t height, uint32_t world_size,
                     uint32_t bytes_per_element);

/* ---- world upload -------------------------------------------------------------
 * Replaces initVoxelDataAllocator()+updateSsboData()
 * (src/voxel_data/voxel_allocator.hpp:38-91; called at src/main.cpp:182,212).
 * Copies out of the host blocks (the host keeps ownership), flattens the two
 * pools into the packed device format and uploads it.  Call again after the
 * host edited blocks (putBlock/deleteBlock set BlockPtr.modified). */
int rc_upload_world(const rc_world_desc* desc);
int rc_world_info_get(rc_world_info* out);

/* Incremental form of the same call, mirroring the BlockPtr.modified protocol: updateSsboData()
 * pushes exactly the 4 MiB blocks whose `modified` flag putBlock / deleteBlock set
 * (src/voxel_data/voxel_allocator.hpp:50-56,69-75; setters tetrahexa_tree.cpp:205,215,245,
 * 260-262,275,278) -- pass the indices of those blocks (i of `glBufferSubData(i << 22, ...)`).
 * The library diffs them against its copy of the pools, re-flattens only the packed subtrees that
 * depend on nodes / child arrays that really changed, appends them to the device arrays and patches
 * the subtree roots in place (n_dirty_* = RC_ALL_BLOCKS: diff every block -- for hosts that do not
 * trust the flags: deleteBlock's final branch, tetrahexa_tree.cpp:348-353, never sets one): a click costs a few memcmp's and a few KB of upload instead of a full
 * flatten.  Falls back to a full upload by itself (info->full_rebuild = 1) on the first call, when
 * the root or the depth changed, or when more than half of the device arrays has become garbage. */
#define RC_ALL_BLOCKS 0xFFFFFFFFu
typedef struct rc_update_info {
    uint32_t full_rebuild;        /* 1: did a full flatten + upload instead */
    uint32_t changed_ref_nodes;   /* 16-byte nodes that differ from the previous upload */
    uint32_t changed_ref_arrays;  /* 256-byte child arrays that differ */
    uint32_t reflattened;         /* packed subtrees re-emitted */
    uint64_t appended_nodes, appended_voxels;
    uint64_t garbage_nodes, garbage_voxels;  /* abandoned since the last full upload */
    uint64_t uploaded_bytes;      /* host -> device bytes of this update */
    double host_ms;               /* diff + re-flatten */
    double upload_ms;
} rc_update_info;
int rc_update_world(const rc_world_desc* desc, const uint32_t* dirty_node_blocks, uint32_t n_dirty_node_blocks,
                    const uint32_t* dirty_array_blocks, uint32_t n_dirty_array_blocks, rc_update_info* info);

/* ---- flattened world as a file ---------------------------------------------------
 * The reference has no on-disk format (its world lives only in the two malloc'd pools and is
 * regenerated by genWorld() at every start, src/main.cpp:188-190).  rc_save_world writes the
 * flattened (device) form of the uploaded world -- 64-byte header, the three arrays, FNV-1a
 * checksum; rc_load_world installs such a file directly, without the host pools and without
 * a flatten (checksum and every child run are validated first).  A loaded world has no update
 * baseline: the first rc_update_world after it is a full upload. */
int rc_save_world(const char* path);
int rc_load_world(const char* path);

/* ---- world generation on the device ----------------------------------------------
 * Replaces initTetraHexaTree() + genWorld() + the SSBO upload (src/main.cpp:181-212,
 * src/voxel_data/tetrahexa_tree.cpp:14-27, src/world_gen.cpp:13-42): the same world -- three
 * octaves of 2-D OpenSimplex noise (the reference's seeds 42, 64, 100 by default), water below
 * y = 20, grass / 3 x dirt / stone down each column, over the start-up test blocks -- is built
 * straight into the flattened device form, without the host tree (the reference spends one
 * putBlock per voxel on one core).  Voxel for voxel equal to the host-built world, and node for
 * node equal to rc_upload_world's flattening of it.  Like a loaded world it has no update
 * baseline.  Columns whose height falls below y = 0 (the reference wraps them around the torus;
 * the noise amplitudes make that practically unreachable) are refused with RC_ERR_UNSUPPORTED. */
typedef struct rc_gen_params {
    uint32_t width, length;  /* columns along x and z; reference WIDTH = LENGTH = 200 (world_gen.cpp:8-9) */
    uint32_t max_depth;      /* reference maxDepth = 6 (1024^3 torus); [3,8] */
    uint32_t startup_blocks; /* 1 = also place initTetraHexaTree's test blocks */
    int64_t seeds[3];        /* reference: 42, 64, 100 (world_gen.cpp:15-17) */
} rc_gen_params;
typedef struct rc_gen_info {
    uint64_t columns, candidate_bricks, bricks, nodes, voxels;
    int32_t min_height, max_height;
    float heights_ms; /* noise + height map */
    float build_ms;   /* brick masks, sort, tree levels, voxel fill */
    float total_ms;   /* wall clock of the call, allocations included */
} rc_gen_info;
int rc_generate_world(const rc_gen_params* params, rc_gen_info* info);

/* ---- multi-GPU: one frame written by every GPU ----------------------------------
 * With rc_set_shard each process renders only its bands.  Instead of gathering the band buffers
 * afterwards, the composing process (the "owner", usually rank 0) allocates the whole RC_RGBA8 frame
 * with rc_frame_target_create and hands the returned CUDA IPC handles (RC_IPC_HANDLE_BYTES each, one
 * per slot) to the other processes of the box, which map them with rc_frame_target_open.  From then
 * on every rc_render_frame of that resolution ALSO stores each finished pixel at its place in the
 * owner's frame, straight from the kernels over NVLink peer memory: the transfer rides along with the
 * tracing, pixel by pixel; there is no gather and no compose pass.  The frame is complete on the owner
 * once every process's frame has finished (order the streams with any cross-process signal, e.g. a
 * one-element all-reduce enqueued after rc_render_frame).  slots = 2 alternates between two frames
 * (frame N can be read while frame N+1 is written); all processes must render the same sequence of
 * frames.  The reference has no counterpart (single GPU). */
#define RC_IPC_HANDLE_BYTES 64
int rc_frame_target_create(uint32_t width, uint32_t height, uint32_t slots, void* handles_out);
int rc_frame_target_open(uint32_t width, uint32_t height, uint32_t slots, const void* handles);
int rc_frame_target_close(void);
void* rc_frame_target_device(uint32_t slot); /* owner only: the frame, bottom row first */
int rc_frame_target_last_slot(void);         /* slot the last rc_render_frame wrote, -1 = none */

/* Frame completion across processes without a collective per frame.  With a frame target every process's kernels store into
 * the owner's frame; what is left is telling the owner's stream that a frame is complete, and the peers' streams that a
 * frame slot may be written again.  The owner allocates a block of flags next to the frame (rc_frame_sync_create, one more
 * CUDA IPC handle), the peers map it (rc_frame_sync_open).  Frames are counted from 0 per sync block, the same on every
 * process.  For frame f:
 *   peer : rc_frame_sync_begin()  (its stream waits until the owner has allowed frame f)  -> rc_render_frame ->
 *          rc_frame_sync_end()    (a flag store behind its kernels: "rank r has finished frame f")
 *   owner: rc_frame_sync_allow(f + 1, stream) as soon as the slot frame f + 1 will use is free, queued on the stream that
 *          frees it (behind the read-back of the frame that was there; NULL = the library's stream);
 *          rc_render_frame -> rc_frame_sync_end() (its stream waits for every peer's flag: the frame is complete from there)
 * The waits are one-warp kernels polling LOCAL memory on the owner and one word of the owner's memory on the peers, with a
 * back-off; a wait that sees nothing for ~2 s gives up (rc_frame_sync_timeouts) instead of hanging the device.  Replaces the
 * one-element NCCL all-reduce used as the completion signal (~50 us per frame at 8 GPUs).  rc_init_multi does not need it
 * (one process: CUDA events). */
int rc_frame_sync_create(void* handle_out);       /* RC_IPC_HANDLE_BYTES */
int rc_frame_sync_open(const void* handle);
int rc_frame_sync_close(void);
int rc_frame_sync_begin(void);
int rc_frame_sync_allow(uint32_t frame, void* cuda_stream);
int rc_frame_sync_end(void);
uint32_t rc_frame_sync_frames(void);              /* frames ended so far = the index of the next one */
int rc_frame_sync_timeouts(uint32_t* out);

/* The last frame (RC_RGBA8) as a binary PPM, top row first: the headless stand-in for looking at
 * the window. Synchronises like rc_read_image. Sharded frames: this rank's rows, others black. */
int rc_write_ppm(const char* path);

/* ---- render -------------------------------------------------------------------
 * Replaces the glUniform* block + glDrawArrays of render() (src/main.cpp:83-107).
 * Asynchronous: enqueues the kernels and returns. */
int rc_render_frame(const rc_frame_params* params);

/* rc_render_frame for callers that cannot allocate device memory -- code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc.  No counterpart in the reference.
 * Same parameters, same pixels as rc_render_frame, but the RC_RGBA8 frame (bottom row first, width * height * 4 bytes) is
 * written into DEVICE memory the caller owns (rgba8_device), and everything else a frame needs -- counters, sky-gate
 * table, AO direction table, AO seed list -- lives in a fixed array inside the library (static __device__, 12.9 MB,
 * held by every GPU that loads librc.so until the process exits) instead of buffers grown with cudaMalloc.  The call
 * never calls cudaMalloc / cudaFree or any allocating routine; it only queues kernels on the library's stream
 * (rc_set_stream) and returns.  Limits: width * height <= RC_STATIC_MAX_PIXELS (e.g. 640 x 360 or 512 x 512; larger
 * frames are refused with RC_ERR_INVALID_ARGUMENT and a message, use rc_render_frame); RGBA8 only (RC_F_HITS, RC_F_RGBA32F,
 * RC_F_KERNEL_TIMES are refused); one device, unsharded (rc_init, not rc_init_multi / rc_set_shard); no frame target.
 * The first call with a new ao_samples or max_steps (or after a world change) also rebuilds the small tables (a kernel
 * and a few host-to-device copies; still no allocation).  rc_frame_stats_get, rc_read_image, rc_device_image keep
 * describing the last rc_render_frame, not this call.  Every call shares the one scratch array, which keeps the last
 * call's contents (what must start at zero is cleared by the call itself): calls that could overlap on the GPU -- issued
 * on different streams through rc_set_stream without ordering between them -- race on it; calls on one stream are
 * ordered and safe.  rc_sync (or a synchronise of the stream) makes the frame visible. */
#define RC_STATIC_MAX_PIXELS 262144u
int rc_render_frame_static(const rc_frame_params* params, void* rgba8_device, size_t rgba8_bytes);

/* Replaces glfwSwapBuffers / glReadPixels (src/main.cpp:216,38): waits for the
 * frame and copies it to host memory.  dst_bytes >= width*height*(4|16). */
int rc_read_image(void* dst, size_t dst_bytes, int format);

/* The same swap, double-buffered like the window's front / back buffers (glfwSwapBuffers returns
 * while the front buffer is still being scanned out): queues the copy of the last rendered RC_RGBA8
 * frame into dst on a separate copy stream and returns at once (dst should be pinned host memory,
 * cudaHostAlloc / cudaHostRegister, for the copy to be asynchronous).  The next rc_render_frame
 * renders into the other of two device frame buffers, so frame N+1 is traced while frame N travels
 * to the host; a third frame waits for the first one's copy.  rc_wait_image(ticket) blocks until
 * that frame is complete in dst.  One queued read-back per rendered frame. */
int rc_read_image_async(void* dst, size_t dst_bytes, int* ticket);
int rc_wait_image(int ticket);

/* Page-locks host memory the caller owns (malloc, mmap, POSIX shared memory) for asynchronous copies.
 * Multi-GPU hosts: map ONE shared-memory frame into every process, register it in each, and let every
 * process rc_read_image_async into it -- a sharded frame's read-back writes only this rank's rows, at
 * their place in the whole frame (rc_read_image does the same), so the frame assembles itself in host
 * memory over all the GPUs' PCIe links at once, with no device-side gather at all. */
int rc_host_register(void* ptr, size_t bytes);
int rc_host_unregister(void* ptr);
/* Full memory fence (std::atomic_thread_fence, seq_cst) for hosts whose processes signal "my bands have landed" through
 * counters that live next to the shared frame: fence before the counter store, fence after the counter load.  No device. */
void rc_host_fence(void);

/* Parity records of the last frame rendered with RC_F_HITS; n >= width*height. */
int rc_read_hits(rc_hit* dst, size_t n);

int rc_sync(void);
int rc_frame_stats_get(rc_frame_stats* out);

/* Device-side views of the last frame for host code that composes frames across
 * GPUs (NCCL gather over NVLink) without a host round trip. NULL if absent. */
void* rc_device_image(int format);
void* rc_device_hits(void);

/* ---- picking -------------------------------------------------------------------
 * RAY_CASTER::castRayFromCam(steps) (src/ray_caster.cpp:54-87) on the device, over the
 * uploaded world: the ray from the camera that feeds lookingAtBlock (src/main.cpp:81,89)
 * and the click handlers (src/input.cpp:143,156).  Same arithmetic: start voxel
 * trunc(cam_pos), DOUBLE deltas of the fp32 quotients 1/dir, one voxel per step, stops at
 * the first block that is neither empty nor LIQUID.  getBlock there is the host's
 * (src/voxel_data/tetrahexa_tree.cpp:113-157): coordinates wrap as a plain torus.
 * Synchronous (one 1-thread kernel + a 28-byte read-back). */
typedef struct rc_pick_result {
    int32_t pos[3];       /* RayResult.pos: pos.round when the loop ended */
    int32_t last_pos[3];  /* RayResult.lastPos: the voxel before the last step (where a block is placed) */
    int32_t steps;        /* RayResult.steps: steps left, 0 = nothing hit */
} rc_pick_result;
int rc_pick(const float cam_pos[3], const float cam_dir[3], int steps, rc_pick_result* out);

/* The hemisphere direction table the AO kernel uses: gen_hemisphare_distrib.py:4-12
 * with num_pts = n, as 3*n floats (x, y, up) in tangent space. */
int rc_hemisphere_table(uint32_t n, float* out);

#ifdef __cplusplus
}
#endif
#endif /* RC_H */
