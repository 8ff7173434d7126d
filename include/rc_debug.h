/* include/rc_debug.h -- test hooks exported by librc.so next to the rc.h ABI.
 *
 * Not part of the drop-in boundary: nothing in the reference corresponds to
 * these.  They let the parity tests reach inside the library (the flattened
 * world, the device lookup routine, the deterministic sin/exp) through the same
 * shared object the host links.
 */
#ifndef RC_DEBUG_H
#define RC_DEBUG_H

#include "rc.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Flatten on the host only (no CUDA device needed, rc_init not required): the
 * flattener half of rc_upload_world.  Result is kept for rc_debug_lookup_host(1,..). */
int rc_debug_flatten_host(const rc_world_desc* desc, rc_world_info* info);

/* Lookup in a host copy of the packed world. which: 0 = the world uploaded with
 * rc_upload_world, 1 = the result of rc_debug_flatten_host.
 * out_word: material word (0 = empty); out_shift: log2 edge of the uniform cell;
 * out_leaf[4]: the reference Leaf {u64 packedColor, u32 flags, f32 metadata} when solid. */
int rc_debug_lookup_host(int which, int32_t x, int32_t y, int32_t z, uint32_t* out_word, int32_t* out_shift,
                         uint32_t* out_leaf);
int rc_debug_lookup_host_batch(int which, const int32_t* xyz, uint64_t n, uint32_t* out_words, int32_t* out_shifts,
                               uint32_t* out_leaves);

/* The same lookup executed by the device routine the kernels use. Host pointers. */
int rc_debug_lookup_device(const int32_t* xyz, uint32_t n, uint32_t* out_words, int32_t* out_shifts);

/* Palette entry of a material word of the uploaded world. */
int rc_debug_palette(uint32_t word, uint32_t* out_leaf);

/* Device det_sin / det_exp (DESIGN.md "deterministic transcendentals"). Host pointers. */
int rc_debug_detmath(const float* in, uint32_t n, float* out_sin, float* out_exp);

/* Host-only halves of rc_upload_world / rc_update_world with tracking (no CUDA needed): flatten into the
 * debug world (which = 1 of rc_debug_lookup_host) keeping the update maps, then patch it incrementally. */
int rc_debug_flatten_host_tracked(const rc_world_desc* desc, rc_world_info* info);
int rc_debug_update_host(const rc_world_desc* desc, const uint32_t* dirty_node_blocks, uint32_t n_dirty_node_blocks,
                         const uint32_t* dirty_array_blocks, uint32_t n_dirty_array_blocks, rc_update_info* info);

/* Host-only halves of rc_save_world / rc_load_world on the debug world (which = 1). */
int rc_debug_save_host(const char* path);
int rc_debug_load_host(const char* path, rc_world_info* info);

/* Scheduling / specialisation switches for A/B measurements and parity tests; every setting renders the same pixels.
 *   "schedule": 0 primary kernel + AO kernel, 1 primary kernel + persistent AO grid (atomic chunk cursor, programmatic
 *               dependent launch), 2 one fused kernel, 3 persistent fused kernel over a global atomic queue of
 *               Morton-ordered 8x4 tiles, 4 the same with the top tree levels staged in shared memory, 5 (default): 0 or 1,
 *               whichever is faster for the frame's size
 *   "ao_opts" : bit 0 empty sample octants skipped (default 1) */
int rc_debug_set_option(const char* name, int value);

/* Copy-only ceiling of the read-back path (measurements): copies the last frame's RGBA8 rows -- this rank's bands of a
 * sharded frame, at their place in the whole frame -- to dst `repeats` times back to back, as rc_read_image_async copies
 * them, with no rendering in between; *ms = wall-clock time of the batch.  dst NULL: a page-locked frame of the call's own. */
int rc_debug_reread_image(void* dst, size_t dst_bytes, uint32_t repeats, float* ms);

/* sizeof of the ABI structs as the library was compiled: rc_world_desc, rc_world_info,
 * rc_frame_params, rc_hit, rc_frame_stats (bindings check their layouts against it). */
int rc_debug_struct_sizes(uint32_t out[5]);

#ifdef __cplusplus
}
#endif
#endif /* RC_DEBUG_H */
