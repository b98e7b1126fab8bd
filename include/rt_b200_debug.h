/*
 * rt_b200_debug.h -- test and measurement hooks of librt_b200.so.  NOT part of the drop-in ABI
 * (include/rt_b200.h): a host binding has no reason to call these, and the Rust `-sys` crate
 * does not declare them.
 *
 * The library never reads the environment.  Its scheduling constants (work items per warp,
 * rays per thread, table chunking, ...; rust_tracer_b200/csrc/tuning.h) are compile-time
 * defaults that the GPU tests override through rt_debug_set to drive every code path of the
 * sweep (several table chunks, pair path on / off, one launch per light, ...) and the tools
 * under tools/ to A/B them.  No knob can change a result: the correctness margins of the bound
 * filter are constants of the library and cannot be set from here.
 */
#ifndef RT_B200_DEBUG_H
#define RT_B200_DEBUG_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* name: a field of rt::Tuning ("rays_per_thread", "small_tree_rays", "tree_rays_per_thread", "small_shadow_rays", "want_light", "want_heavy", "want_heavy_jobs", "tail_min", "slices",
 * "tail_pct", "tail_mult", "item_max", "table_cap", "merge_lights", "job_rotate", "level_pipeline", "primary_tiles", "heavy_allow", "node_bits",
 * "band_samples", "origin_tables", "origin_min_slots", "host_threads", "scene_cache", "shadow_anyhit",
 * "trace").  Process-wide; takes effect at the next call (scene knobs: at the next
 * rt_scene_create).  RT_ERR_INVALID for an unknown name or a value out of range.          */
int rt_debug_set(const char *name, int64_t value);
/* Current value, or INT64_MIN for an unknown name. */
int64_t rt_debug_get(const char *name);
/* Back to the compile-time defaults. */
void rt_debug_reset(void);

/* Host-side logic of the sweep's scheduling, callable without a GPU (tests/test_sweep_plan_cpu.py): the plan of work items for
 * `total` rays (`heavy` of them expensive) against a resident table of `per` slots on `warps` persistent warps --
 * out = {items, heavy_batches, heavy_slices, heavy_slice_slots, reg_batches, n_slices, slice_slots, tail_slices,
 * tail_slice_slots, merge} -- and the (batch, slice, slots per slice) an item number decodes to in the kernel
 * (returns 1, or 0 past the plan).  -1: bad argument.                                                              */
int rt_debug_plan_items(int32_t rays_per_thread, int64_t total, int32_t heavy, int32_t per, int64_t warps, int32_t n_chunks,
                        int32_t jobs, int64_t out[10]);
/* The bound hierarchy the library derives from a scene description, as built on the host (no GPU): level 0 bound slots,
 * 1 groups of 16, 2 quads of 64, 3 supers of 256; five doubles per sphere (cx, cy, cz, radius incl. slack, kind: 0 sphere,
 * 1 pad, 2 passes every ray).  Returns the number of spheres of the level (at most `cap` are written) or a negative
 * rt_status.  rt_debug_scene_scalars: out = {n_slots, cmax, slot_rmax, level_rmax of groups, quads, supers}.
 * (rt_scene_desc: include/rt_b200.h, to be included first.)                                                        */
struct rt_scene_desc;
int rt_debug_scene_spheres(const struct rt_scene_desc *desc, int32_t level, double *out, int32_t cap);
int rt_debug_scene_scalars(const struct rt_scene_desc *desc, double out[6]);
/* The origin table of point light `light` (in the order of the point lights of the description), unpacked: four doubles per
 * sphere (cx, cy, cz, W' = |C'|^2 - R'^2, all relative to the light) -- the slots, then groups, quads, supers --
 * and out_origin = {light x, y, z, reach, rmax}.  Returns the number of spheres or a negative rt_status.             */
int rt_debug_scene_light_table(const struct rt_scene_desc *desc, int32_t light, double *out, int32_t cap, double out_origin[5]);
int rt_debug_decode_item(uint32_t item, int64_t total, int32_t batch_rays, const int64_t plan[10], int64_t out[3]);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_DEBUG_H */
