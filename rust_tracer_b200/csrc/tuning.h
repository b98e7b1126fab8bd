// tuning.h -- scheduling constants of the library, in one place.
//
// The library never reads the environment: a drop-in must not change its behaviour because of
// what the caller's process happens to export.  Every value below is a compile-time default;
// tests and measurement tools can override the scheduling ones through rt_debug_set
// (include/rt_b200_debug.h), which is not part of the drop-in ABI.  None of them can change a
// result: correctness margins (the bound radii and filter margins of api_scene.cu / sweep.cu)
// are constants and are NOT here.
#pragma once

namespace rt {

struct Tuning {
    int rays_per_thread = 0;    // sweep: rays per thread (1, 2, 4); 0 = by the number of rays
    int small_tree_rays = 1 << 30; // sweep: general-filter sweeps of fewer rays than this run ...
    int small_shadow_rays = 0;     // sweep: a level's shadow launch with fewer rays per light than this runs 2 rays per thread (0: never)
    int tree_rays_per_thread = 2;  // ... this many rays per thread instead of 4
    int want_light = 4;         // sweep: work items per warp aimed at for ordinary batches ...
    int want_heavy = 4;         // ... and for the expensive first batches of a launch (rays that start far outside the scene)
    int want_heavy_jobs = 4;    // ... the same for the shadow rays of a level (one launch for all lights)
    int tail_min = 0;           // sweep: the tail of a launch is cut finer only if it has >= tail_min batches per warp
    int slices = 0;             // sweep: force this many table slices per batch; 0 = plan
    int tail_pct = 15;          // sweep: the last tail_pct % of a launch's batches ...
    int tail_mult = 4;          // ... are cut tail_mult times finer
    int item_max = 16;          // sweep: lanes with a touching ray up to which a group is swept pair by pair
    int table_cap = 0;          // sweep: bound slots per shared-memory chunk; 0 = what fits (tests force chunks)
    int merge_lights = 1;       // wavefront: all lights' shadow rays of a level in one launch
    int primary_tiles = 2;      // wavefront: primary rays in tiles of 16 x 8 pixels (a warp's 128 rays form a narrow pyramid, not a one-pixel-high
                                // fan); 2: at 4 samples per pixel the tiles are 8 x 4 pixels (again one batch); 1: 16 x 8 always; 0: row by row
    int level_pipeline = 1;     // wavefront: a level's shadow sweep is not joined before the next level starts (only before shading)
    int job_rotate = 1;         // sweep: in that launch, a CTA whose light has run out of work items goes on with another light's
    int heavy_allow = 65536;    // wavefront: slack on the count of expensive child rays
    int node_bits = 0;          // wavefront: bits per axis of the node order's cells (3..7); 0 = by the size of the level
    int band_samples = 1 << 23; // render: samples per band
    int origin_tables = 1;      // scene: build per-light / camera origin tables
    int origin_min_slots = 256; // scene: ... only for tables of at least this many slots
    int host_threads = 4;       // scene: host threads of the table build
    int scene_cache = 1;        // scene: keep the tables of the last description built (api_scene.cu: host_scene_for)
    int shadow_anyhit = 1;      // sweep: a shadow ray stops at the first hit that occludes its light
    int trace = 0;              // 1: print per-phase host laps and per-sweep counters to stderr (sweeps one by one);
                                // 2: the same per level, the level's sweeps overlapped as in a normal call
};

Tuning &tuning();

} // namespace rt
