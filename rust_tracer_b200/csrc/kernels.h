// kernels.h -- launch wrappers shared by the translation units of librt_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "scene_dev.cuh"

namespace rt {

// A run of n rays: a[i * stride] = (ox, oy, oz, dx), b[i * stride] = (dy, dz, -, -);
// key[i * stride] receives the nearest hit (exact.cuh: make_key) or KEY_MISS.  stride > 1
// picks the rays of one light out of a [node][light] pool (the ordered wavefront lays shadow rays
// out [light][node] and needs none); only the first segment of a sweep may be strided.
struct RaySeg {
    const float4 *a;
    const float4 *b;
    unsigned long long *key;
    int n;
    int stride = 1;
    int preset = 0; // the caller has already set every key to KEY_MISS (required when stride > 1)
    int heavy = 0;  // the first `heavy` rays are expensive (origins far outside the scene: large filter
                    // margins); their batches are cut into more table slices so that no work item runs long
};

// The bound table seen from one point O' (a light, the camera): slot j holds
// (C_j - O', W'_j = |C_j - O'|^2 - R_j^2 rounded down), same slot order, pairing and padding
// as DevScene::bound (and DevScene::group for the group spheres).  Every ray of a launch that uses it lies on a line through O' (up to
// rounding), so the filter needs the ray's direction only: 4 FMA per pair instead of 7.
struct OriginTable {
    const float4 *bound = nullptr; // null: no table, use the general filter
    const float4 *group = nullptr; // the group spheres (DevScene::group) re-centred likewise
    const float4 *quad = nullptr;  // and the quad spheres (DevScene::quad)
    const float4 *super = nullptr; // and the super spheres (DevScene::super)
    float reach = 0.f;             // |O'| + cmax, rounded up
    float ox = 0.f, oy = 0.f, oz = 0.f; // O'
    // > 0: the rays are shadow rays travelling TOWARDS O' (a light); a bound whose centre lies
    // more than |O' - ray origin| + rmax from O' along the ray cannot hold a hit between the ray
    // origin and the light and is skipped (rmax = largest bound radius, rounded up).  0: no clip.
    float rmax = 0.f;
    int shadow = 0; // the rays are the shadow rays of the light at O': any occluding hit ends a ray (sweep.cu: SHADOW)
};

// Nearest-hit query for every ray of s0 then s1 (Scene::intersect, scene/mod.rs:98-116).
// batch_counters (device,
// n_counters zeroed words) hand ray batches to the persistent warps, one word per launch;
// *used receives the number consumed (one per shared-memory chunk of the bound table).
// exact_counter (device, may be null; N_SWEEP_COUNTERS words) accumulates the number of exact pair
// tests [0], [7], [8] and of filter evaluations [1..6] (sweep.cu: sweep_kernel).
// origin.bound != null: every ray passes through origin's point; s1 must be empty.
cudaError_t launch_sweep(const DevScene &sc, RaySeg s0, RaySeg s1, OriginTable origin, int sm_count,
                         unsigned int *batch_counters, int n_counters, int *used, unsigned long long *exact_counter,
                         cudaStream_t stream);
// Several sweeps against origin tables in ONE launch (the shadow rays of every light of a
// wavefront level): job j = rays segs[j] (keys preset to KEY_MISS) against tabs[j]; CTA b serves job
// b % n_jobs.  Consumes n_jobs batch counters.  cudaErrorNotSupported (nothing launched) when the
// case does not fit -- fewer than 2 or more than 8 jobs, a table too large to stay resident --
// and the caller falls back to one launch_sweep per job.
cudaError_t launch_sweep_jobs(const DevScene &sc, const RaySeg *segs, const OriginTable *tabs, int n_jobs,
                              int sm_count, unsigned int *batch_counters, int n_counters, int *used,
                              unsigned long long *exact_counter, cudaStream_t stream);
// Launches one sweep of this scene's table takes (shared-memory chunks of the bound table; 1 when it is resident).
int sweep_table_chunks(const DevScene &sc);
// FMA per (ray, bound) pair of the two filters (for the executed-flop count)
constexpr int FILTER_FMA_GENERAL = 7, FILTER_FMA_ORIGIN = 4;
// Same result, every (ray, shape) pair through the exact test in shape order (audit).
void launch_sweep_all_pairs(const DevScene &sc, RaySeg s0, RaySeg s1, unsigned long long *exact_counter,
                            cudaStream_t stream);

// ---- wavefront state -----------------------------------------------------------------
// Tree rays (primary / reflect / refract) and shadow rays live in two pools; level k of
// the ray tree occupies a contiguous range of each.
struct Pools {
    // tree rays
    float4 *t_a, *t_b;            // ray
    unsigned long long *t_key;    // nearest hit
    int *t_parent;                // level 0: sample index; else (parent tree ray << 1) | slot (0 reflect, 1 refract)
    int *t_sbase;                 // first shadow ray of this node, -1 if the ray missed
    float4 *t_crefl, *t_crefr;    // colour returned by the reflect / refract child (w = 1 if the child is a node)
    int *t_root;                  // sample index of the tree this ray belongs to (forest mode)
    int *t_hitid;                 // Intersection::id of the node (cube: inner triangle), -1 on a miss; may be null
    int t_cap;
    // node order (may be null: tree-ray order): the tree rays of the level being spawned, hits
    // first in Morton order of their hit points, then the misses; bins[] is its histogram
    int *t_order;
    unsigned int *bins;           // node_bins(NODE_BIN_BITS_MAX) + 2 words (16-byte aligned)
    unsigned int *bin_sums;       // NODE_BIN_CHUNKS_MAX words: totals / prefixes of the bins' scan chunks (wavefront.cu)
    int node_bits;                // bits per axis of the node cells inside the shapes' box (tuning.h: node_bits)
    // shadow rays
    float4 *s_a, *s_b;            // per level: [light][node] (ordered spawn) or [node][light]
    unsigned long long *s_key;
    int s_cap;
    // device counters: [0] tree rays allocated, [1] shadow rays allocated, [2] nodes, [3] overflow flag,
    // [4] nodes of the level being spawned whose hit point lies outside the node grid's box (they
    // come first in t_order), [5] child rays those nodes spawned
    unsigned int *counters;
};

struct FrameParams {
    float cam_ox, cam_oy, cam_oz;
    float x_min, x_max, y_min, y_max;
    int x_res, y_res;
    int depth, spp;
    unsigned long long seed;
    const int *rows; // device: image rows rendered by this call
    int n_rows;
    int forest_shading;
    int primary_tiles; // primary rays are laid out in tiles of 16 x 8 pixels rather than row by row (wavefront.cu: raygen_kernel)
};

constexpr int N_COUNTERS = 6;
constexpr int N_SWEEP_COUNTERS = 5 + MAX_SEGS; // launch_sweep's exact_counter: exact tests, general evals, origin evals, slot evals per segment, exact sphere / triangle tests
constexpr int NODE_BIN_BITS_MAX = 7, NODE_FAR_BITS = 4; // per axis: hit points inside (Pools::node_bits <= MAX) / outside the shapes' box
constexpr int NODE_FAR_BINS = 1 << (3 * NODE_FAR_BITS);
// bins of the node order for `bits` bits per axis; bin node_bins(bits) itself counts the rays that missed
__host__ __device__ constexpr int node_bins(int bits) { return NODE_FAR_BINS + (1 << (3 * bits)); }
constexpr int NODE_BIN_CHUNKS_MAX = 1024; // scan chunks of 4096 bins: covers node_bins(NODE_BIN_BITS_MAX) + 1

// Primary rays for samples [s0, s0 + n) of this call into tree rays [0, n).
void launch_raygen(const FrameParams &fp, const Pools &p, long long sample0, int n, cudaStream_t stream);
// Level k: for every tree ray in [begin, end) that hit, allocate its shadow rays and (if
// spawn_children) its reflect / refract children.
// ordered: first sort the level's nodes by the cell of their hit point (p.t_order, p.bins), so
// that consecutive nodes -- and with them consecutive shadow rays of each light and consecutive
// child rays -- are neighbours in space; the level's shadow rays are then laid out light by
// light: node number r owns ray s_begin + l hits + r of light l, and counters[1] = s_begin +
// hits L.  Not ordered: [node][light] blocks are handed out by atomicAdd on counters[1] (which
// must hold s_begin).
void launch_spawn(const DevScene &sc, const FrameParams &fp, const Pools &p, int begin, int end, int spawn_children,
                  int ordered, int s_begin, cudaStream_t stream);
// Level k, deepest first: colour of every node in [begin, end) -> parent slot, or
// sample_rgb[sample - sample0] at level 0.
// light_stride: the shadow ray of point light l of a node is s_key[t_sbase + l * light_stride] -- the
// level's node count where spawn ran ordered ([light][node] layout), 1 otherwise ([node][light]).
void launch_shade(const DevScene &sc, const FrameParams &fp, const Pools &p, int begin, int end, int level, int light_stride,
                  float4 *sample_rgb, const unsigned char *sample_mask, cudaStream_t stream);
// Average spp samples per pixel (ordered sum * 0.25), write f32 and/or quantised u8.  Pixel
// pixel0 + i of the rendered rows goes to position pixel0 + i of the output, or, with frame_rows
// (device: the image row of every rendered row), to its own place in a whole width-wide frame --
// which may be another GPU's memory: the stores then are the multi-GPU gather.
void launch_resolve(const float4 *sample_rgb, long long pixel0, int n_pixels, int spp, float *rgb_f32,
                    unsigned char *rgb_u8, const unsigned char *pixel_mask, const int *frame_rows, int width,
                    cudaStream_t stream);

// render_forest_filter's write-back: the pixels with pixel_mask set, compacted as out_idx[k] = pixel number and
// its colour at 3 k of rgb_f32 / rgb_u8 (either may be null), in no particular order; *count = how many.
void launch_resolve_compact(const float4 *sample_rgb, int n_pixels, const unsigned char *pixel_mask, int *out_idx,
                            unsigned int *count, float *rgb_f32, unsigned char *rgb_u8, cudaStream_t stream);

// Forest bookkeeping over tree rays [0, n): sizes[root] += 1 per node (RayTree::size,
// render_tree.rs:39-48); mask[root] = 1 if a node's Intersection::id is in ids[]
// (RayTree::shapes, render_tree.rs:26,182).
void launch_forest_sizes(const Pools &p, int n, int *sizes, cudaStream_t stream);
void launch_forest_mark(const Pools &p, int n, const int *ids, int n_ids, unsigned char *mask, cudaStream_t stream);

// FMA-saturation micro-kernel; returns flops executed.
double launch_fp32_peak(int blocks, int iters, float *sink, cudaStream_t stream);

} // namespace rt
