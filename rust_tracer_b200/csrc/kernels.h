// kernels.h -- launch wrappers shared by the translation units of librt_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "scene_dev.cuh"

namespace rt {

// A contiguous run of rays: a[i] = (ox, oy, oz, dx), b[i] = (dy, dz, -, -); key[i] receives
// the nearest hit (exact.cuh: make_key) or KEY_MISS.
struct RaySeg {
    const float4 *a;
    const float4 *b;
    unsigned long long *key;
    int n;
};

// Nearest-hit query for every ray of s0 then s1 (Scene::intersect, scene/mod.rs:98-116).
// rays_per_thread: 1, 2 or 4, or 0 to choose by the number of rays.  batch_counters (device,
// n_counters zeroed words) hand ray batches to the persistent warps, one word per launch;
// *used receives the number consumed (one per shared-memory chunk of the bound table).
// exact_counter (device, may be null) accumulates the number of exact pair tests.
cudaError_t launch_sweep(const DevScene &sc, RaySeg s0, RaySeg s1, int rays_per_thread, int sm_count,
                         unsigned int *batch_counters, int n_counters, int *used, unsigned long long *exact_counter,
                         cudaStream_t stream);
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
    // shadow rays
    float4 *s_a, *s_b;
    unsigned long long *s_key;
    int s_cap;
    // device counters: [0] tree rays allocated, [1] shadow rays allocated, [2] nodes, [3] overflow flag
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
};

// Primary rays for samples [s0, s0 + n) of this call into tree rays [0, n).
void launch_raygen(const FrameParams &fp, const Pools &p, long long sample0, int n, cudaStream_t stream);
// Level k: for every tree ray in [begin, end) that hit, allocate its shadow rays and (if
// spawn_children) its reflect / refract children.
void launch_spawn(const DevScene &sc, const FrameParams &fp, const Pools &p, int begin, int end, int spawn_children,
                  cudaStream_t stream);
// Level k, deepest first: colour of every node in [begin, end) -> parent slot, or
// sample_rgb[sample - sample0] at level 0.
void launch_shade(const DevScene &sc, const FrameParams &fp, const Pools &p, int begin, int end, int level,
                  float4 *sample_rgb, const unsigned char *sample_mask, cudaStream_t stream);
// Average spp samples per pixel (ordered sum * 0.25), write f32 and/or quantised u8.
void launch_resolve(const float4 *sample_rgb, long long pixel0, int n_pixels, int spp, float *rgb_f32,
                    unsigned char *rgb_u8, const unsigned char *pixel_mask, cudaStream_t stream);

// Forest bookkeeping over tree rays [0, n): sizes[root] += 1 per node (RayTree::size,
// render_tree.rs:39-48); mask[root] = 1 if a node's Intersection::id is in ids[]
// (RayTree::shapes, render_tree.rs:26,182).
void launch_forest_sizes(const Pools &p, int n, int *sizes, cudaStream_t stream);
void launch_forest_mark(const Pools &p, int n, const int *ids, int n_ids, unsigned char *mask, cudaStream_t stream);

// FMA-saturation micro-kernel; returns flops executed.
double launch_fp32_peak(int blocks, int iters, float *sink, cudaStream_t stream);

} // namespace rt
