// scene_dev.cuh -- device-resident scene tables (structure of arrays, 16-byte records).
//
// Layout in HBM (all immutable during a render except `materials`, which
// rt_scene_update_materials may replace between calls):
//   bound[2p], bound[2p+1]  float4 x 2        conservative bounding spheres of bound slots 2p and
//                                             2p+1 (sphere / triangle / cube), interleaved for
//                                             packed FFMA2:  (cx0, cx1, cy0, cy1) (cz0, cz1, W0, W1)
//                                             with W = |C|^2 - R^2 rounded down.  Slots are grouped
//                                             into segments (SweepSegs) by error class, each padded
//                                             to SEG_PAD slots with W = FLT_MAX (passes no ray).
//                                             The ONLY table the sweep's inner loop reads (whole
//                                             table resident in shared memory).
//   group[2q], group[2q+1]  float4 x 2        bounding spheres of slot groups 2q and 2q+1, same
//                                             packing: group g encloses the bounds of slots
//                                             [GROUP_SLOTS g, GROUP_SLOTS (g + 1)), which are
//                                             neighbours in space (Morton order within a segment).
//                                             A warp skips a group none of its rays can touch.
//   quad[2q], quad[2q+1]    float4 x 2        the same one level up: spheres around QUAD_SLOTS slots
//                                             (4 consecutive groups).
//   super[2q], super[2q+1]  float4 x 2        and one more: spheres around SUPER_SLOTS slots (4 consecutive
//                                             quads), tested first.
//   bound_shape[j] int32                      shape id of bound slot j, -1 for a pad
//   slot_meta[j]   int2  (kind, shape id)     and
//   slot_geom[3*j..] 3 x float4               copies of shape_meta / geom in SLOT order: the exact
//                                             stage of the sweep reads them straight from the queued
//                                             slot number, in one round trip instead of three
//   shape_meta[s]  int4  (kind, material, plane index, 0)
//   geom[3*s..]    3 x float4                 sphere/cube/plane: rows 0..2 of inv_transform
//                                             triangle: v0.xyz v1.x | v1.yz v2.xy | v2.z n.xyz
//   plane_data[7*p..] 7 x float4              origin, normal, u axis, v axis, then rows 0..2 of the
//                                             forward transform (xyz used, w unused)
//   cube_tris[36]  float4                     the 12 unit-cube triangles, same packing as geom
//   materials[m]   DevMaterial (64 B)
//   lights[l]      DevLight (32 B)
#pragma once
#include <cstdio>

#include "exact.cuh"

// Checked build (tools/build_variant.sh checked "-DRT_DEBUG_CHECKS=1"): every index the kernels compute into
// a queue, a table, a pool or a key array is tested against its bounds; a violation prints its place and
// traps, which fails the launch.  compute-sanitizer is closed on this pool (profiles/
// r02_compute_sanitizer_closed.txt), so this is how the GPU tests look for bad accesses.
#ifndef RT_DEBUG_CHECKS
#define RT_DEBUG_CHECKS 0
#endif
#if RT_DEBUG_CHECKS
#define RT_CHECK(cond)                                                                  \
    do {                                                                                \
        if (!(cond)) {                                                                  \
            printf("RT_CHECK failed: %s  (%s:%d, block %d thread %d)\n", #cond, __FILE__, __LINE__, (int)blockIdx.x, (int)threadIdx.x); \
            __trap();                                                                   \
        }                                                                               \
    } while (0)
#else
#define RT_CHECK(cond) ((void)0)
#endif

namespace rt {

enum : int { KIND_SPHERE = 0, KIND_PLANE = 1, KIND_TRIANGLE = 2, KIND_CUBE = 3 };

struct DevColorFn {
    int kind; // 0 const, 1 checkerboard
    float r, g, b;
};

struct DevMaterial { // 64 bytes
    DevColorFn ambient, diffuse, specular;
    float power, reflectivity, refraction_index;
    int kind;
};

struct DevLight { // 32 bytes
    float px, py, pz;
    int kind; // 0 point, 1 ambient
    float cr, cg, cb;
    int point_index; // index among the point lights (shadow slot), -1 for ambient lights
};

constexpr int GROUP_SLOTS = 16; // bound slots per group
constexpr int QUAD_SLOTS = 64;  // bound slots per quad = 4 consecutive groups
constexpr int SUPER_SLOTS = 256; // bound slots per super = 4 consecutive quads
constexpr int SEG_PAD = 512;    // every segment of the bound table is a multiple of this many slots (a pair of supers)
constexpr int MAX_SEGS = 4;

// Bound slots [begin, end) share one margin factor k: a ray passes slot j iff
//   b^2 - |L|^2 + |C|^2 + k * 2^-24 * (|O| + cmax)^2  >=  W_j        (sweep.cu)
// With an origin table (kernels.h: OriginTable; A = |O| + cmax, B = |O'| + cmax):
//   (C'.dh)^2 + 2^-24 * (kx A^2 + 26 B^2 + 8 A B)  >=  W'_j
struct SweepSegs {
    int n;
    int begin[MAX_SEGS], end[MAX_SEGS];
    float k[MAX_SEGS];
    float kx[MAX_SEGS]; // the reference's exact-test share of k: 1.5 x (25 + 15 aniso^2)
};
// Segments hold one kind of shape each (the sweeps count their filter evaluations per segment); what the
// filter needs is coarser: consecutive segments with the same margin factors form one RUN, and a ray's
// filter coefficients are built once per run (m10k: spheres and triangles share k = 108: one run).

struct DevScene {
    int n_shapes;
    int n_bounded;  // bounded shapes (without pads)
    int n_slots;    // bound slots including pads (multiple of SEG_PAD)
    float cmax;     // max over bounds of |C| + R, rounded up
    float slot_rmax;     // largest radius of a bound slot, rounded up (3e38: one passes every ray)
    float level_rmax[3]; // largest radius of a group / quad / super sphere, rounded up (3e38: one of them passes every ray): sweep.cu, distance clip
    SweepSegs segs;
    SweepSegs runs; // segs merged where k and kx agree
    int n_planes;
    int n_lights;
    int n_point_lights;
    int first_nan_shape; // first sphere/triangle/cube in id order, -1 if none (NaN rays)
    float amb_r, amb_g, amb_b;
    float bin_lo[3], bin_scale[3]; // (hit point - bin_lo) * bin_scale - 1 is in [-1, 1] inside the box of the bounded shapes (wavefront.cu: node cells)
    const float4 *bound;
    const float4 *group;    // n_slots / GROUP_SLOTS group spheres, paired like bound
    const float4 *quad;     // n_slots / QUAD_SLOTS spheres around 4 consecutive groups each, paired like bound
    const float4 *super;    // n_slots / SUPER_SLOTS spheres around 4 consecutive quads each, paired like bound
    const int *bound_shape;
    const int2 *slot_meta;
    const float4 *slot_geom;
    const int4 *shape_meta; // (kind, material, plane index or -1, 0)
    const float4 *geom;
    const int *plane_shape;
    const float4 *plane_data;
    const float4 *cube_tris;
    const DevMaterial *materials;
    const DevLight *lights;
};

} // namespace rt
