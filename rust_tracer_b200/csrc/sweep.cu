// sweep.cu -- the hot kernel: nearest hit of every ray against every shape.
//
// Replaces Scene::intersect's linear scan (reference src/scene/mod.rs:98-116, called from
// src/render.rs:47 and, once per light, src/scene/mod.rs:193).  Every (ray, shape) pair is
// resolved, most of them wholesale:
//   phase 1  FILTER   one persistent CTA per SM keeps the bounding-sphere table of the shapes
//                     (slots), of every 16 slots (groups), every 64 (quads) and every 256 (supers) in
//                     shared memory; each warp fetches batches of 32 R rays (R = 4; 2 for reflected /
//                     refracted rays), each thread keeps the filter coefficients of R rays in registers
//                     and tests them against two spheres at a time with packed FFMA2 (fma.rn.f32x2):
//                     supers first, then the quads of a super some ray of the warp may touch, its
//                     groups, then the slots of such a group -- all 32 R rays against each slot pair
//                     (brute loop), or, when few lanes hold a touching ray, one (ray, group) pair per
//                     half warp with lane = slot (pair path).  7 FMA per (ray, sphere) decide "this ray
//                     certainly misses everything in this sphere"; 4 FMA for rays that share a
//                     point (origin tables, below).  The tests are conservative: their margins
//                     cover both their own rounding and the reference's ("filter
//                     margins"), so they only ever reject pairs the exact test would reject too.
//                     Shadow rays leave the sweep at their first occluding hit (any-hit); nearest-hit
//                     rays skip the spheres of the hierarchy that lie wholly beyond their best hit so
//                     far (distance clip).
//   phase 2  EXACT    surviving (ray, slot) pairs go to a per-warp queue in shared memory;
//                     whenever 32 are waiting the warp runs 32 exact tests, one per lane,
//                     with the reference's own operand order (exact.cuh), and folds the
//                     result into the ray's 64-bit (t, id) key with an integer atomicMin.
//   phase 3  PLANES   infinite planes have no bound; every ray tests them exactly.
// The winner is therefore decided by the exact arithmetic alone, in every case.
//
// The filter, for a ray (O, d) and a sphere (C, R), with dh = d / |d|, L = C - O:
//     dist^2(C, line) <= R^2   <=>   (L.dh)^2 - |L|^2 + |C|^2  >=  |C|^2 - R^2 =: W
// and  (L.dh)^2 - |L|^2 + |C|^2 = b^2 + s  with  b = C.dh - O.dh,  s = C.(2 O) - |O|^2:
// two 3-FMA chains seeded per ray, one FMA to combine, one compare against the per-sphere W.
// The per-ray seed of s also carries the margin  k 2^-24 (|O| + cmax)^2, and b^2 is taken as
// b max(b, 0): a sphere centred behind the origin passes only if the origin is inside it.
//
// Rays that share a point O' -- primary rays (the camera) and the shadow rays of one light --
// are swept against the tables re-centred on O' (kernels.h: OriginTable): with C' = C - O',
//     dist^2(C, line) <= R^2   <=>   (C'.dh)^2  >=  |C'|^2 - R^2 =: W'
// one 3-FMA chain and one FMA that squares it and adds the per-ray margin: 4 FMA per pair.
//
// Scheduling (host side, plan_items / launch_sweep / launch_sweep_jobs): a launch's rays are cut
// into work items = (batch of 32 R rays) x (slice of the resident table) handed out by an atomic
// counter, slice by slice -- finer for the first, expensive batches and for the last 8 % of the launch -- and the
// shadow rays of all lights of a wavefront level go into ONE launch whose CTAs start on one light's rays and
// origin table each and move on to another light's when theirs run out (JOBS).
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cstdio>
#include <type_traits>

#include "kernels.h"
#include "tuning.h"

namespace rt {

#ifndef RT_SWEEP_PRELOAD
#define RT_SWEEP_PRELOAD 1 // the prologue loads the lane's R rays before it works on any of them
#endif
#ifndef RT_SWEEP_FLUSH_PIPE
#define RT_SWEEP_FLUSH_PIPE 0 // 1: the exact stage fetches the next 32 pairs' operands while it tests the current 32 (A/B builds)
#endif
#ifndef RT_SWEEP_SUPERS
#define RT_SWEEP_SUPERS 1 // 0: the sweep starts at the quad spheres (A/B builds); 1: at the super spheres above them
#endif
#ifndef RT_SWEEP_BRUTE_UNROLL
#define RT_SWEEP_BRUTE_UNROLL 1 // measured: 1 beats 2 by 0.3 ms per c4 frame, 4 loses 5.6 ms (instruction cache)
#endif
#ifndef RT_SWEEP_FLUSH_AT
#define RT_SWEEP_FLUSH_AT 32 // queued pairs at which a group's end runs the exact stage
#endif
#ifndef RT_SWEEP_FLUSH_AT_SHADOW
#define RT_SWEEP_FLUSH_AT_SHADOW 32 // the same for shadow rays (sooner = rays leave the sweep sooner)
#endif
#ifndef RT_SWEEP_CLIP
#define RT_SWEEP_CLIP 1 // nearest-hit sweeps: group / quad / super spheres wholly beyond a ray's best hit so far are not descended into
#endif
#ifndef RT_SWEEP_NEAR_FIRST
#define RT_SWEEP_NEAR_FIRST 0 // 1: a work item visits the super pairs of a run nearest first (as seen from its first ray's origin). Measured: no gain on c4
                               // (35.99 vs 35.69 ms, profiles/r02_ab_near_first_c4.log) -- in a dense scene the first hit found already clips well
#endif
#ifndef RT_SWEEP_SLICE_MAJOR
#define RT_SWEEP_SLICE_MAJOR 1 // items of a sliced launch are handed out slice by slice (all batches against slice 0, then slice 1, ...) and an item
                               // starts from what earlier slices have left in the rays' keys: a shadow ray found occluded does not enter, a tree
                               // ray starts with its clip.  0: batch by batch, every slice from scratch
#endif
#ifndef RT_SWEEP_SKIP_OCCLUDED
#define RT_SWEEP_SKIP_OCCLUDED 0 // 1: the exact stage drops queued pairs of shadow rays already found occluded.  Measured: 3 % fewer exact tests,
                                  // 0.1 ms slower per c4 frame (the check sits on the flush's critical path): profiles/r02_ab_skip_occluded_pairs_c4.log
#endif
#ifndef RT_SWEEP_PRIMARY_TF
#define RT_SWEEP_PRIMARY_TF 0 // 1: primary rays also clip bound slots (not only the hierarchy) beyond their best hit (A/B builds)
#endif
#ifndef RT_SWEEP_R8
#define RT_SWEEP_R8 0 // 1: also build 8 rays per thread x 256 threads (rays_per_thread=8; A/B builds)
#endif
#ifndef RT_SWEEP_BULK_TABLE
#define RT_SWEEP_BULK_TABLE 0 // how the table reaches shared memory, measured on c4 (profiles/r02_ab_table_staging_heavy_c4.log):
                              // 0 = copied by the CTA's 512 threads behind a block barrier                       38.08 ms
                              // 1 = cp.async.bulk + mbarrier, one copy per table issued by one thread, each
                              //     warp waits at its first use                                                   38.55 ms
                              // 2 = cp.async.bulk in 96 chunks issued by the lanes of warp 0, wait before the
                              //     first item                                                                    53.35 ms
#endif
#ifndef RT_SWEEP_QCAP
#define RT_SWEEP_QCAP 256
#endif
#ifndef RT_SWEEP_T4
#define RT_SWEEP_T4 512 // threads per CTA of the 4-rays-per-thread kernels (A/B builds: 576 = 18 warps under 112 registers)
#endif
#ifndef RT_SWEEP_T2
#define RT_SWEEP_T2 512 // threads per CTA of the 2-rays-per-thread kernels (A/B builds: 1024 with RT_SWEEP_QCAP=128)
#endif
constexpr int QCAP = RT_SWEEP_QCAP;      // queued pairs per warp
constexpr int BRUTE_UNROLL = RT_SWEEP_BRUTE_UNROLL;
constexpr float U24 = 5.9604645e-08f;    // 2^-24, unit round-off of f32
constexpr float GROUP_MARGIN = 17.f;     // margin of a group sphere / margin of a bound (make_group: a / rho <= 16)
constexpr float QUAD_MARGIN = 3.f;       // margin of a quad sphere / margin of a group sphere (a / rho <= 2)
constexpr float SUPER_MARGIN = 3.f;      // margin of a super sphere / margin of a quad sphere (a / rho <= 2)
// seed_q = seed_g + QUAD_SEED (seed_g - seed): (QUAD_MARGIN GROUP_MARGIN - GROUP_MARGIN) / (GROUP_MARGIN - 1)
constexpr float QUAD_SEED = (QUAD_MARGIN * GROUP_MARGIN - GROUP_MARGIN) / (GROUP_MARGIN - 1.f) * 1.0001f;
constexpr float SUPER_SEED = (SUPER_MARGIN * QUAD_MARGIN * GROUP_MARGIN - GROUP_MARGIN) / (GROUP_MARGIN - 1.f) * 1.0001f; // likewise for a super
constexpr int MIN_SLICE_SLOTS = SEG_PAD; // smallest table slice of a work item
constexpr int SMEM_LIMIT = 232448;       // 227 KB opt-in dynamic shared memory per CTA on sm_100

struct RayRegs {
    float ox, oy, oz, dx, dy, dz;
};

RT_DEV bool ray_has_nan(const RayRegs &r) {
    return !(r.ox == r.ox && r.oy == r.oy && r.oz == r.oz && r.dx == r.dx && r.dy == r.dy && r.dz == r.dz);
}

RT_DEV RayRegs load_ray(const RaySeg &s0, const RaySeg &s1, int gi) {
    RT_CHECK(gi >= 0 && gi < s0.n + s1.n);
    float4 a, b;
    if (gi < s0.n) { a = s0.a[(long long)gi * s0.stride]; b = s0.b[(long long)gi * s0.stride]; }
    else { a = s1.a[gi - s0.n]; b = s1.b[gi - s0.n]; }
    RayRegs r;
    r.ox = a.x; r.oy = a.y; r.oz = a.z; r.dx = a.w; r.dy = b.x; r.dz = b.y;
    return r;
}

// One exact (ray, shape) test; returns the key or KEY_MISS.
RT_DEV unsigned long long exact_pair(const int4 *shape_meta, const float4 *geom, const float4 *cube_tris, int shape, V3 o,
                                     V3 d) {
    int kind = shape_meta[shape].x;
    const float4 *g = geom + 3 * shape;
    float t;
    if (kind == KIND_SPHERE) {
        bool entering;
        if (!exact_sphere(load_m34(g), o, d, t, entering)) return KEY_MISS;
    } else if (kind == KIND_TRIANGLE) {
        float u, v, det;
        if (!exact_triangle(load_tri(g), o, d, t, u, v, det)) return KEY_MISS;
    } else { // cube
        float u, v, det;
        int which;
        if (!exact_cube(load_m34(g), cube_tris, o, d, t, which, u, v, det)) return KEY_MISS;
    }
    return make_key(t, shape);
}

// A cube's 12 triangle tests, out of line: cubes are rare, and the exact stage's code is fetched by warps
// that come from all over the filter loops (its instruction-cache misses were 45 % of its stall samples)
__device__ __noinline__ bool exact_cube_t(const float4 *g, const float4 *cube_tris, V3 o, V3 d, float *t_out) {
    float t, u, v, det;
    int which;
    if (!exact_cube(load_m34(g), cube_tris, o, d, t, which, u, v, det)) return false;
    *t_out = t;
    return true;
}

// The same from a slot's own copy of (kind, shape id, geometry).
RT_DEV unsigned long long exact_slot(int kind, int shape, const float4 *g, const float4 *cube_tris, V3 o, V3 d) {
    float t;
    if (kind == KIND_SPHERE) {
        bool entering;
        if (!exact_sphere(load_m34(g), o, d, t, entering)) return KEY_MISS;
    } else if (kind == KIND_TRIANGLE) {
        float u, v, det;
        if (!exact_triangle(load_tri(g), o, d, t, u, v, det)) return KEY_MISS;
    } else { // cube
        if (!exact_cube_t(g, cube_tris, o, d, &t)) return KEY_MISS;
    }
    return make_key(t, shape);
}

__device__ __noinline__ unsigned long long plane_keys(int n_planes, const int *plane_shape, const float4 *plane_data,
                                                      const float4 *geom, V3 o, V3 d, unsigned long long best) {
    for (int p = 0; p < n_planes; ++p) {
        int shape = plane_shape[p];
        const float4 *pd = plane_data + 7 * p;
        float4 po = pd[0], pn = pd[1];
        float t;
        if (exact_plane(load_m34(geom + 3 * shape), v3(po.x, po.y, po.z), v3(pn.x, pn.y, pn.z), o, d, t)) {
            unsigned long long k = make_key(t, shape);
            best = k < best ? k : best;
        }
    }
    return best;
}

// What the exact stage needs, kept in shared memory so that the filter loop holds none of
// it in registers (and the compiler cannot hoist the loads of these kernel parameters into it).
struct SweepCtx {
    RaySeg s0, s1;
    const int2 *slot_meta;
    const float4 *slot_geom;
    const float4 *cube_tris;
    float lx, ly, lz; // SHADOW: the light the rays of this CTA travel towards
    float cmax;       // DevScene::cmax
    int n_slots;      // (checked build)
};

// One ray set of a JOBS launch (sweep_kernel) with its origin table, work counter and item shapes.
struct SweepJob {
    RaySeg s0;
    OriginTable ot;
    unsigned int *batch_counter;
    int n_slices, slice_slots, merge;
    int heavy_batches, heavy_slices, heavy_slice_slots;
    int reg_batches, tail_slices, tail_slice_slots;
    unsigned int n_items; // work items of the plan (the counter runs past it as warps find the job dry)
};
constexpr int MAX_SWEEP_JOBS = 8;
constexpr unsigned int JOB_SWITCH_MIN = 64; // a CTA moves to another job only if that one has at least this many items left
struct SweepJobs {
    SweepJob j[MAX_SWEEP_JOBS];
    int n;
    int rotate; // a CTA whose job has run dry reloads the table of the job with the most items left and goes on there
};

// PointLight::get_energy (reference src/scene/mod.rs:195-204) calls the light occluded iff the
// nearest hit point is nearer to the ray's origin than the light:
//     (ray.at(t) - origin).len2() < (light - origin).len2()
// -- the same operations in the same order as shade_kernel (wavefront.cu), so the two agree bit
// for bit.  For t >= 0 the left side is a non-decreasing function of t even in f32 (a chain of
// monotone roundings), so once ANY hit with t >= 0 occludes, the nearest hit occludes too.
RT_DEV bool occludes(V3 o, V3 d, float t, float light_d2) { return vlen2(vsub(ray_at(o, d, t), o)) < light_d2; }

// sqrt.approx / rsqrt.approx (2 ulp): the filter's margins allow dh 4 u of error per component and every length
// here is rounded UP by a factor far above 2 ulp, so the IEEE sqrt and the three IEEE divisions (some 90
// instructions per ray, once per work item) are not needed
RT_DEV float sqrt_approx(float x) {
    float y;
    asm("sqrt.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
RT_DEV float rsqrt_approx(float x) {
    float y;
    asm("rsqrt.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// per-warp counters of the exact stage
struct ExactCount {
    unsigned int all, sphere, triangle;
};

// Run queued pairs 32 at a time, one exact test per lane; a partial batch stays queued unless
// `all`.  Called by a converged warp.  Returns the number of pairs left in the queue.
// SHADOW: wd2[local] = squared distance from ray `local` of the batch to its light; a hit that
// occludes the light sets the ray's bit in wocc[] (the caller then drops the ray from the sweep).
// The operands of one queued pair: its ray, and kind / shape id / geometry of its slot.
struct PairOperands {
    RayRegs rr;
    int2 sm;
    float4 g0, g1, g2;
    int local;
};
RT_DEV PairOperands fetch_pair(const SweepCtx *ctx, const unsigned int *wqueue, int idx, int ray0) {
    PairOperands p;
    const unsigned int pair = wqueue[idx];
    p.local = pair >> 24; // (ray of the warp batch << 24) | bound slot
    const int j = pair & 0xFFFFFF;
    RT_CHECK(p.local < 128 && j < ctx->n_slots && idx < QCAP);
    // ray, kind and geometry are all addressed by the queue entry: independent loads, one round trip
    p.sm = ctx->slot_meta[j];
    const float4 *gp = ctx->slot_geom + 3 * j;
    p.g0 = gp[0]; p.g1 = gp[1]; p.g2 = gp[2];
    p.rr = load_ray(ctx->s0, ctx->s1, ray0 + p.local);
    return p;
}

template <bool SHADOW>
__device__ __noinline__ int sweep_flush(const SweepCtx *ctx, unsigned int *wqueue, unsigned long long *wbest, int ray0,
                                        int cnt, int all, ExactCount *n_exact, float *wd2, unsigned int *wocc) {
    const int lane = threadIdx.x & 31;
    int done = 0;
    unsigned int n_sphere = 0, n_triangle = 0, n_run = 0;
#if RT_SWEEP_FLUSH_PIPE
    PairOperands next;
    bool have_next = (cnt >= 32 || (all && cnt > 0)) && lane < cnt;
    if (have_next) next = fetch_pair(ctx, wqueue, lane, ray0);
#endif
    while (cnt - done >= 32 || (all && done < cnt)) {
        int idx = done + lane;
        int kind = -1;
#if RT_SWEEP_FLUSH_PIPE
        const bool have = have_next;
        PairOperands cur;
        if (have) cur = next;
        { // the batch after this one, if it will run: its loads fly while this one is tested
            const int nd = done + 32;
            have_next = (cnt - nd >= 32 || (all && nd < cnt)) && nd + lane < cnt;
            if (have_next) next = fetch_pair(ctx, wqueue, nd + lane, ray0);
        }
        if (have) {
#else
        bool run = idx < cnt;
        if constexpr (SHADOW && RT_SWEEP_SKIP_OCCLUDED) { // a ray an earlier round of this flush (or an earlier flush) found occluded needs no more tests
            if (run) {
                const unsigned int local = wqueue[idx] >> 24;
                run = ((wocc[local >> 5] >> (local & 31)) & 1u) == 0u;
            }
        }
        if (run) {
            const PairOperands cur = fetch_pair(ctx, wqueue, idx, ray0);
#endif
            const float4 g[3] = {cur.g0, cur.g1, cur.g2};
            const int local = cur.local;
            kind = cur.sm.x;
            const V3 o = v3(cur.rr.ox, cur.rr.oy, cur.rr.oz), d = v3(cur.rr.dx, cur.rr.dy, cur.rr.dz);
            unsigned long long k = exact_slot(cur.sm.x, cur.sm.y, g, ctx->cube_tris, o, d);
            if (k != KEY_MISS) {
                atomicMin(&wbest[local], k);
                if constexpr (SHADOW) {
                    if (occludes(o, d, key_t(k), wd2[local])) atomicOr(&wocc[local >> 5], 1u << (local & 31));
                } else {
#if RT_SWEEP_CLIP
                    // distance clip (wd2 = the rays' clip distances here): nothing further along the ray than this hit can win.
                    // 1.001 and 1e-4 (|O| + cmax) are far above the rounding of t, of |d| and of the filters' b (<= 24 u lambda)
                    const float n2 = vlen2(d);
                    const float s_hit = key_t(k) * sqrt_approx(n2);
                    const float clip = fmaf(s_hit, 1.001f, 1e-4f * (sqrt_approx(vlen2(o)) + ctx->cmax));
                    if (n2 > 1e-30f && n2 < 1e30f && clip >= 0.f) // (|d|^2 a normal f32; false for NaN)
                        atomicMin((unsigned int *)&wd2[local], __float_as_uint(clip));
#endif
                }
            }
        }
        (void)idx;
        n_sphere += __popc(__ballot_sync(0xffffffffu, kind == KIND_SPHERE));
        n_triangle += __popc(__ballot_sync(0xffffffffu, kind == KIND_TRIANGLE));
        n_run += __popc(__ballot_sync(0xffffffffu, kind >= 0)); // pairs tested (queued pairs of occluded rays are dropped untested)
        done += 32;
    }
    if (lane == 0) {
        n_exact->all += n_run;
        n_exact->sphere += n_sphere;
        n_exact->triangle += n_triangle;
    }
    int left = cnt - done;
    unsigned int keep = 0;
    if (left > 0 && lane < left) keep = wqueue[done + lane];
    __syncwarp();
    if (left > 0 && lane < left) wqueue[lane] = keep;
    __syncwarp();
    return left > 0 ? left : 0;
}

// ---- table staging: one bulk copy per table, completion on an mbarrier ------------------------
RT_DEV uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
RT_DEV void mbar_init(unsigned long long *bar, unsigned int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
RT_DEV void mbar_expect_tx(unsigned long long *bar, unsigned int bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
// global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned (cp.async.bulk: the
// TMA unit moves the block; no thread holds the data in registers on the way)
RT_DEV void bulk_load(void *dst, const void *src, unsigned int bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}
RT_DEV void mbar_wait(unsigned long long *bar, unsigned int phase) {
    unsigned int ok = 0;
    while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(smem_addr(bar)), "r"(phase)
                     : "memory");
    }
}

// Filter coefficients of one ray (8 registers).  A ray the filter cannot bound (zero or
// non-finite direction length, coordinates beyond 1e15) passes every bound; a dead lane or a
// NaN ray (handled apart, see KEY_NAN_PREFIX) passes none.
// seed_g: the seed for a group sphere, its margin GROUP_MARGIN times larger (see make_group in
// api_scene.cu: a ray within a member's margin must also be within its group's).
struct RayCoef {
    float dx, dy, dz, kb, ox2, oy2, oz2, seed, seed_g;
};

RT_DEV RayCoef ray_coef(const RayRegs &r, float k_seg, float cmax, bool usable) {
    RayCoef c;
    c.dx = c.dy = c.dz = c.kb = c.ox2 = c.oy2 = c.oz2 = 0.f;
    c.seed = c.seed_g = -3.0e38f;
    if (!usable) return c;
    float n = sqrtf(fmaf(r.dx, r.dx, fmaf(r.dy, r.dy, r.dz * r.dz)));
    float o2 = fmaf(r.ox, r.ox, fmaf(r.oy, r.oy, r.oz * r.oz));
    float lam = (sqrtf(o2) + cmax) * 1.000001f;
    if (!(n > 0.f) || !(n < 1e18f) || !(lam < 1e15f)) {
        c.seed = c.seed_g = 3.0e38f;
        return c;
    }
    c.dx = __fdiv_rn(r.dx, n); c.dy = __fdiv_rn(r.dy, n); c.dz = __fdiv_rn(r.dz, n);
    // b = C.dh + kb is the distance from the origin to the centre's foot on the ray; kb carries
    // 12 u lam (1.5 x the rounding of b) so that the computed b is never below the true one
    c.kb = fmaf(12.f * U24, lam, -fmaf(r.ox, c.dx, fmaf(r.oy, c.dy, r.oz * c.dz)));
    c.ox2 = 2.f * r.ox; c.oy2 = 2.f * r.oy; c.oz2 = 2.f * r.oz;
    c.seed = fmaf(k_seg * U24, lam * lam, -o2);
    c.seed_g = fmaf(GROUP_MARGIN * k_seg * U24, lam * lam, -o2);
    return c;
}

// Same for a launch with an origin table: direction and margin only (4 registers).
// mu = 2^-24 (kx A^2 + 32 B^2 + 8 A B), A = |O| + cmax, B = |O'| + cmax: the reference's exact
// test, the filter's own rounding, and the distance of O' from the ray's line (a shadow ray's
// direction is the rounded, normalised  light - origin).  DESIGN.md, "filter margins".
// tf: far clip of a shadow ray in units of e = (C'.dh)^2 + mu -- a bound with e > tf lies wholly
// behind the ray's origin, or beyond the light by more than the origin-light distance; neither
// can shadow (PointLight::get_energy, scene/mod.rs:195-204, compares the hit with the light).
struct RayCoefO {
    float dx, dy, dz, mu, mu_g, tf;
};

RT_DEV RayCoefO ray_coef_origin(const RayRegs &r, float kx_seg, float cmax, const OriginTable &ot, bool usable) {
    const float reach = ot.reach;
    RayCoefO c;
    c.dx = c.dy = c.dz = 0.f;
    c.mu = c.mu_g = -3.0e38f;
    c.tf = 3.4e38f;
    if (!usable) return c;
    const float n2 = fmaf(r.dx, r.dx, fmaf(r.dy, r.dy, r.dz * r.dz));
    const float o2 = fmaf(r.ox, r.ox, fmaf(r.oy, r.oy, r.oz * r.oz));
    const float A = (sqrt_approx(o2) + cmax) * 1.000002f;
    // |d| in (1e-18, 1e18): its square is a normal f32 and rsqrt.approx is good to 2 ulp
    if (!(n2 > 1e-36f) || !(n2 < 1e36f) || !(A < 1e15f) || !(reach < 1e15f)) {
        c.mu = c.mu_g = 3.0e38f;
        return c;
    }
    const float inv = rsqrt_approx(n2); // dh = d / |d| to 5.3 u (2^-24) per component, which the margin below budgets
    c.dx = r.dx * inv; c.dy = r.dy * inv; c.dz = r.dz * inv;
    // 32 >= 1.5 x (2 x (5.3 [dh: n2 0.75 u, rsqrt.approx 2 ulp = 4 u, the product 0.5 u] + 3 [the FMA chain of b]) + 1 [the square] + 2)
    c.mu = U24 * 1.000001f * fmaf(kx_seg * A, A, fmaf(32.f * reach, reach, 8.f * A * reach));
    c.mu_g = GROUP_MARGIN * c.mu;
    if (ot.rmax > 0.f) {
        float lx = ot.ox - r.ox, ly = ot.oy - r.oy, lz = ot.oz - r.oz;
        float D = fmaf(1.0001f, sqrt_approx(fmaf(lx, lx, fmaf(ly, ly, lz * lz))) + ot.rmax, 1e-4f * reach);
        float t = fmaf(D, D, c.mu) * 1.0001f;
        if (t < 3.0e38f) c.tf = t; // NaN or overflow: no clip
    }
    return c;
}

RT_DEV float2 f2(float x) { return make_float2(x, x); }

// Work item number -> (ray batch, table slice, slots per slice).  Items come in three runs -- the batches of the first `heavy`
// rays (heavy_slices slices each), the bulk (reg_batches batches, n_slices each) and the last batches of the launch
// (tail_slices each) -- and within a run slice-major (item = slice * batches + batch; RT_SWEEP_SLICE_MAJOR) or batch-major.
// False past the last item of the plan.  Host and device: tests/test_sweep_plan_cpu.py walks every item of a plan through it.
__host__ __device__ inline bool decode_item(unsigned int item, long long total, int batch_rays, int heavy_batches, int heavy_slices,
                                            int heavy_slice_slots, int reg_batches, int n_slices, int slice_slots, int tail_slices,
                                            int tail_slice_slots, unsigned int &batch, int &slice, int &ss) {
    // items [0, heavy_batches * heavy_slices): the expensive first batches, cut finer
    const unsigned int heavy_items = (unsigned int)heavy_batches * (unsigned int)heavy_slices;
#if RT_SWEEP_SLICE_MAJOR
    // within each of the three runs of items: slice-major (item = slice * batches + batch)
    if (item < heavy_items) {
        slice = (int)(item / (unsigned int)heavy_batches);
        batch = item - (unsigned int)slice * (unsigned int)heavy_batches;
        ss = heavy_slice_slots;
    } else {
        const unsigned int li = item - heavy_items, reg_items = (unsigned int)reg_batches * (unsigned int)n_slices;
        if (li < reg_items) {
            slice = (int)(li / (unsigned int)reg_batches);
            batch = li - (unsigned int)slice * (unsigned int)reg_batches + (unsigned int)heavy_batches;
            ss = slice_slots;
        } else { // the last batches of the launch, cut finer: the launch ends on short items
            const unsigned int lt = li - reg_items;
            const long long tail_batches = (total + batch_rays - 1) / batch_rays - heavy_batches - reg_batches;
            if (tail_batches <= 0 || lt >= (unsigned long long)tail_batches * (unsigned int)tail_slices) return false; // past the plan
            slice = (int)(lt / (unsigned int)tail_batches);
            batch = lt - (unsigned int)slice * (unsigned int)tail_batches + (unsigned int)heavy_batches + (unsigned int)reg_batches;
            ss = tail_slice_slots;
        }
    }
#else
    if (item < heavy_items) {
        batch = item / (unsigned int)heavy_slices;
        slice = (int)(item - batch * (unsigned int)heavy_slices);
        ss = heavy_slice_slots;
    } else {
        const unsigned int li = item - heavy_items, reg_items = (unsigned int)reg_batches * (unsigned int)n_slices;
        if (li < reg_items) {
            batch = li / (unsigned int)n_slices;
            slice = (int)(li - batch * (unsigned int)n_slices);
            batch += (unsigned int)heavy_batches;
            ss = slice_slots;
        } else { // the last batches of the launch, cut finer: the launch ends on short items
            const unsigned int lt = li - reg_items;
            batch = lt / (unsigned int)tail_slices;
            slice = (int)(lt - batch * (unsigned int)tail_slices);
            batch += (unsigned int)heavy_batches + (unsigned int)reg_batches;
            ss = tail_slice_slots;
        }
    }
#endif
    return (long long)batch * batch_rays < total;
}

// One persistent CTA per SM; slots [slot_begin, slot_end) of the bound table stay in shared
// memory for the whole launch (copied by the CTA's threads behind one block barrier; RT_SWEEP_BULK_TABLE
// selects cp.async.bulk + mbarrier instead, measured slower).  A work item is (batch of
// 32 R rays) x (slice of the resident slots); each WARP fetches its own items from
// batch_counter, so no block barrier follows the table load, and small launches still spread
// over every warp of the GPU (n_slices > 1).
// Items come in three runs: the batches of the first s0.heavy rays (heavy_slices slices each), the
// bulk (reg_batches batches, n_slices each), and the last batches of the launch cut finer
// (tail_slices each) so that the warps run out of work at about the same time.
// `merge`: keys are folded into global memory with atomicMin (they were initialised to
// KEY_MISS or hold the result of earlier table chunks) instead of stored; `planes`: this
// launch also tests the planes (slice 0 does).
// ORIGIN: the table is `ot` (all rays pass through its point), filter of 4 FMA per pair.
//
// SHADOW (with ORIGIN): the rays are the shadow rays of the light at ot's point.  All
// PointLight::get_energy asks of their nearest hit is whether it is nearer than the light
// (scene/mod.rs:195-204), so a ray leaves the sweep at the first exact hit that says so (see
// occludes()): its filter coefficients are neutralised, later quads and groups are only visited
// for the rays still undecided, and a warp whose rays are all decided drops the rest of its
// item.  The planes are tested FIRST: a plane may return t < 0 (plane.rs:59-64 keeps it), which
// no other shape can undercut, so such a hit is the nearest one and decides alone.  The key
// stored for an occluded ray is the smallest hit found so far, not necessarily the nearest one
// of all: both occlude, and shading reads nothing else from a shadow key.
//
// JOBS: one launch sweeps several independent ray sets, each against its own origin table (the
// shadow rays of every light of one wavefront level): CTA b works on job b % jobs.n only -- its
// rays, table, work counter and item shapes come from jobs.j[] (copied to shared memory) instead
// of the parameters of the same name, which are unused.  A level's L small launches, each
// needing its table cut into slices to occupy every warp of the GPU, become one launch in which
// a light has 1/L of the CTAs and L times the batches per warp.
//
// exact_counter (N_SWEEP_COUNTERS words): [0] exact tests, [1] spheres evaluated by the general
// filter, [2] by the origin filter (bound, group and quad spheres alike), [3 + g] evaluations of
// bound slots of segment g, [7] exact tests of spheres, [8] of triangles.
template <int R, int SWEEP_THREADS, bool ORIGIN, bool JOBS, bool SHADOW>
__global__ void __launch_bounds__(SWEEP_THREADS, 1)
sweep_kernel(DevScene sc, RaySeg s0_, RaySeg s1, OriginTable ot_, int slot_begin, int slot_end, int n_slices_, int slice_slots_,
             int merge_, int planes, int item_max, int heavy_batches_, int heavy_slices_, int heavy_slice_slots_,
             int reg_batches_, int tail_slices_, int tail_slice_slots_,
             unsigned int *batch_counter_, unsigned long long *exact_counter, const __grid_constant__ SweepJobs jobs) {
    static_assert(!SHADOW || ORIGIN, "shadow rays are swept against their light's origin table");
    constexpr int SWEEP_WARPS = SWEEP_THREADS / 32;
    constexpr unsigned int FULL = 0xffffffffu;
    extern __shared__ float4 smem[];
    __shared__ SweepCtx ctx;
    __shared__ SweepJob job;
    __shared__ unsigned long long tab_bar; // mbarrier: the table's bulk copies have landed
    __shared__ ExactCount n_exact[SWEEP_WARPS];
    __shared__ unsigned long long n_hier[SWEEP_WARPS], n_slot[SWEEP_WARPS][MAX_SEGS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __shared__ int job_index; // JOBS: the job this CTA serves now, -1 when none has enough left
    // JOBS: the lights of a level do not cost the same.  A CTA starts on job blockIdx.x % n; when that job's items
    // are out it restages the table of the job with the most items left (~2 us for 210 KB out of L2) and goes on
    // there, so that the launch ends on all SMs at about the same time.  (Body not re-indented.)
    for (int round = 0;; ++round) {
    if constexpr (JOBS) {
        if (round) {
            if (!jobs.rotate || RT_SWEEP_BULK_TABLE != 0) break;
            __syncthreads(); // every warp has left the previous job's table and `job`
        }
        if (threadIdx.x == 0) {
            int pick = -1;
            if (round == 0) pick = (int)(blockIdx.x % (unsigned int)jobs.n);
            else {
                unsigned int most = JOB_SWITCH_MIN - 1;
                for (int j = 0; j < jobs.n; ++j) {
                    const unsigned int taken = *(volatile unsigned int *)jobs.j[j].batch_counter, n = jobs.j[j].n_items;
                    const unsigned int left = taken < n ? n - taken : 0u;
                    if (left > most) { most = left; pick = j; }
                }
            }
            job_index = pick;
            if (pick >= 0) job = jobs.j[pick];
        }
        __syncthreads();
        if (job_index < 0) break;
    } else {
        if (round) break;
    }
    const RaySeg &s0 = JOBS ? job.s0 : s0_;
    const OriginTable &ot = JOBS ? job.ot : ot_;
    const int n_slices = JOBS ? job.n_slices : n_slices_, slice_slots = JOBS ? job.slice_slots : slice_slots_;
    const int merge = JOBS ? job.merge : merge_;
    const int heavy_batches = JOBS ? job.heavy_batches : heavy_batches_, heavy_slices = JOBS ? job.heavy_slices : heavy_slices_;
    const int heavy_slice_slots = JOBS ? job.heavy_slice_slots : heavy_slice_slots_;
    const int reg_batches = JOBS ? job.reg_batches : reg_batches_, tail_slices = JOBS ? job.tail_slices : tail_slices_;
    const int tail_slice_slots = JOBS ? job.tail_slice_slots : tail_slice_slots_;
    unsigned int *const batch_counter = JOBS ? job.batch_counter : batch_counter_;
    const int n_tab = slot_end - slot_begin, n_gtab = n_tab / GROUP_SLOTS, n_qtab = n_tab / QUAD_SLOTS, n_stab = n_tab / SUPER_SLOTS;
    float4 *tab = smem;                                                         // n_tab slots + one slack pair
    float4 *gtab = tab + n_tab + 2;                                             // n_gtab groups
    float4 *qtab = gtab + n_gtab;                                               // n_qtab quads
    float4 *stab = qtab + n_qtab;                                               // n_stab supers
    unsigned long long *best = (unsigned long long *)(stab + n_stab);           // SWEEP_THREADS * R
    unsigned int *queue = (unsigned int *)(best + SWEEP_THREADS * R);           // SWEEP_WARPS * QCAP
    float *light_d2 = (float *)(queue + SWEEP_WARPS * QCAP);                    // SHADOW: SWEEP_THREADS * R
    unsigned int *occluded = (unsigned int *)(light_d2 + SWEEP_THREADS * R);    // SHADOW: SWEEP_WARPS * R bit masks

    const int total = s0.n + s1.n;
    unsigned long long *wbest = best + warp * (32 * R);
    unsigned int *wqueue = queue + warp * QCAP;
    float *wd2 = light_d2 + warp * (32 * R); // SHADOW: squared distance to the light; else: the distance clip of each ray
    unsigned int *wocc = SHADOW ? occluded + warp * R : nullptr;
    const unsigned int lt_mask = (1u << lane) - 1u;

    if (tid == 0) {
        if (round == 0) mbar_init(&tab_bar, 1);
        ctx.s0 = s0; ctx.s1 = s1;
        ctx.slot_meta = sc.slot_meta; ctx.slot_geom = sc.slot_geom; ctx.cube_tris = sc.cube_tris;
        ctx.lx = ot.ox; ctx.ly = ot.oy; ctx.lz = ot.oz;
        ctx.n_slots = sc.n_slots;
        ctx.cmax = sc.cmax;
    }
    if (tid < 2) tab[n_tab + tid] = make_float4(0.f, 0.f, 0.f, 0.f); // prefetched past the end, never used
    if (lane == 0 && round == 0) {
        n_exact[warp].all = n_exact[warp].sphere = n_exact[warp].triangle = 0;
        n_hier[warp] = 0;
#pragma unroll
        for (int g = 0; g < MAX_SEGS; ++g) n_slot[warp][g] = 0;
    }
    __syncthreads();
#if RT_SWEEP_BULK_TABLE == 0
    {
        const float4 *src = (ORIGIN ? ot.bound : sc.bound) + slot_begin;
        for (int k = tid; k < n_tab; k += SWEEP_THREADS) tab[k] = src[k];
        const float4 *gsrc = (ORIGIN ? ot.group : sc.group) + slot_begin / GROUP_SLOTS;
        for (int k = tid; k < n_gtab; k += SWEEP_THREADS) gtab[k] = gsrc[k];
        const float4 *qsrc = (ORIGIN ? ot.quad : sc.quad) + slot_begin / QUAD_SLOTS;
        for (int k = tid; k < n_qtab; k += SWEEP_THREADS) qtab[k] = qsrc[k];
        const float4 *ssrc = (ORIGIN ? ot.super : sc.super) + slot_begin / SUPER_SLOTS;
        for (int k = tid; k < n_stab; k += SWEEP_THREADS) stab[k] = ssrc[k];
    }
    __syncthreads();
    bool table_ready = true;
#elif RT_SWEEP_BULK_TABLE == 2
    if (warp == 0) { // the table (slots, groups, quads: three contiguous runs of float4) in chunks, one bulk copy per lane and run
        const unsigned int b_tab = (unsigned int)n_tab * sizeof(float4), b_g = (unsigned int)n_gtab * sizeof(float4),
                           b_q = (unsigned int)n_qtab * sizeof(float4), b_s = (unsigned int)n_stab * sizeof(float4);
        if (lane == 0) mbar_expect_tx(&tab_bar, b_tab + b_g + b_q + b_s);
        __syncwarp();
        auto chunked = [&](void *dst, const void *src, unsigned int bytes) {
            const unsigned int per = ((bytes + 31) / 32 + 15) & ~15u, off = (unsigned int)lane * per; // 16-byte multiples
            if (off < bytes) bulk_load((char *)dst + off, (const char *)src + off, min(per, bytes - off), &tab_bar);
        };
        chunked(tab, (ORIGIN ? ot.bound : sc.bound) + slot_begin, b_tab);
        chunked(gtab, (ORIGIN ? ot.group : sc.group) + slot_begin / GROUP_SLOTS, b_g);
        chunked(qtab, (ORIGIN ? ot.quad : sc.quad) + slot_begin / QUAD_SLOTS, b_q);
        chunked(stab, (ORIGIN ? ot.super : sc.super) + slot_begin / SUPER_SLOTS, b_s);
    }
    mbar_wait(&tab_bar, 0); // (the copies of a 171 KB table take ~2 us; waiting here keeps the item loop free of the flag)
    bool table_ready = true;
#else
    if (tid == 0) { // the table: slots, groups, quads -- three contiguous runs of float4
        const unsigned int b_tab = (unsigned int)n_tab * sizeof(float4), b_g = (unsigned int)n_gtab * sizeof(float4),
                           b_q = (unsigned int)n_qtab * sizeof(float4), b_s = (unsigned int)n_stab * sizeof(float4);
        mbar_expect_tx(&tab_bar, b_tab + b_g + b_q + b_s);
        bulk_load(tab, (ORIGIN ? ot.bound : sc.bound) + slot_begin, b_tab, &tab_bar);
        bulk_load(gtab, (ORIGIN ? ot.group : sc.group) + slot_begin / GROUP_SLOTS, b_g, &tab_bar);
        bulk_load(qtab, (ORIGIN ? ot.quad : sc.quad) + slot_begin / QUAD_SLOTS, b_q, &tab_bar);
        bulk_load(stab, (ORIGIN ? ot.super : sc.super) + slot_begin / SUPER_SLOTS, b_s, &tab_bar);
    }
    bool table_ready = false; // this warp has seen the copies land
#endif
    const float4 *tab0 = tab - slot_begin;                 // indexed by absolute slot
    const float4 *gtab0 = gtab - slot_begin / GROUP_SLOTS; // indexed by absolute group
    const float4 *qtab0 = qtab - slot_begin / QUAD_SLOTS;  // indexed by absolute quad
    const float4 *stab0 = stab - slot_begin / SUPER_SLOTS; // indexed by absolute super

    for (;;) {
        unsigned int item = 0;
        if (lane == 0) item = atomicAdd(batch_counter, 1u);
        item = __shfl_sync(FULL, item, 0);
        unsigned int batch;
        int slice, ss;
        if (!decode_item(item, total, R * 32, heavy_batches, heavy_slices, heavy_slice_slots, reg_batches, n_slices, slice_slots, tail_slices,
                         tail_slice_slots, batch, slice, ss))
            break; // past the plan
        const long long ray0_ll = (long long)batch * (R * 32);
        if (ray0_ll >= total) break;
        RT_CHECK(slice >= 0 && (slot_begin + slice * ss < slot_end || slot_end == slot_begin) && ss % SEG_PAD == 0); // (an empty table has one empty slice)
        const int ray0 = (int)ray0_ll;
        const int sb = slot_begin + slice * ss, se = min(slot_end, sb + ss);

        // vmask bits 2r, 2r+1: ray r of this lane exists, has no NaN component and is still
        // undecided (it can be queued); nanmask bit r: it exists and has a NaN component
        unsigned int vmask = 0, nanmask = 0;
        if constexpr (SHADOW) {
            if (lane < R) wocc[lane] = 0u;
        }
#if RT_SWEEP_PRELOAD
        // all R rays of the lane first, in one round trip: the plane tests below are calls the loads of the
        // later rays would otherwise queue up behind (long-scoreboard stalls were 60 % of this prologue's samples)
        RayRegs pre[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int gi = ray0 + r * 32 + lane;
            if (gi < total) pre[r] = load_ray(s0, s1, gi);
            else pre[r].ox = pre[r].oy = pre[r].oz = pre[r].dx = pre[r].dy = pre[r].dz = 0.f;
        }
#endif
#pragma unroll
        for (int r = 0; r < R; ++r) {
            int gi = ray0 + r * 32 + lane;
            unsigned long long first = KEY_MISS;
            if (gi < total) {
#if RT_SWEEP_PRELOAD
                const RayRegs rr = pre[r];
#else
                const RayRegs rr = load_ray(s0, s1, gi);
#endif
                if (ray_has_nan(rr)) nanmask |= 1u << r;
                else {
                    vmask |= 3u << (2 * r);
                    if constexpr (SHADOW) { // planes first: a hit at t < 0 is the nearest of all and decides alone
                        const V3 o = v3(rr.ox, rr.oy, rr.oz), d = v3(rr.dx, rr.dy, rr.dz);
                        const float d2 = vlen2(vsub(v3(ctx.lx, ctx.ly, ctx.lz), o));
                        wd2[r * 32 + lane] = d2;
                        first = plane_keys(sc.n_planes, sc.plane_shape, sc.plane_data, sc.geom, o, d, KEY_MISS);
                        if (first != KEY_MISS) {
                            const float tp = key_t(first);
                            if (tp < 0.f || occludes(o, d, tp, d2)) vmask &= ~(3u << (2 * r));
                        }
#if RT_SWEEP_SLICE_MAJOR
                        if (merge && slice > 0) { // what the items of earlier slices have found for this ray
                            const unsigned long long seen = __ldcg(s0.key + (long long)gi * s0.stride);
                            if (seen != KEY_MISS && occludes(o, d, key_t(seen), d2)) vmask &= ~(3u << (2 * r));
                        }
#endif
                    }
                }
            }
            wbest[r * 32 + lane] = first;
            if constexpr (!SHADOW) {
                float clip0 = 3.0e38f; // no hit yet: no clip
#if RT_SWEEP_SLICE_MAJOR && RT_SWEEP_CLIP && RT_SWEEP_PRELOAD
                if (merge && slice > 0 && gi < total && ((vmask >> (2 * r)) & 1u)) {
                    const unsigned long long seen = __ldcg(gi < s0.n ? s0.key + (long long)gi * s0.stride : s1.key + (gi - s0.n));
                    if (seen != KEY_MISS) { // the same clip sweep_flush derives from a hit
                        const RayRegs rr = pre[r];
                        const float n2 = fmaf(rr.dx, rr.dx, fmaf(rr.dy, rr.dy, rr.dz * rr.dz));
                        const float c0 = fmaf(key_t(seen) * sqrt_approx(n2), 1.001f,
                                              1e-4f * (sqrt_approx(fmaf(rr.ox, rr.ox, fmaf(rr.oy, rr.oy, rr.oz * rr.oz))) + sc.cmax));
                        if (n2 > 1e-30f && n2 < 1e30f && c0 >= 0.f) clip0 = c0;
                    }
                }
#endif
                wd2[r * 32 + lane] = clip0;
            }
        }
        __syncwarp();
        int qn = 0; // queued pairs of this warp (warp-uniform)
        bool alive = true; // SHADOW: some ray of the warp is still undecided
        if constexpr (SHADOW) alive = __any_sync(FULL, vmask != 0u);

        // ---- phase 1 + 2 --------------------------------------------------------------
        typename std::conditional<ORIGIN, RayCoefO, RayCoef>::type c[R];
        // e[r] = (filter value of ray r against the two spheres packed in T0, T1)
        // level: 0 bound slots, 1 group spheres, 2 quad spheres, 3 super spheres (wider margins, see GROUP_MARGIN)
        auto pair_e = [&](const float4 &T0, const float4 &T1, float2 (&e)[R], int level) {
            const float2 cx = make_float2(T0.x, T0.y), cy = make_float2(T0.z, T0.w), cz = make_float2(T1.x, T1.y);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if constexpr (ORIGIN) {
                    float2 b = __ffma2_rn(cx, f2(c[r].dx), __ffma2_rn(cy, f2(c[r].dy), __fmul2_rn(cz, f2(c[r].dz))));
                    e[r] = __ffma2_rn(b, b, f2(level == 0 ? c[r].mu : (level == 1 ? c[r].mu_g : (level == 2 ? QUAD_MARGIN : SUPER_MARGIN * QUAD_MARGIN) * c[r].mu_g)));
#if RT_SWEEP_CLIP
                    if constexpr (!SHADOW) { // (the rays start at O': b is the distance along the ray to the centre's foot)
                        if (level >= 1) {
                            const float lim = wd2[r * 32 + lane] + sc.level_rmax[level - 1];
                            if (b.x > lim) e[r].x = -3.0e38f; // the whole sphere lies beyond this ray's best hit
                            if (b.y > lim) e[r].y = -3.0e38f;
                        }
                    }
#endif
                } else {
                    float2 b = __ffma2_rn(cx, f2(c[r].dx), __ffma2_rn(cy, f2(c[r].dy), __ffma2_rn(cz, f2(c[r].dz), f2(c[r].kb))));
                    float2 s = __ffma2_rn(cx, f2(c[r].ox2), __ffma2_rn(cy, f2(c[r].oy2), __ffma2_rn(cz, f2(c[r].oz2),
                                           f2(level == 0 ? c[r].seed : (level == 1 ? c[r].seed_g : fmaf(level == 2 ? QUAD_SEED : SUPER_SEED, c[r].seed_g - c[r].seed, c[r].seed_g))))));
                    // a centre behind the origin (b < 0): only s decides, i.e. whether the origin is inside the sphere
                    e[r] = __ffma2_rn(b, make_float2(fmaxf(b.x, 0.f), fmaxf(b.y, 0.f)), s);
#if RT_SWEEP_CLIP
                    if (level >= 1) {
                        const float lim = wd2[r * 32 + lane] + sc.level_rmax[level - 1];
                        if (b.x > lim) e[r].x = -3.0e38f;
                        if (b.y > lim) e[r].y = -3.0e38f;
                    }
#endif
                }
            }
        };
        // SHADOW, after a flush: rays the exact stage has just found occluded leave the sweep -- they
        // pass no quad, group or slot any more (margin -> -3e38) and can no longer be queued
        auto drop_occluded = [&]() {
            if constexpr (SHADOW) {
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (((vmask >> (2 * r)) & 1u) && ((wocc[r] >> lane) & 1u)) {
                        vmask &= ~(3u << (2 * r));
                        c[r].mu = c[r].mu_g = -3.0e38f;
                    }
                alive = __any_sync(FULL, vmask != 0u);
            }
#if RT_SWEEP_PRIMARY_TF
            if constexpr (!SHADOW && ORIGIN) { // primary rays: the far clip of the slot tests follows the best hit (e = b^2 + mu, b from the camera)
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const float lim = wd2[r * 32 + lane] + sc.slot_rmax;
                    const float t = fmaf(lim, lim, c[r].mu) * 1.0001f;
                    if (t < c[r].tf) c[r].tf = t; // (false for NaN; no hit yet: lim = 3e38, t = inf)
                }
            }
#endif
        };
        if (!table_ready) { // first use of the table by this warp: the copies ran while it fetched and loaded its rays
            mbar_wait(&tab_bar, 0);
            table_ready = true;
        }
        // filter evaluations of this item, in units of 32 (one per lane): against group / quad spheres,
        // and (per segment of the table = per kind of shape) against bound slots
        unsigned int evals_hier = 0;
        // the segment of slot j (segments hold one kind of shape each; ends of unused segments are INT_MAX)
        auto seg_of = [&](int j) { return (j >= sc.segs.end[0] ? 1 : 0) + (j >= sc.segs.end[1] ? 1 : 0) + (j >= sc.segs.end[2] ? 1 : 0); };
        for (int g = 0; g < sc.runs.n && alive; ++g) { // runs of segments that share their margin factors
            const int gb = max(sc.runs.begin[g], sb), ge = min(sc.runs.end[g], se);
            if (gb >= ge) continue;
            // the coefficients depend on the run through its margin factor only
            const float k_seg = ORIGIN ? sc.runs.kx[g] : sc.runs.k[g];
            float ref_x = 0.f, ref_y = 0.f, ref_z = 0.f; // origin of this lane's first ray
            {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    bool usable = (vmask >> (2 * r)) & 1u;
                    RayRegs rr;
                    rr.ox = rr.oy = rr.oz = rr.dx = rr.dy = rr.dz = 0.f;
                    if (usable) rr = load_ray(s0, s1, ray0 + r * 32 + lane);
                    if (r == 0) { ref_x = rr.ox; ref_y = rr.oy; ref_z = rr.oz; }
                    if constexpr (ORIGIN) c[r] = ray_coef_origin(rr, k_seg, sc.cmax, ot, usable);
                    else c[r] = ray_coef(rr, k_seg, sc.cmax, usable);
                }
            }
#if RT_SWEEP_SUPERS
#if RT_SWEEP_NEAR_FIRST
            // Visiting order of the run's super pairs (at most 21 in a resident table): by the power of the item's first ray
            // origin with respect to the nearer sphere of the pair, |G - O|^2 - R^2 = W - 2 G.O + |O|^2 -- a heuristic for
            // "nearest first" only, nothing depends on it but how soon hits are found (SHADOW: rays leave; else: the clip).
            // Lane p holds the key of pair p with p in its low 5 bits; the minimum of the warp is visited and retired.
            unsigned int order_key = 0xffffffffu;
            const int n_pairs = (ge - gb) / (2 * SUPER_SLOTS);
            {
                float ox = __shfl_sync(FULL, ref_x, 0), oy = __shfl_sync(FULL, ref_y, 0), oz = __shfl_sync(FULL, ref_z, 0);
                if constexpr (ORIGIN) { ox -= ot.ox; oy -= ot.oy; oz -= ot.oz; } // the table is centred on O'
                if (lane < n_pairs) {
                    const float4 T0 = stab0[(gb >> 8) + 2 * lane], T1 = stab0[(gb >> 8) + 2 * lane + 1];
                    const float o2 = fmaf(ox, ox, fmaf(oy, oy, oz * oz));
                    const float p0 = fmaf(-2.f, fmaf(T0.x, ox, fmaf(T0.z, oy, T1.x * oz)), T1.z) + o2;
                    const float p1 = fmaf(-2.f, fmaf(T0.y, ox, fmaf(T0.w, oy, T1.y * oz)), T1.w) + o2;
                    unsigned int u = __float_as_uint(fminf(p0, p1));
                    u ^= (u >> 31) ? 0xffffffffu : 0x80000000u; // orderable
                    order_key = (u & ~31u) | (unsigned int)lane;
                }
            }
            RT_CHECK(n_pairs <= 31);
            for (int it = 0; it < n_pairs && alive; ++it) {
            const unsigned int nearest = __reduce_min_sync(FULL, order_key);
            if (order_key == nearest) order_key = 0xffffffffu; // retired
            const int js = gb + (int)(nearest & 31u) * 2 * SUPER_SLOTS;
#else
            for (int js = gb; js < ge && alive; js += 2 * SUPER_SLOTS) { // runs and slices are whole pairs of supers (SEG_PAD)
#endif
            // which of the two supers (4 quads each) can any ray of this warp touch?
            unsigned int supers_hit;
            {
                const float4 T0 = stab0[js >> 8], T1 = stab0[(js >> 8) + 1];
                float2 e[R];
                pair_e(T0, T1, e, 3);
                float m0 = e[0].x, m1 = e[0].y;
#pragma unroll
                for (int r = 1; r < R; ++r) { m0 = fmaxf(m0, e[r].x); m1 = fmaxf(m1, e[r].y); }
                supers_hit = (__any_sync(FULL, m0 >= T1.z) ? 1u : 0u) | (__any_sync(FULL, m1 >= T1.w) ? 2u : 0u);
                evals_hier += 2 * R;
            }
            if (!supers_hit) continue;
            const int jq_begin = (supers_hit & 1u) ? js : js + SUPER_SLOTS, jq_end = (supers_hit & 2u) ? js + 2 * SUPER_SLOTS : js + SUPER_SLOTS;
#else
            {
            const int jq_begin = gb, jq_end = ge;
#endif
            for (int jq = jq_begin; jq < jq_end && alive; jq += 2 * QUAD_SLOTS) {
            RT_CHECK(jq >= slot_begin && jq + 2 * QUAD_SLOTS <= slot_end && (jq & (2 * QUAD_SLOTS - 1)) == 0);
            // which of the two quads (4 groups each) can any ray of this warp touch?
            unsigned int quads_hit;
            {
                const float4 T0 = qtab0[jq >> 6], T1 = qtab0[(jq >> 6) + 1];
                float2 e[R];
                pair_e(T0, T1, e, 2);
                float m0 = e[0].x, m1 = e[0].y;
#pragma unroll
                for (int r = 1; r < R; ++r) { m0 = fmaxf(m0, e[r].x); m1 = fmaxf(m1, e[r].y); }
                quads_hit = (__any_sync(FULL, m0 >= T1.z) ? 1u : 0u) | (__any_sync(FULL, m1 >= T1.w) ? 2u : 0u);
                evals_hier += 2 * R;
            }
            // most quad pairs are touched by no ray of the warp: leave before the group loop is set up
            // (its control flow for skipped quads was 12 % of the instructions executed)
            if (!quads_hit) continue;
            constexpr int GP_PER_QUAD = QUAD_SLOTS / (2 * GROUP_SLOTS); // pairs of groups in a quad
            const int gp_begin = (quads_hit & 1u) ? 0 : GP_PER_QUAD, gp_end = (quads_hit & 2u) ? 2 * GP_PER_QUAD : GP_PER_QUAD;
            for (int gp = gp_begin; gp < gp_end && alive; ++gp) { // the pairs of groups of the touched quads
                const int j0 = jq + gp * 2 * GROUP_SLOTS;
                // which of the two groups can any ray of this warp touch?
                unsigned int ball0, ball1, pm = 0;
                {
                    const float4 T0 = gtab0[j0 >> 4], T1 = gtab0[(j0 >> 4) + 1];
                    float2 e[R];
                    pair_e(T0, T1, e, 1);
                    float m0 = e[0].x, m1 = e[0].y;
#pragma unroll
                    for (int r = 1; r < R; ++r) { m0 = fmaxf(m0, e[r].x); m1 = fmaxf(m1, e[r].y); }
                    ball0 = __ballot_sync(FULL, m0 >= T1.z);
                    ball1 = __ballot_sync(FULL, m1 >= T1.w);
                    evals_hier += 2 * R;
                    const int p0 = __popc(ball0), p1 = __popc(ball1);
                    if ((p0 && p0 <= item_max) || (p1 && p1 <= item_max)) { // bit 2r + h of pm: ray r of this lane may touch group h
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            if (e[r].x >= T1.z) pm |= 1u << (2 * r);
                            if (e[r].y >= T1.w) pm |= 2u << (2 * r);
                        }
                        pm &= vmask;
                    }
                }
                for (int h = 0; h < 2 && alive; ++h) {
                    const unsigned int gball = h ? ball1 : ball0; // lanes with a ray that may touch group h
                    if (!gball) continue;
                    const int jg = j0 + h * GROUP_SLOTS;
                    if (__popc(gball) <= item_max) {
                        // Few (ray, group) pairs: one pair per half warp, lane = slot of the group.
                        const int js = jg + (lane & (GROUP_SLOTS - 1));
                        const float *ts = (const float *)(tab0 + (js & ~1)) + (js & 1);
                        const float scx = ts[0], scy = ts[2], scz = ts[4], sw = ts[6];
                        unsigned int evals = 0;
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            unsigned int ball = __ballot_sync(FULL, (pm >> (2 * r + h)) & 1u);
                            while (ball) {
                                const int l0 = __ffs(ball) - 1;
                                ball &= ball - 1;
                                int l1 = -1;
                                if (ball) { l1 = __ffs(ball) - 1; ball &= ball - 1; }
                                const int src = (lane < 16 || l1 < 0) ? l0 : l1;
                                float ev, clip = 3.4e38f;
                                if constexpr (ORIGIN) {
                                    const float dx = __shfl_sync(FULL, c[r].dx, src), dy = __shfl_sync(FULL, c[r].dy, src),
                                                dz = __shfl_sync(FULL, c[r].dz, src), mu = __shfl_sync(FULL, c[r].mu, src);
                                    clip = __shfl_sync(FULL, c[r].tf, src);
                                    const float b = fmaf(scx, dx, fmaf(scy, dy, scz * dz));
                                    ev = fmaf(b, b, mu);
                                } else {
                                    const float dx = __shfl_sync(FULL, c[r].dx, src), dy = __shfl_sync(FULL, c[r].dy, src),
                                                dz = __shfl_sync(FULL, c[r].dz, src), kb = __shfl_sync(FULL, c[r].kb, src);
                                    const float ox2 = __shfl_sync(FULL, c[r].ox2, src), oy2 = __shfl_sync(FULL, c[r].oy2, src),
                                                oz2 = __shfl_sync(FULL, c[r].oz2, src), seed = __shfl_sync(FULL, c[r].seed, src);
                                    const float b = fmaf(scx, dx, fmaf(scy, dy, fmaf(scz, dz, kb)));
                                    const float sv = fmaf(scx, ox2, fmaf(scy, oy2, fmaf(scz, oz2, seed)));
                                    ev = fmaf(b, fmaxf(b, 0.f), sv);
                                }
                                ++evals;
                                const bool cand = ev >= sw && ev <= clip && (lane < 16 || l1 >= 0);
                                const unsigned int cball = __ballot_sync(FULL, cand);
                                if (cball) {
                                    if (qn > QCAP - 32) {
                                        qn = sweep_flush<SHADOW>(&ctx, wqueue, wbest, ray0, qn, 0, &n_exact[warp], wd2, wocc);
                                        drop_occluded(); // (a ray dropped here may still be queued once more by this loop: harmless)
                                    }
                                    RT_CHECK(qn + 32 <= QCAP && js >= sb && js < se);
                                    if (cand) wqueue[qn + __popc(cball & lt_mask)] = ((unsigned int)(r * 32 + src) << 24) | (unsigned int)js;
                                    qn += __popc(cball);
                                    __syncwarp();
                                }
                            }
                        }
                        if (lane == 0) n_slot[warp][seg_of(jg)] += evals;
                        if (qn >= (SHADOW ? RT_SWEEP_FLUSH_AT_SHADOW : RT_SWEEP_FLUSH_AT)) {
                            qn = sweep_flush<SHADOW>(&ctx, wqueue, wbest, ray0, qn, 0, &n_exact[warp], wd2, wocc);
                            drop_occluded();
                        }
                        continue;
                    }
                    const float4 *tp = tab0 + jg;
                    float4 N0 = tp[0], N1 = tp[1]; // next slot pair, fetched one iteration ahead
                    if (lane == 0) n_slot[warp][seg_of(jg)] += (GROUP_SLOTS / 2) * 2 * R;
#pragma unroll BRUTE_UNROLL
                    for (int i = 0; i < GROUP_SLOTS / 2; ++i) {
                        const float4 T0 = N0, T1 = N1;
                        N0 = tp[2 * i + 2]; N1 = tp[2 * i + 3]; // past the group: the next one's, or the table's slack pair
                        float2 e[R];
                        pair_e(T0, T1, e, 0);
                        float m0 = e[0].x, m1 = e[0].y;
#pragma unroll
                        for (int r = 1; r < R; ++r) { m0 = fmaxf(m0, e[r].x); m1 = fmaxf(m1, e[r].y); }
                        if (__any_sync(FULL, m0 >= T1.z || m1 >= T1.w)) { // warp-uniform: some lane has a candidate
                            unsigned int mask = 0;
#pragma unroll
                            for (int r = 0; r < R; ++r) {
                                if constexpr (ORIGIN) {
                                    if (e[r].x >= T1.z && e[r].x <= c[r].tf) mask |= 1u << (2 * r);
                                    if (e[r].y >= T1.w && e[r].y <= c[r].tf) mask |= 2u << (2 * r);
                                } else {
                                    if (e[r].x >= T1.z) mask |= 1u << (2 * r);
                                    if (e[r].y >= T1.w) mask |= 2u << (2 * r);
                                }
                            }
                            mask &= vmask;
                            unsigned int ball;
                            while ((ball = __ballot_sync(FULL, mask != 0)) != 0) { // one queue slot per lane per round
                                if (qn > QCAP - 32) {
                                    qn = sweep_flush<SHADOW>(&ctx, wqueue, wbest, ray0, qn, 0, &n_exact[warp], wd2, wocc);
                                    drop_occluded();
                                    mask &= vmask;
                                }
                                RT_CHECK(qn + 32 <= QCAP);
                                if (mask) {
                                    int bit = __ffs(mask) - 1;
                                    mask &= mask - 1;
                                    unsigned int local = (unsigned int)((bit >> 1) * 32 + lane);
                                    wqueue[qn + __popc(ball & lt_mask)] = (local << 24) | (unsigned int)(jg + 2 * i + (bit & 1));
                                }
                                qn += __popc(ball);
                                __syncwarp();
                            }
                        }
                    }
                    if (qn >= (SHADOW ? RT_SWEEP_FLUSH_AT_SHADOW : RT_SWEEP_FLUSH_AT)) {
                        qn = sweep_flush<SHADOW>(&ctx, wqueue, wbest, ray0, qn, 0, &n_exact[warp], wd2, wocc);
                        drop_occluded();
                    }
                }
            }
            }
            } // supers
        }
        if (lane == 0) n_hier[warp] += evals_hier;
        __syncwarp();
        sweep_flush<SHADOW>(&ctx, wqueue, wbest, ray0, qn, 1, &n_exact[warp], wd2, wocc);

        // ---- phase 3 + output -----------------------------------------------------------
#pragma unroll
        for (int r = 0; r < R; ++r) {
            int gi = ray0 + r * 32 + lane;
            if (gi >= total) continue;
            unsigned long long k;
            if ((nanmask >> r) & 1u) {
                k = sc.first_nan_shape >= 0
                        ? (((unsigned long long)KEY_NAN_PREFIX << 32) | (unsigned long long)(uint32_t)sc.first_nan_shape)
                        : KEY_MISS;
            } else {
                k = wbest[r * 32 + lane];
                if (!SHADOW && planes && slice == 0) { // (SHADOW: the planes went first, into wbest)
                    RayRegs rr = load_ray(s0, s1, gi);
                    k = plane_keys(sc.n_planes, sc.plane_shape, sc.plane_data, sc.geom, v3(rr.ox, rr.oy, rr.oz),
                                   v3(rr.dx, rr.dy, rr.dz), k);
                }
            }
            RT_CHECK(gi >= 0 && gi < total);
            unsigned long long *kp = gi < s0.n ? s0.key + (long long)gi * s0.stride : s1.key + (gi - s0.n);
            if (!merge) *kp = k;
            else if (k != KEY_MISS) atomicMin(kp, k);
        }
        __syncwarp();
    }
    if (!table_ready) mbar_wait(&tab_bar, 0); // never leave a CTA with its bulk copies in flight
    } // rounds (JOBS)
    if (exact_counter && lane == 0) {
        if (n_exact[warp].all) {
            atomicAdd(exact_counter, (unsigned long long)n_exact[warp].all);
            atomicAdd(exact_counter + 3 + MAX_SEGS, (unsigned long long)n_exact[warp].sphere);
            atomicAdd(exact_counter + 4 + MAX_SEGS, (unsigned long long)n_exact[warp].triangle);
        }
        // (ray, sphere) pairs the filter evaluated: 2 spheres x 32 R rays per iteration
        unsigned long long slots = 0;
#pragma unroll
        for (int g = 0; g < MAX_SEGS; ++g) {
            if (n_slot[warp][g]) atomicAdd(exact_counter + 3 + g, n_slot[warp][g] * 32ull); // evaluations of bound slots, by segment
            slots += n_slot[warp][g];
        }
        if (slots + n_hier[warp]) atomicAdd(exact_counter + (ORIGIN ? 2 : 1), (slots + n_hier[warp]) * 32ull);
    }
}

// ---- audit: every pair exact, in shape order (what the reference literally does) --------
__global__ void __launch_bounds__(256)
sweep_all_pairs_kernel(DevScene sc, RaySeg s0, RaySeg s1, unsigned long long *exact_counter) {
    int gi = blockIdx.x * blockDim.x + threadIdx.x;
    int total = s0.n + s1.n;
    if (gi >= total) return;
    RayRegs rr = load_ray(s0, s1, gi);
    V3 o = v3(rr.ox, rr.oy, rr.oz), d = v3(rr.dx, rr.dy, rr.dz);
    unsigned long long best = KEY_MISS;
    if (ray_has_nan(rr)) {
        if (sc.first_nan_shape >= 0)
            best = ((unsigned long long)KEY_NAN_PREFIX << 32) | (unsigned long long)(uint32_t)sc.first_nan_shape;
    } else {
        for (int j = 0; j < sc.n_slots; ++j) {
            int shape = sc.bound_shape[j];
            if (shape < 0) continue;
            unsigned long long k = exact_pair(sc.shape_meta, sc.geom, sc.cube_tris, shape, o, d);
            best = k < best ? k : best;
        }
        best = plane_keys(sc.n_planes, sc.plane_shape, sc.plane_data, sc.geom, o, d, best);
    }
    if (gi < s0.n) s0.key[(long long)gi * s0.stride] = best;
    else s1.key[gi - s0.n] = best;
    (void)exact_counter; // rays x n_bounded, added by the host
}

template <int R, int T, bool SHADOW> static size_t sweep_smem_fixed() {
    return 2 * sizeof(float4) + (size_t)T * R * sizeof(unsigned long long) + (size_t)(T / 32) * QCAP * sizeof(unsigned int) +
           (size_t)T * R * sizeof(float) + (SHADOW ? (size_t)(T / 32) * R * sizeof(unsigned int) : 0);
}

// Bound slots one launch can keep in shared memory.
template <int R, int T, bool SHADOW> static int sweep_table_cap() {
    // per slot: its float4, 1/GROUP_SLOTS of its group's, 1/QUAD_SLOTS of its quad's and 1/SUPER_SLOTS of its super's
    int cap = (int)((SMEM_LIMIT - 1024 - sweep_smem_fixed<R, T, SHADOW>()) * SUPER_SLOTS /
                    (sizeof(float4) * (SUPER_SLOTS + SUPER_SLOTS / GROUP_SLOTS + SUPER_SLOTS / QUAD_SLOTS + 1))) /
              SEG_PAD * SEG_PAD;
    const int forced = tuning().table_cap / SEG_PAD * SEG_PAD; // tests force several table chunks
    if (forced >= SEG_PAD && forced < cap) cap = forced;
    return cap;
}

// How the rays of one sweep are cut into work items (ray batches x table slices), for `warps`
// persistent warps and a resident table of `per` slots.
struct ItemPlan {
    int n_slices = 1, slice_slots = SEG_PAD, merge = 0;
    long long heavy_batches = 0;
    int heavy_slices = 1, heavy_slice_slots = SEG_PAD;
    long long reg_batches = 0; // batches after the heavy ones cut into n_slices; the rest into tail_slices
    int tail_slices = 1, tail_slice_slots = SEG_PAD;
    long long items = 0;
};

template <int R>
static ItemPlan plan_items(long long total, int heavy, int per, long long warps, int n_chunks, bool jobs = false) {
    const Tuning &tn = tuning();
    ItemPlan pl;
    const long long batches = (total + 32ll * R - 1) / (32ll * R);
    // Work items = ray batches x table slices.  A launch with few rays is cut into slices of
    // the table so that every warp of the GPU still gets several items; the batches of the
    // first `heavy` rays (origins far outside the scene: filter margins so wide that they
    // pass most of the table on to the exact test) are cut as finely as the table allows, since
    // one of them whole can take longer than the rest of the launch.
    auto slicing = [&](long long want, long long n_batches, int *slots) {
        int n = (int)std::min<long long>((want + n_batches - 1) / std::max<long long>(n_batches, 1), std::max(1, per / MIN_SLICE_SLOTS));
        n = std::max(n, 1);
        *slots = std::max(SEG_PAD, ((per + n - 1) / n + SEG_PAD - 1) / SEG_PAD * SEG_PAD);
        return std::max(1, (per + *slots - 1) / *slots);
    };
    pl.heavy_batches = std::min<long long>(batches, ((long long)std::max(heavy, 0) + 32ll * R - 1) / (32ll * R));
    const long long want_light = (long long)tn.want_light * warps, want_heavy = (long long)(jobs ? tn.want_heavy_jobs : tn.want_heavy) * warps; // measured (tuning.h)
    pl.n_slices = slicing(want_light, batches - pl.heavy_batches, &pl.slice_slots);
    if (tn.slices > 0) {
        pl.n_slices = std::max(1, std::min(tn.slices, std::max(1, per / SEG_PAD)));
        pl.slice_slots = std::max(SEG_PAD, ((per + pl.n_slices - 1) / pl.n_slices + SEG_PAD - 1) / SEG_PAD * SEG_PAD);
        pl.n_slices = std::max(1, (per + pl.slice_slots - 1) / pl.slice_slots);
    }
    if (pl.heavy_batches > 0) {
        pl.heavy_slices = slicing(want_heavy, pl.heavy_batches, &pl.heavy_slice_slots);
        if (pl.heavy_slices < pl.n_slices) { pl.heavy_slices = pl.n_slices; pl.heavy_slice_slots = pl.slice_slots; }
    }
    // A warp's last item ends, on average, half an item before the launch does.  The last
    // tail_pct % of the batches (at most 2 per warp) are therefore cut tail_mult times finer (measured: tuning.h).
    const long long regular = batches - pl.heavy_batches;
    const int tail_pct = tn.tail_pct, tail_mult = tn.tail_mult;
    long long tail_batches = std::min<long long>(regular * tail_pct / 100, 2 * warps);
    if (regular < (long long)tn.tail_min * warps) tail_batches = 0; // a short launch: finer items cost more (prologues, shadow rays stopping later) than its tail
    pl.tail_slices = pl.n_slices; pl.tail_slice_slots = pl.slice_slots;
    if (tail_batches > 0 && tail_mult > 1) {
        const int n = std::min(pl.n_slices * tail_mult, std::max(1, per / MIN_SLICE_SLOTS));
        pl.tail_slice_slots = std::max(SEG_PAD, ((per + n - 1) / n + SEG_PAD - 1) / SEG_PAD * SEG_PAD);
        pl.tail_slices = std::max(1, (per + pl.tail_slice_slots - 1) / pl.tail_slice_slots);
    }
    if (pl.tail_slices <= pl.n_slices) { tail_batches = 0; pl.tail_slices = pl.n_slices; pl.tail_slice_slots = pl.slice_slots; }
    pl.reg_batches = regular - tail_batches;
    pl.merge = (pl.n_slices > 1 || n_chunks > 1 || (pl.heavy_batches > 0 && pl.heavy_slices > 1) || tail_batches > 0) ? 1 : 0;
    pl.items = pl.heavy_batches * pl.heavy_slices + pl.reg_batches * pl.n_slices + tail_batches * pl.tail_slices;
    return pl;
}

template <int R, int T, bool SHADOW> static size_t sweep_smem_bytes(int slots) {
    return (size_t)(slots + slots / GROUP_SLOTS + slots / QUAD_SLOTS + slots / SUPER_SLOTS) * sizeof(float4) + sweep_smem_fixed<R, T, SHADOW>();
}

// Opt in to SMEM_LIMIT bytes of dynamic shared memory, once per kernel and device rather than per launch
// (a driver call that takes a process-wide lock: it showed when 8 threads drive 8 devices).
template <typename K> static cudaError_t allow_big_smem(K kernel, std::atomic<unsigned long long> &done) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64 && ((done.load(std::memory_order_acquire) >> dev) & 1ull)) return cudaSuccess;
    cudaFuncAttributes fa{};
    if ((e = cudaFuncGetAttributes(&fa, kernel)) != cudaSuccess) return e;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT - (int)fa.sharedSizeBytes); // static + dynamic <= 227 KB
    if (e == cudaSuccess && dev >= 0 && dev < 64) done.fetch_or(1ull << dev, std::memory_order_release);
    return e;
}

template <int R, int T, bool ORIGIN, bool SHADOW>
static cudaError_t launch_sweep_r(const DevScene &sc, RaySeg s0, RaySeg s1, OriginTable ot, long long total, int sm_count,
                                  unsigned int *batch_counters, int n_counters, int *used, unsigned long long *exact_counter,
                                  cudaStream_t stream) {
    constexpr int SWEEP_WARPS = T / 32;
    const int cap = sweep_table_cap<R, T, SHADOW>();
    const int n_chunks = std::max(1, (sc.n_slots + cap - 1) / cap);
    if (n_chunks > n_counters) return cudaErrorInvalidValue;
    // chunks of equal size (rounded to SEG_PAD) rather than cap, cap, ..., rest
    const int per = ((sc.n_slots + n_chunks - 1) / n_chunks + SEG_PAD - 1) / SEG_PAD * SEG_PAD;
    const ItemPlan pl = plan_items<R>(total, s0.heavy, per, (long long)sm_count * SWEEP_WARPS, n_chunks);
    if (pl.merge) { // every item folds its keys in with atomicMin: start from "no hit"
        cudaError_t err = cudaSuccess;
        if (s0.n > 0 && !s0.preset) {
            if (s0.stride != 1) return cudaErrorInvalidValue; // the caller presets the keys of a strided segment
            err = cudaMemsetAsync(s0.key, 0xFF, sizeof(unsigned long long) * (size_t)s0.n, stream);
        }
        if (err == cudaSuccess && s1.n > 0) err = cudaMemsetAsync(s1.key, 0xFF, sizeof(unsigned long long) * (size_t)s1.n, stream);
        if (err != cudaSuccess) return err;
    }
    unsigned int grid = (unsigned int)std::min<long long>((pl.items + SWEEP_WARPS - 1) / SWEEP_WARPS, (long long)sm_count);
    if (tuning().trace)
        std::fprintf(stderr, "[rt_b200]   plan %s%s: %lld rays, R %d, %lld items = heavy %lld x %d + %lld x %d + tail x %d slices, grid %u\n",
                     ORIGIN ? "origin" : "general", SHADOW ? " shadow" : "", total, R, pl.items, pl.heavy_batches, pl.heavy_slices,
                     pl.reg_batches, pl.n_slices, pl.tail_slices, grid);
    static const SweepJobs no_jobs{};
    for (int c = 0; c < n_chunks; ++c) {
        int b = c * per, e = std::min(sc.n_slots, b + per);
        const size_t smem = sweep_smem_bytes<R, T, SHADOW>(e - b);
        static std::atomic<unsigned long long> smem_allowed{0}; // per instantiation: bit d = done on device d
        cudaError_t err = allow_big_smem(sweep_kernel<R, T, ORIGIN, false, SHADOW>, smem_allowed);
        if (err != cudaSuccess) return err;
        sweep_kernel<R, T, ORIGIN, false, SHADOW><<<grid, T, smem, stream>>>(sc, s0, s1, ot, b, e, pl.n_slices, pl.slice_slots, pl.merge,
                                                                             c == n_chunks - 1 ? 1 : 0, tuning().item_max, (int)pl.heavy_batches,
                                                                             pl.heavy_slices, pl.heavy_slice_slots, (int)pl.reg_batches,
                                                                             pl.tail_slices, pl.tail_slice_slots, batch_counters + c,
                                                                             exact_counter, no_jobs);
        if ((err = cudaGetLastError()) != cudaSuccess) return err;
    }
    *used = n_chunks;
    return cudaSuccess;
}

// Rays per thread of a launch: fewer only when even table slicing cannot occupy the warps.
static int pick_rays_per_thread(long long total, int n_slots, int sm_count) {
    if (tuning().rays_per_thread > 0) return tuning().rays_per_thread;
    const long long want = 2ll * sm_count * 16;
    const long long max_slices = std::max(1, n_slots / MIN_SLICE_SLOTS);
    int r = 4;
    while (r > 1 && ((total + 32ll * r - 1) / (32ll * r)) * max_slices < want) r >>= 1;
    return r;
}

template <int R, int T, bool SHADOW>
static cudaError_t launch_sweep_jobs_r(const DevScene &sc, const RaySeg *segs, const OriginTable *tabs, int n_jobs, int sm_count,
                                       unsigned int *batch_counters, unsigned long long *exact_counter, cudaStream_t stream) {
    constexpr int SWEEP_WARPS = T / 32;
    if (sc.n_slots > sweep_table_cap<R, T, SHADOW>()) return cudaErrorNotSupported; // one resident table only
    SweepJobs jobs{};
    jobs.n = n_jobs;
    jobs.rotate = tuning().job_rotate;
    const unsigned int grid = (unsigned int)sm_count;
    for (int j = 0; j < n_jobs; ++j) {
        const long long ctas = (sm_count - j + n_jobs - 1) / n_jobs; // CTAs b with b % n_jobs == j
        const ItemPlan pl = plan_items<R>(segs[j].n, segs[j].heavy, sc.n_slots, ctas * SWEEP_WARPS, 1, true);
        SweepJob &J = jobs.j[j];
        J.s0 = segs[j]; J.ot = tabs[j]; J.batch_counter = batch_counters + j;
        J.n_slices = pl.n_slices; J.slice_slots = pl.slice_slots; J.merge = pl.merge;
        J.heavy_batches = (int)pl.heavy_batches; J.heavy_slices = pl.heavy_slices; J.heavy_slice_slots = pl.heavy_slice_slots;
        J.reg_batches = (int)pl.reg_batches; J.tail_slices = pl.tail_slices; J.tail_slice_slots = pl.tail_slice_slots;
        J.n_items = (unsigned int)std::min<long long>(pl.items, 0xffffffffll);
        if (tuning().trace)
            std::fprintf(stderr, "[rt_b200]   plan job %d: %d rays, R %d, %lld items = heavy %lld x %d + %lld x %d + tail x %d slices, %lld CTAs\n", j,
                         segs[j].n, R, pl.items, pl.heavy_batches, pl.heavy_slices, pl.reg_batches, pl.n_slices, pl.tail_slices, ctas);
    }
    const size_t smem = sweep_smem_bytes<R, T, SHADOW>(sc.n_slots);
    static std::atomic<unsigned long long> smem_allowed{0};
    cudaError_t err = allow_big_smem(sweep_kernel<R, T, true, true, SHADOW>, smem_allowed);
    if (err != cudaSuccess) return err;
    const RaySeg none{nullptr, nullptr, nullptr, 0};
    sweep_kernel<R, T, true, true, SHADOW><<<grid, T, smem, stream>>>(sc, none, none, OriginTable{}, 0, sc.n_slots, 1, SEG_PAD, 0, 1, tuning().item_max,
                                                                      0, 1, SEG_PAD, 0, 1, SEG_PAD, nullptr, exact_counter, jobs);
    return cudaGetLastError();
}

cudaError_t launch_sweep_jobs(const DevScene &sc, const RaySeg *segs, const OriginTable *tabs, int n_jobs, int sm_count,
                              unsigned int *batch_counters, int n_counters, int *used, unsigned long long *exact_counter,
                              cudaStream_t stream) {
    *used = 0;
    if (n_jobs < 2 || n_jobs > MAX_SWEEP_JOBS || n_jobs > n_counters || sm_count < n_jobs) return cudaErrorNotSupported;
    long long most = 0;
    for (int j = 0; j < n_jobs; ++j) {
        // keys are folded in with atomicMin when a job's items are table slices: the caller presets them
        if (!tabs[j].bound || segs[j].n <= 0 || !segs[j].preset) return cudaErrorNotSupported;
        most = std::max<long long>(most, segs[j].n);
    }
    int r = pick_rays_per_thread(most, sc.n_slots, std::max(1, sm_count / n_jobs));
    if (tuning().rays_per_thread == 0 && most < (long long)tuning().small_shadow_rays) r = std::min(r, 2); // (A/B knob; measured: no gain)
    const bool anyhit = tuning().shadow_anyhit != 0;
    cudaError_t err;
#define RT_JOBS_T(R_, T_) \
    (anyhit ? launch_sweep_jobs_r<R_, T_, true>(sc, segs, tabs, n_jobs, sm_count, batch_counters, exact_counter, stream) \
            : launch_sweep_jobs_r<R_, T_, false>(sc, segs, tabs, n_jobs, sm_count, batch_counters, exact_counter, stream))
#define RT_JOBS(R_) RT_JOBS_T(R_, 512)
#if RT_SWEEP_R8
    if (r >= 8) err = RT_JOBS_T(8, 256);
    else
#endif
    if (r >= 4) err = RT_JOBS_T(4, RT_SWEEP_T4);
    else if (r >= 2) err = RT_JOBS_T(2, RT_SWEEP_T2);
    else err = RT_JOBS(1);
#undef RT_JOBS
#undef RT_JOBS_T
    if (err == cudaSuccess) *used = n_jobs;
    return err;
}

cudaError_t launch_sweep(const DevScene &sc, RaySeg s0, RaySeg s1, OriginTable ot, int sm_count,
                         unsigned int *batch_counters, int n_counters, int *used, unsigned long long *exact_counter,
                         cudaStream_t stream) {
    long long total = (long long)s0.n + s1.n;
    *used = 0;
    if (total <= 0) return cudaSuccess;
    if (ot.bound && s1.n > 0) return cudaErrorInvalidValue;
    if (sm_count < 1) sm_count = 1;
    int r = pick_rays_per_thread(total, sc.n_slots, sm_count);
    // sweeps with the general filter (reflected / refracted rays, far less coherent than primary or shadow rays): batches of
    // 64 rays make narrower beams (and a small sweep needs half the table slices to occupy the warps)
    if (!ot.bound && tuning().rays_per_thread == 0 && total < (long long)tuning().small_tree_rays) r = std::min(r, std::max(1, tuning().tree_rays_per_thread));
#define RT_SWEEP(R_, T_, O_, S_) \
    launch_sweep_r<R_, T_, O_, S_>(sc, s0, s1, ot, total, sm_count, batch_counters, n_counters, used, exact_counter, stream)
#if RT_SWEEP_R8
    if (r >= 8) {
        if (ot.bound && ot.shadow && tuning().shadow_anyhit) return RT_SWEEP(8, 256, true, true);
        if (ot.bound) return RT_SWEEP(8, 256, true, false);
        return RT_SWEEP(8, 256, false, false);
    }
#endif
    if (ot.bound && ot.shadow && tuning().shadow_anyhit) { // the shadow rays of one light
        if (r >= 4) return RT_SWEEP(4, RT_SWEEP_T4, true, true);
        if (r >= 2) return RT_SWEEP(2, RT_SWEEP_T2, true, true);
        return RT_SWEEP(1, 512, true, true);
    }
    if (ot.bound) {
        if (r >= 4) return RT_SWEEP(4, RT_SWEEP_T4, true, false);
        if (r >= 2) return RT_SWEEP(2, RT_SWEEP_T2, true, false);
        return RT_SWEEP(1, 512, true, false);
    }
    if (r >= 4) return RT_SWEEP(4, RT_SWEEP_T4, false, false); // 16 warps x <= 128 registers; more warps spill the filter loop
    if (r >= 2) return RT_SWEEP(2, RT_SWEEP_T2, false, false);
    return RT_SWEEP(1, 512, false, false);
#undef RT_SWEEP
}

} // namespace rt

// Test hook (include/rt_b200_debug.h): the item plan of a sweep of `total` rays (`heavy` of them expensive) against a resident
// table of `per` slots on `warps` persistent warps, and the decoding of one of its items -- host code only, no GPU.
extern "C" int rt_debug_plan_items(int32_t rays_per_thread, int64_t total, int32_t heavy, int32_t per, int64_t warps, int32_t n_chunks,
                                   int32_t jobs, int64_t out[10]) {
    using namespace rt;
    if (!out || total < 0 || per < SEG_PAD || per % SEG_PAD != 0 || warps < 1 || n_chunks < 1) return -1;
    ItemPlan pl;
    if (rays_per_thread == 4) pl = plan_items<4>(total, heavy, per, warps, n_chunks, jobs != 0);
    else if (rays_per_thread == 2) pl = plan_items<2>(total, heavy, per, warps, n_chunks, jobs != 0);
    else if (rays_per_thread == 1) pl = plan_items<1>(total, heavy, per, warps, n_chunks, jobs != 0);
    else return -1;
    out[0] = pl.items; out[1] = pl.heavy_batches; out[2] = pl.heavy_slices; out[3] = pl.heavy_slice_slots; out[4] = pl.reg_batches;
    out[5] = pl.n_slices; out[6] = pl.slice_slots; out[7] = pl.tail_slices; out[8] = pl.tail_slice_slots; out[9] = pl.merge;
    return 0;
}
extern "C" int rt_debug_decode_item(uint32_t item, int64_t total, int32_t batch_rays, const int64_t plan[10], int64_t out[3]) {
    if (!plan || !out || batch_rays < 1) return -1;
    unsigned int batch = 0;
    int slice = 0, ss = 0;
    const bool ok = rt::decode_item(item, total, batch_rays, (int)plan[1], (int)plan[2], (int)plan[3], (int)plan[4], (int)plan[5], (int)plan[6],
                                    (int)plan[7], (int)plan[8], batch, slice, ss);
    out[0] = batch; out[1] = slice; out[2] = ss;
    return ok ? 1 : 0;
}

namespace rt {

int sweep_table_chunks(const DevScene &sc) {
    const int cap = sweep_table_cap<4, RT_SWEEP_T4, true>(); // the smallest of the kernels' capacities
    return std::max(1, (sc.n_slots + cap - 1) / cap);
}

void launch_sweep_all_pairs(const DevScene &sc, RaySeg s0, RaySeg s1, unsigned long long *exact_counter,
                            cudaStream_t stream) {
    long long total = (long long)s0.n + s1.n;
    if (total <= 0) return;
    unsigned int blocks = (unsigned int)((total + 255) / 256);
    sweep_all_pairs_kernel<<<blocks, 256, 0, stream>>>(sc, s0, s1, exact_counter);
}

} // namespace rt
