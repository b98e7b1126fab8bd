// sweep.cu -- the hot kernel: nearest hit of every ray against every shape.
//
// Replaces Scene::intersect's linear scan (reference src/scene/mod.rs:98-116, called from
// src/render.rs:47 and, once per light, src/scene/mod.rs:193).  All-pairs, like an N-body
// sweep:
//   phase 1  FILTER   each thread keeps R rays in registers and streams the bounding-sphere
//                     table (one float4 per shape) through shared memory; a 13-instruction
//                     FMA test decides "this ray certainly misses this shape".  The test is
//                     conservative: margins cover both its own rounding and the
//                     reference's, so it only ever rejects pairs the exact
//                     test would reject too.
//   phase 2  EXACT    surviving (ray, shape) pairs go to a per-warp queue in shared memory;
//                     whenever 32 are waiting the warp runs 32 exact tests, one per lane,
//                     with the reference's own operand order (exact.cuh), and folds the
//                     result into the ray's 64-bit (t, id) key with an integer atomicMin.
//   phase 3  PLANES   infinite planes have no bound; every ray tests them exactly.
// The winner is therefore decided by the exact arithmetic alone, in every case.
#include "kernels.h"

namespace rt {

constexpr int SWEEP_BLOCK = 256;
constexpr int SWEEP_WARPS = SWEEP_BLOCK / 32;
constexpr int TILE = 1024;  // bounds per shared-memory tile (16 KB)
constexpr int CHUNK = 128;  // bounds between two queue checks (== BOUND_PAD)
constexpr int QCAP = 512;   // queued pairs per warp
static_assert(CHUNK == BOUND_PAD, "bound table padding must match the sweep chunk");

// conservative-filter margin: 128 units of f32 round-off (2^-24) on b^2 and a*|L|^2
constexpr float FILTER_MARGIN = 128.0f * 5.9604645e-08f;

struct RayRegs {
    float ox, oy, oz, dx, dy, dz;
};

RT_DEV bool ray_has_nan(const RayRegs &r) {
    return !(r.ox == r.ox && r.oy == r.oy && r.oz == r.oz && r.dx == r.dx && r.dy == r.dy && r.dz == r.dz);
}

RT_DEV RayRegs load_ray(const RaySeg &s0, const RaySeg &s1, int gi) {
    float4 a, b;
    if (gi < s0.n) { a = s0.a[gi]; b = s0.b[gi]; }
    else { a = s1.a[gi - s0.n]; b = s1.b[gi - s0.n]; }
    RayRegs r;
    r.ox = a.x; r.oy = a.y; r.oz = a.z; r.dx = a.w; r.dy = b.x; r.dz = b.y;
    return r;
}

// One exact (ray, shape) test; returns the key or KEY_MISS.
RT_DEV unsigned long long exact_pair(const DevScene &sc, int shape, V3 o, V3 d) {
    int kind = sc.shape_meta[shape].x;
    const float4 *g = sc.geom + 3 * shape;
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
        if (!exact_cube(load_m34(g), sc.cube_tris, o, d, t, which, u, v, det)) return KEY_MISS;
    }
    return make_key(t, shape);
}

RT_DEV unsigned long long plane_keys(const DevScene &sc, V3 o, V3 d, unsigned long long best) {
    for (int p = 0; p < sc.n_planes; ++p) {
        int shape = sc.plane_shape[p];
        const float4 *pd = sc.plane_data + 7 * p;
        float4 po = pd[0], pn = pd[1];
        float t;
        if (exact_plane(load_m34(sc.geom + 3 * shape), v3(po.x, po.y, po.z), v3(pn.x, pn.y, pn.z), o, d, t)) {
            unsigned long long k = make_key(t, shape);
            best = k < best ? k : best;
        }
    }
    return best;
}

template <int R>
__global__ void __launch_bounds__(SWEEP_BLOCK)
sweep_kernel(DevScene sc, RaySeg s0, RaySeg s1, int n_bound_padded, unsigned long long *exact_counter) {
    __shared__ float4 tile[TILE];
    __shared__ unsigned int queue[SWEEP_WARPS][QCAP];
    __shared__ unsigned long long best[SWEEP_WARPS][32 * R];
    __shared__ int qcount[SWEEP_WARPS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int total = s0.n + s1.n;
    const int block_ray0 = blockIdx.x * R * SWEEP_BLOCK;

    // ---- rays into registers --------------------------------------------------------
    float ox[R], oy[R], oz[R], dx[R], dy[R], dz[R], a1[R], a2[R];
    bool live[R], nanray[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        int gi = block_ray0 + r * SWEEP_BLOCK + tid;
        live[r] = gi < total;
        nanray[r] = false;
        RayRegs rr;
        if (live[r]) {
            rr = load_ray(s0, s1, gi);
            nanray[r] = ray_has_nan(rr);
        }
        if (!live[r] || nanray[r]) {
            // dummy ray that passes no bound: far origin, zero direction
            ox[r] = oy[r] = oz[r] = 1e18f;
            dx[r] = dy[r] = dz[r] = 0.f;
            a1[r] = 1.f; a2[r] = 0.f;
        } else {
            ox[r] = rr.ox; oy[r] = rr.oy; oz[r] = rr.oz; dx[r] = rr.dx; dy[r] = rr.dy; dz[r] = rr.dz;
            float a = fmaf(rr.dx, rr.dx, fmaf(rr.dy, rr.dy, rr.dz * rr.dz));
            a1[r] = a * ((1.f - FILTER_MARGIN) / (1.f + FILTER_MARGIN));
            a2[r] = a; // a / (1+m) rounded up: errs on the passing side
        }
        best[warp][r * 32 + lane] = KEY_MISS;
    }
    if (lane == 0) qcount[warp] = 0;
    __syncwarp();

    unsigned int n_exact = 0;

    // one queued pair: (local ray << 24) | bound index
    auto run_pair = [&](unsigned int pair) {
        int local = pair >> 24;
        int j = pair & 0xFFFFFF;
        int gi = block_ray0 + (local >> 5) * SWEEP_BLOCK + warp * 32 + (local & 31);
        RayRegs rr = load_ray(s0, s1, gi);
        int shape = sc.bound_shape[j];
        unsigned long long k = exact_pair(sc, shape, v3(rr.ox, rr.oy, rr.oz), v3(rr.dx, rr.dy, rr.dz));
        if (k != KEY_MISS) atomicMin(&best[warp][local], k);
    };

    // run queued pairs 32 at a time; keep a partial batch unless `all`
    auto flush = [&](bool all) {
        __syncwarp();
        int cnt = min(*(volatile int *)&qcount[warp], QCAP);
        int done = 0;
        while (cnt - done >= 32 || (all && done < cnt)) {
            int idx = done + lane;
            if (idx < cnt) { run_pair(queue[warp][idx]); ++n_exact; }
            done += 32;
        }
        int left = cnt - done;
        unsigned int keep = 0;
        if (left > 0 && lane < left) keep = queue[warp][done + lane];
        __syncwarp();
        if (left > 0 && lane < left) queue[warp][lane] = keep;
        if (lane == 0) qcount[warp] = left > 0 ? left : 0;
        __syncwarp();
    };

    // ---- phase 1 + 2 ------------------------------------------------------------------
    for (int base = 0; base < n_bound_padded; base += TILE) {
        __syncthreads();
        for (int k = tid; k < TILE; k += SWEEP_BLOCK)
            if (base + k < n_bound_padded) tile[k] = sc.bound[base + k];
        __syncthreads();
        const int tile_n = min(TILE, n_bound_padded - base);
        for (int c0 = 0; c0 < tile_n; c0 += CHUNK) {
#pragma unroll 4
            for (int k = c0; k < c0 + CHUNK; ++k) {
                const float4 B = tile[k];
                bool pass[R];
                bool any = false;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float lx = B.x - ox[r], ly = B.y - oy[r], lz = B.z - oz[r];
                    float b = fmaf(lz, dz[r], fmaf(ly, dy[r], lx * dx[r]));
                    float ll = fmaf(lz, lz, fmaf(ly, ly, lx * lx));
                    // b^2 (1+m) - a |L|^2 (1-m) + a R^2 >= 0, divided through by (1+m)
                    float t = fmaf(a2[r], B.w, fmaf(-a1[r], ll, b * b));
                    pass[r] = t >= 0.f;
                    any |= pass[r];
                }
                if (any) {
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        if (pass[r]) {
                            unsigned int pair = ((unsigned int)(r * 32 + lane) << 24) | (unsigned int)(base + k);
                            int slot = atomicAdd(&qcount[warp], 1);
                            if (slot < QCAP) queue[warp][slot] = pair;
                            else { run_pair(pair); ++n_exact; } // queue full: resolve in place
                        }
                    }
                }
            }
            __syncwarp();
            if (*(volatile int *)&qcount[warp] >= 32) flush(false);
        }
    }
    flush(true);

    // ---- phase 3 + output ---------------------------------------------------------------
#pragma unroll
    for (int r = 0; r < R; ++r) {
        int gi = block_ray0 + r * SWEEP_BLOCK + tid;
        if (!live[r]) continue;
        unsigned long long k;
        if (nanray[r]) {
            k = sc.first_nan_shape >= 0
                    ? (((unsigned long long)KEY_NAN_PREFIX << 32) | (unsigned long long)(uint32_t)sc.first_nan_shape)
                    : KEY_MISS;
        } else {
            k = best[warp][r * 32 + lane];
            k = plane_keys(sc, v3(ox[r], oy[r], oz[r]), v3(dx[r], dy[r], dz[r]), k);
        }
        if (gi < s0.n) s0.key[gi] = k;
        else s1.key[gi - s0.n] = k;
    }
    if (exact_counter) {
        unsigned int tot = n_exact;
        for (int off = 16; off > 0; off >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, off);
        if (lane == 0 && tot) atomicAdd(exact_counter, (unsigned long long)tot);
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
        for (int j = 0; j < sc.n_bounded; ++j) {
            unsigned long long k = exact_pair(sc, sc.bound_shape[j], o, d);
            best = k < best ? k : best;
        }
        best = plane_keys(sc, o, d, best);
    }
    if (gi < s0.n) s0.key[gi] = best;
    else s1.key[gi - s0.n] = best;
    (void)exact_counter; // rays x n_bounded, added by the host
}

static int padded(int n) { return (n + BOUND_PAD - 1) / BOUND_PAD * BOUND_PAD; }

void launch_sweep(const DevScene &sc, RaySeg s0, RaySeg s1, int rays_per_thread, unsigned long long *exact_counter,
                  cudaStream_t stream) {
    long long total = (long long)s0.n + s1.n;
    if (total <= 0) return;
    int nb = padded(sc.n_bounded);
    auto blocks = [&](int r) { return (unsigned int)((total + (long long)r * SWEEP_BLOCK - 1) / ((long long)r * SWEEP_BLOCK)); };
    switch (rays_per_thread) {
    case 1: sweep_kernel<1><<<blocks(1), SWEEP_BLOCK, 0, stream>>>(sc, s0, s1, nb, exact_counter); break;
    case 2: sweep_kernel<2><<<blocks(2), SWEEP_BLOCK, 0, stream>>>(sc, s0, s1, nb, exact_counter); break;
    default: sweep_kernel<4><<<blocks(4), SWEEP_BLOCK, 0, stream>>>(sc, s0, s1, nb, exact_counter); break;
    }
}

void launch_sweep_all_pairs(const DevScene &sc, RaySeg s0, RaySeg s1, unsigned long long *exact_counter,
                            cudaStream_t stream) {
    long long total = (long long)s0.n + s1.n;
    if (total <= 0) return;
    unsigned int blocks = (unsigned int)((total + 255) / 256);
    sweep_all_pairs_kernel<<<blocks, 256, 0, stream>>>(sc, s0, s1, exact_counter);
}

} // namespace rt
