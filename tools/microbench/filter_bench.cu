// filter_bench.cu -- developer micro-benchmarks behind the sweep kernel's design
// ("sweep filter").  Standalone: nvcc -gencode arch=compute_100a,code=sm_100a.
//   1. issue-rate probes: FFMA, packed FFMA2 (fma.rn.f32x2), FFMA+FSETP mixes
//   2. candidate (ray x bounding-sphere) reject filters on a synthetic 10k-bound table:
//        A   project the centre on a per-ray basis (u, v) of the plane normal to the ray
//        B   expanded quadratic  b^2 + 2 C.O + (R^2 - |C|^2) >= |O|^2 - mu
//        A2 / B2  the same with two bounds per packed FFMA2
//      each checked against a double-precision distance test: a filter may pass too much,
//      never too little.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

static const float U24 = 5.9604645e-08f; // 2^-24

// ------------------------------------------------------------------ issue-rate probes
template <int MODE>
__global__ void __launch_bounds__(256) probe_kernel(int iters, float *sink, float a, float b) {
    float acc[16];
    float2 acc2[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) { acc[c] = (float)c + threadIdx.x * 1e-9f; acc2[c] = make_float2(acc[c], acc[c] + 1.f); }
    float2 a2 = make_float2(a, a * 1.0000001f), b2 = make_float2(b, b * 0.5f);
    int cnt = 0;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (MODE == 0) {
#pragma unroll
                for (int c = 0; c < 16; ++c) acc[c] = fmaf(acc[c], a, b);
            } else if (MODE == 1) {
#pragma unroll
                for (int c = 0; c < 16; ++c) acc2[c] = __ffma2_rn(acc2[c], a2, b2);
            } else if (MODE == 2) { // 8 FFMA : 1 FSETP (+ predicate combine), branch every 16
                bool any = false;
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    acc[c] = fmaf(acc[c], a, b);
                    if ((c & 7) == 7) any |= acc[c] > 1e30f;
                }
                if (any) cnt++;
            } else if (MODE == 3) { // 4 FFMA2 : 2 FSETP
                bool any = false;
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    acc2[c] = __ffma2_rn(acc2[c], a2, b2);
                    if ((c & 3) == 3) any |= (acc2[c].x > 1e30f) | (acc2[c].y > 1e30f);
                }
                if (any) cnt++;
            }
        }
    }
    float s = (float)cnt;
#pragma unroll
    for (int c = 0; c < 16; ++c) s += acc[c] + acc2[c].x + acc2[c].y;
    if (s == 123.456f) sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ------------------------------------------------------------------ filter candidates
struct RayA { float ux, uy, uz, ku, vx, vy, vz, kv, nmu; };
struct RayB { float dx, dy, dz, kb, ox2, oy2, oz2, thr; };

__device__ __forceinline__ void basis(float dx, float dy, float dz, float &ux, float &uy, float &uz, float &vx, float &vy,
                                      float &vz) {
    // u = norm(d x e_k), e_k the axis least aligned with d;  v = d x u
    float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
    if (ax <= ay && ax <= az) { ux = 0.f; uy = dz; uz = -dy; }
    else if (ay <= az) { ux = -dz; uy = 0.f; uz = dx; }
    else { ux = dy; uy = -dx; uz = 0.f; }
    float n = __fsqrt_rn(ux * ux + uy * uy + uz * uz);
    ux = __fdiv_rn(ux, n); uy = __fdiv_rn(uy, n); uz = __fdiv_rn(uz, n);
    vx = dy * uz - dz * uy; vy = dz * ux - dx * uz; vz = dx * uy - dy * ux;
}

__device__ __forceinline__ RayA make_rayA(float4 a, float4 b, float cmax) {
    float ox = a.x, oy = a.y, oz = a.z, dx = a.w, dy = b.x, dz = b.y;
    float n = __fsqrt_rn(dx * dx + dy * dy + dz * dz);
    dx = __fdiv_rn(dx, n); dy = __fdiv_rn(dy, n); dz = __fdiv_rn(dz, n);
    RayA r;
    basis(dx, dy, dz, r.ux, r.uy, r.uz, r.vx, r.vy, r.vz);
    r.ku = -(ox * r.ux + oy * r.uy + oz * r.uz);
    r.kv = -(ox * r.vx + oy * r.vy + oz * r.vz);
    float on = sqrtf(ox * ox + oy * oy + oz * oz);
    float lmax = on + cmax;
    float eps = U24 * (3.f * cmax + 6.f * on + 12.f * lmax);
    r.nmu = -2.f * (15.f * U24 * lmax * lmax + 2002.f * eps * eps);
    return r;
}
__device__ __forceinline__ RayB make_rayB(float4 a, float4 b, float cmax) {
    float ox = a.x, oy = a.y, oz = a.z, dx = a.w, dy = b.x, dz = b.y;
    float n = __fsqrt_rn(dx * dx + dy * dy + dz * dz);
    RayB r;
    r.dx = __fdiv_rn(dx, n); r.dy = __fdiv_rn(dy, n); r.dz = __fdiv_rn(dz, n);
    r.kb = -(ox * r.dx + oy * r.dy + oz * r.dz);
    r.ox2 = 2.f * ox; r.oy2 = 2.f * oy; r.oz2 = 2.f * oz;
    float o2 = ox * ox + oy * oy + oz * oz;
    float lmax = sqrtf(o2) + cmax;
    r.thr = 2.f * 55.f * U24 * lmax * lmax - o2; // seed of the s chain: mu - |O|^2
    return r;
}

constexpr int QCAP = 512;
constexpr int CHUNK = 64;

// VARIANT 0 = A, 1 = B, 2 = A2 (two bounds per FFMA2), 3 = B2
template <int VARIANT, int R, int THREADS>
__global__ void __launch_bounds__(THREADS) filter_kernel(const float4 *__restrict__ table, int n_bounds, const float4 *ray_a,
                                                         const float4 *ray_b, int n_rays, float cmax, int *work_counter,
                                                         unsigned long long *pass_total, unsigned char *pass_bits) {
    extern __shared__ float4 smem[];
    constexpr int WARPS = THREADS / 32;
    float4 *tab = smem;                                          // n_bounds float4
    unsigned int *queue = (unsigned int *)(smem + n_bounds);     // WARPS * QCAP
    int *qcount = (int *)(queue + WARPS * QCAP);                 // WARPS
    __shared__ int batch_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int k = tid; k < n_bounds; k += THREADS) tab[k] = table[k];
    if (lane == 0) qcount[warp] = 0;
    __syncthreads();
    unsigned long long passes = 0;
    for (;;) {
        if (tid == 0) batch_s = atomicAdd(work_counter, 1);
        __syncthreads();
        const int batch = batch_s;
        __syncthreads();
        const int ray0 = batch * (R * THREADS);
        if (ray0 >= n_rays) break;
        RayA ra[R];
        RayB rb[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            int gi = ray0 + r * THREADS + tid;
            float4 a = make_float4(1e18f, 1e18f, 1e18f, 1.f), b = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gi < n_rays) { a = ray_a[gi]; b = ray_b[gi]; }
            if (VARIANT == 0 || VARIANT == 2) ra[r] = make_rayA(a, b, cmax);
            else rb[r] = make_rayB(a, b, cmax);
        }
        auto push = [&](int r, int j) {
            unsigned int pair = ((unsigned int)(r * 32 + lane) << 24) | (unsigned int)j;
            int slot = atomicAdd(&qcount[warp], 1);
            if (slot < QCAP) queue[warp * QCAP + slot] = pair;
            if (pass_bits) {
                int gi = ray0 + r * THREADS + tid;
                if (gi < n_rays) pass_bits[(size_t)gi * n_bounds + j] = 1;
            }
        };
        for (int c0 = 0; c0 < n_bounds; c0 += CHUNK) {
            if (VARIANT == 0) {
#pragma unroll 4
                for (int k = c0; k < c0 + CHUNK; ++k) {
                    const float4 B = tab[k];
                    bool pass[R], any = false;
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        float p = fmaf(B.x, ra[r].ux, fmaf(B.y, ra[r].uy, fmaf(B.z, ra[r].uz, ra[r].ku)));
                        float q = fmaf(B.x, ra[r].vx, fmaf(B.y, ra[r].vy, fmaf(B.z, ra[r].vz, ra[r].kv)));
                        float s = fmaf(q, q, fmaf(p, p, ra[r].nmu));
                        pass[r] = s <= B.w;
                        any |= pass[r];
                    }
                    if (any) {
#pragma unroll
                        for (int r = 0; r < R; ++r) if (pass[r]) push(r, k);
                    }
                }
            } else if (VARIANT == 1) {
#pragma unroll 4
                for (int k = c0; k < c0 + CHUNK; ++k) {
                    const float4 B = tab[k];
                    bool pass[R], any = false;
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        float b = fmaf(B.x, rb[r].dx, fmaf(B.y, rb[r].dy, fmaf(B.z, rb[r].dz, rb[r].kb)));
                        float s = fmaf(B.x, rb[r].ox2, fmaf(B.y, rb[r].oy2, fmaf(B.z, rb[r].oz2, rb[r].thr)));
                        float e = fmaf(b, b, s);
                        pass[r] = e >= B.w;
                        any |= pass[r];
                    }
                    if (any) {
#pragma unroll
                        for (int r = 0; r < R; ++r) if (pass[r]) push(r, k);
                    }
                }
            } else if (VARIANT == 2) {
#pragma unroll 2
                for (int k = c0; k < c0 + CHUNK; k += 2) {
                    const float4 T0 = tab[k], T1 = tab[k + 1]; // (cx0,cx1,cy0,cy1) (cz0,cz1,W0,W1)
                    const float2 cx = make_float2(T0.x, T0.y), cy = make_float2(T0.z, T0.w), cz = make_float2(T1.x, T1.y);
                    bool p0[R], p1[R], any = false;
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        float2 ux = make_float2(ra[r].ux, ra[r].ux), uy = make_float2(ra[r].uy, ra[r].uy),
                               uz = make_float2(ra[r].uz, ra[r].uz), ku = make_float2(ra[r].ku, ra[r].ku);
                        float2 vx = make_float2(ra[r].vx, ra[r].vx), vy = make_float2(ra[r].vy, ra[r].vy),
                               vz = make_float2(ra[r].vz, ra[r].vz), kv = make_float2(ra[r].kv, ra[r].kv);
                        float2 nmu = make_float2(ra[r].nmu, ra[r].nmu);
                        float2 p = __ffma2_rn(cx, ux, __ffma2_rn(cy, uy, __ffma2_rn(cz, uz, ku)));
                        float2 q = __ffma2_rn(cx, vx, __ffma2_rn(cy, vy, __ffma2_rn(cz, vz, kv)));
                        float2 s = __ffma2_rn(q, q, __ffma2_rn(p, p, nmu));
                        p0[r] = s.x <= T1.z; p1[r] = s.y <= T1.w;
                        any |= p0[r] | p1[r];
                    }
                    if (any) {
#pragma unroll
                        for (int r = 0; r < R; ++r) { if (p0[r]) push(r, k); if (p1[r]) push(r, k + 1); }
                    }
                }
            } else {
#pragma unroll 2
                for (int k = c0; k < c0 + CHUNK; k += 2) {
                    const float4 T0 = tab[k], T1 = tab[k + 1]; // (cx0,cx1,cy0,cy1) (cz0,cz1,nw0,nw1)
                    const float2 cx = make_float2(T0.x, T0.y), cy = make_float2(T0.z, T0.w), cz = make_float2(T1.x, T1.y);
                    bool p0[R], p1[R], any = false;
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        float2 dx = make_float2(rb[r].dx, rb[r].dx), dy = make_float2(rb[r].dy, rb[r].dy),
                               dz = make_float2(rb[r].dz, rb[r].dz), kb = make_float2(rb[r].kb, rb[r].kb);
                        float2 ox = make_float2(rb[r].ox2, rb[r].ox2), oy = make_float2(rb[r].oy2, rb[r].oy2),
                               oz = make_float2(rb[r].oz2, rb[r].oz2), sd = make_float2(rb[r].thr, rb[r].thr);
                        float2 b = __ffma2_rn(cx, dx, __ffma2_rn(cy, dy, __ffma2_rn(cz, dz, kb)));
                        float2 s = __ffma2_rn(cx, ox, __ffma2_rn(cy, oy, __ffma2_rn(cz, oz, sd)));
                        float2 e = __ffma2_rn(b, b, s);
                        p0[r] = e.x >= T1.z; p1[r] = e.y >= T1.w;
                        any |= p0[r] | p1[r];
                    }
                    if (any) {
#pragma unroll
                        for (int r = 0; r < R; ++r) { if (p0[r]) push(r, k); if (p1[r]) push(r, k + 1); }
                    }
                }
            }
            __syncwarp();
            int qc = *(volatile int *)&qcount[warp];
            if (qc >= 32) { // stand-in for the exact batch: drain
                if (lane == 0) { passes += (unsigned long long)qc; qcount[warp] = 0; }
                __syncwarp();
            }
        }
    }
    __syncwarp();
    if (lane == 0) { passes += (unsigned long long)qcount[warp]; atomicAdd(pass_total, passes); }
}


// B2 refined: per-warp work fetch (no block barriers after the table load), optional
// register prefetch of the next bound pair, unroll factor U.
template <int R, int THREADS, int U, int PREFETCH>
__global__ void __launch_bounds__(THREADS, 1) filter_b2w_kernel(const float4 *__restrict__ table, int n_bounds, const float4 *ray_a,
                                                             const float4 *ray_b, int n_rays, float cmax, int *work_counter,
                                                             unsigned long long *pass_total, unsigned char *pass_bits) {
    extern __shared__ float4 smem[];
    constexpr int WARPS = THREADS / 32;
    float4 *tab = smem;
    unsigned int *queue = (unsigned int *)(smem + n_bounds);
    int *qcount = (int *)(queue + WARPS * QCAP);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int k = tid; k < n_bounds; k += THREADS) tab[k] = table[k];
    if (lane == 0) qcount[warp] = 0;
    __syncthreads();
    unsigned long long passes = 0;
    for (;;) {
        int batch = 0;
        if (lane == 0) batch = atomicAdd(work_counter, 1);
        batch = __shfl_sync(0xffffffffu, batch, 0);
        const int ray0 = batch * (R * 32);
        if (ray0 >= n_rays) break;
        RayB rb[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            int gi = ray0 + r * 32 + lane;
            float4 a = make_float4(1e18f, 1e18f, 1e18f, 1.f), b = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gi < n_rays) { a = ray_a[gi]; b = ray_b[gi]; }
            rb[r] = make_rayB(a, b, cmax);
        }
        auto push = [&](int r, int j) {
            unsigned int pair = ((unsigned int)(r * 32 + lane) << 24) | (unsigned int)j;
            int slot = atomicAdd(&qcount[warp], 1);
            if (slot < QCAP) queue[warp * QCAP + slot] = pair;
            if (pass_bits) {
                int gi = ray0 + r * 32 + lane;
                if (gi < n_rays) pass_bits[(size_t)gi * n_bounds + j] = 1;
            }
        };
        float4 N0 = tab[0], N1 = tab[1];
        for (int c0 = 0; c0 < n_bounds; c0 += CHUNK) {
#pragma unroll U
            for (int k = c0; k < c0 + CHUNK; k += 2) {
                float4 T0, T1;
                if (PREFETCH) {
                    T0 = N0; T1 = N1;
                    int kn = (k + 2 < n_bounds) ? k + 2 : 0;
                    N0 = tab[kn]; N1 = tab[kn + 1];
                } else { T0 = tab[k]; T1 = tab[k + 1]; }
                const float2 cx = make_float2(T0.x, T0.y), cy = make_float2(T0.z, T0.w), cz = make_float2(T1.x, T1.y);
                float2 e[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float2 dx = make_float2(rb[r].dx, rb[r].dx), dy = make_float2(rb[r].dy, rb[r].dy),
                           dz = make_float2(rb[r].dz, rb[r].dz), kb = make_float2(rb[r].kb, rb[r].kb);
                    float2 ox = make_float2(rb[r].ox2, rb[r].ox2), oy = make_float2(rb[r].oy2, rb[r].oy2),
                           oz = make_float2(rb[r].oz2, rb[r].oz2), sd = make_float2(rb[r].thr, rb[r].thr);
                    float2 b = __ffma2_rn(cx, dx, __ffma2_rn(cy, dy, __ffma2_rn(cz, dz, kb)));
                    float2 s = __ffma2_rn(cx, ox, __ffma2_rn(cy, oy, __ffma2_rn(cz, oz, sd)));
                    e[r] = __ffma2_rn(b, b, s);
                }
                float m0 = e[0].x, m1 = e[0].y;
#pragma unroll
                for (int r = 1; r < R; ++r) { m0 = fmaxf(m0, e[r].x); m1 = fmaxf(m1, e[r].y); }
                if (m0 >= T1.z || m1 >= T1.w) {
#pragma unroll
                    for (int r = 0; r < R; ++r) { if (e[r].x >= T1.z) push(r, k); if (e[r].y >= T1.w) push(r, k + 1); }
                }
            }
            __syncwarp();
            int qc = *(volatile int *)&qcount[warp];
            if (qc >= 32) {
                if (lane == 0) { passes += (unsigned long long)qc; qcount[warp] = 0; }
                __syncwarp();
            }
        }
    }
    __syncwarp();
    if (lane == 0) { passes += (unsigned long long)qcount[warp]; atomicAdd(pass_total, passes); }
}


// B2 with the bound table in __constant__ memory: uniform loads (LDCU) put the bound pair in
// uniform registers, so each FFMA2 reads only per-lane registers from the register file.
#define CTAB_SLOTS 3968
__constant__ float4 c_tab[CTAB_SLOTS];

template <int R, int THREADS, int U>
__global__ void __launch_bounds__(THREADS, 1) filter_b2s_kernel(int n_bounds, const float4 *ray_a, const float4 *ray_b, int n_rays,
                                                             float cmax, int *work_counter, unsigned long long *pass_total) {
    constexpr int WARPS = THREADS / 32;
    __shared__ unsigned int queue[WARPS * QCAP];
    __shared__ int qcount[WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (lane == 0) qcount[warp] = 0;
    __syncwarp();
    unsigned long long passes = 0;
    const int n_batches = (n_rays + R * THREADS - 1) / (R * THREADS);
    for (int batch = blockIdx.x; batch < n_batches; batch += gridDim.x) {
        const int ray0 = batch * (R * THREADS);
        RayB rb[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            int gi = ray0 + r * THREADS + tid;
            float4 a = make_float4(1e18f, 1e18f, 1e18f, 1.f), b = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gi < n_rays) { a = ray_a[gi]; b = ray_b[gi]; }
            rb[r] = make_rayB(a, b, cmax);
        }
        auto push = [&](int r, int j) {
            unsigned int pair = ((unsigned int)(r * 32 + lane) << 24) | (unsigned int)j;
            int slot = atomicAdd(&qcount[warp], 1);
            if (slot < QCAP) queue[warp * QCAP + slot] = pair;
        };
        for (int c0 = 0; c0 < n_bounds; c0 += CHUNK) {
#pragma unroll U
            for (int k = c0; k < c0 + CHUNK; k += 2) {
                const float4 T0 = c_tab[k], T1 = c_tab[k + 1];
                const float2 cx = make_float2(T0.x, T0.y), cy = make_float2(T0.z, T0.w), cz = make_float2(T1.x, T1.y);
                float2 e[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float2 dx = make_float2(rb[r].dx, rb[r].dx), dy = make_float2(rb[r].dy, rb[r].dy),
                           dz = make_float2(rb[r].dz, rb[r].dz), kb = make_float2(rb[r].kb, rb[r].kb);
                    float2 ox = make_float2(rb[r].ox2, rb[r].ox2), oy = make_float2(rb[r].oy2, rb[r].oy2),
                           oz = make_float2(rb[r].oz2, rb[r].oz2), sd = make_float2(rb[r].thr, rb[r].thr);
                    float2 b = __ffma2_rn(cx, dx, __ffma2_rn(cy, dy, __ffma2_rn(cz, dz, kb)));
                    float2 s = __ffma2_rn(cx, ox, __ffma2_rn(cy, oy, __ffma2_rn(cz, oz, sd)));
                    e[r] = __ffma2_rn(b, b, s);
                }
                float m0 = e[0].x, m1 = e[0].y;
#pragma unroll
                for (int r = 1; r < R; ++r) { m0 = fmaxf(m0, e[r].x); m1 = fmaxf(m1, e[r].y); }
                if (m0 >= T1.z || m1 >= T1.w) {
#pragma unroll
                    for (int r = 0; r < R; ++r) { if (e[r].x >= T1.z) push(r, k); if (e[r].y >= T1.w) push(r, k + 1); }
                }
            }
            __syncwarp();
            int qc = *(volatile int *)&qcount[warp];
            if (qc >= 32) {
                if (lane == 0) { passes += (unsigned long long)qc; qcount[warp] = 0; }
                __syncwarp();
            }
        }
    }
    __syncwarp();
    if (lane == 0) { passes += (unsigned long long)qcount[warp]; atomicAdd(pass_total, passes); }
}

template <int R, int THREADS, int U>
__global__ void __launch_bounds__(THREADS, 1) filter_b2c_kernel(int n_bounds, const float4 *ray_a, const float4 *ray_b, int n_rays,
                                                             float cmax, int *work_counter, unsigned long long *pass_total) {
    constexpr int WARPS = THREADS / 32;
    __shared__ unsigned int queue[WARPS * QCAP];
    __shared__ int qcount[WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (lane == 0) qcount[warp] = 0;
    __syncwarp();
    unsigned long long passes = 0;
    for (;;) {
        int batch = 0;
        if (lane == 0) batch = atomicAdd(work_counter, 1);
        batch = __shfl_sync(0xffffffffu, batch, 0);
        const int ray0 = batch * (R * 32);
        if (ray0 >= n_rays) break;
        RayB rb[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            int gi = ray0 + r * 32 + lane;
            float4 a = make_float4(1e18f, 1e18f, 1e18f, 1.f), b = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gi < n_rays) { a = ray_a[gi]; b = ray_b[gi]; }
            rb[r] = make_rayB(a, b, cmax);
        }
        auto push = [&](int r, int j) {
            unsigned int pair = ((unsigned int)(r * 32 + lane) << 24) | (unsigned int)j;
            int slot = atomicAdd(&qcount[warp], 1);
            if (slot < QCAP) queue[warp * QCAP + slot] = pair;
        };
        for (int c0 = 0; c0 < n_bounds; c0 += CHUNK) {
#pragma unroll U
            for (int k = c0; k < c0 + CHUNK; k += 2) {
                const float4 T0 = c_tab[k], T1 = c_tab[k + 1];
                const float2 cx = make_float2(T0.x, T0.y), cy = make_float2(T0.z, T0.w), cz = make_float2(T1.x, T1.y);
                float2 e[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float2 dx = make_float2(rb[r].dx, rb[r].dx), dy = make_float2(rb[r].dy, rb[r].dy),
                           dz = make_float2(rb[r].dz, rb[r].dz), kb = make_float2(rb[r].kb, rb[r].kb);
                    float2 ox = make_float2(rb[r].ox2, rb[r].ox2), oy = make_float2(rb[r].oy2, rb[r].oy2),
                           oz = make_float2(rb[r].oz2, rb[r].oz2), sd = make_float2(rb[r].thr, rb[r].thr);
                    float2 b = __ffma2_rn(cx, dx, __ffma2_rn(cy, dy, __ffma2_rn(cz, dz, kb)));
                    float2 s = __ffma2_rn(cx, ox, __ffma2_rn(cy, oy, __ffma2_rn(cz, oz, sd)));
                    e[r] = __ffma2_rn(b, b, s);
                }
                float m0 = e[0].x, m1 = e[0].y;
#pragma unroll
                for (int r = 1; r < R; ++r) { m0 = fmaxf(m0, e[r].x); m1 = fmaxf(m1, e[r].y); }
                if (m0 >= T1.z || m1 >= T1.w) {
#pragma unroll
                    for (int r = 0; r < R; ++r) { if (e[r].x >= T1.z) push(r, k); if (e[r].y >= T1.w) push(r, k + 1); }
                }
            }
            __syncwarp();
            int qc = *(volatile int *)&qcount[warp];
            if (qc >= 32) {
                if (lane == 0) { passes += (unsigned long long)qc; qcount[warp] = 0; }
                __syncwarp();
            }
        }
    }
    __syncwarp();
    if (lane == 0) { passes += (unsigned long long)qcount[warp]; atomicAdd(pass_total, passes); }
}

// register-file probes: FFMA2 with (a) a rotating register pair, (b) a kernel-parameter pair
template <int MODE>
__global__ void __launch_bounds__(256) rf_probe_kernel(int iters, float *sink, float2 p0, float2 p1, float2 p2, float2 p3) {
    float2 acc[8];
    float sc[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) { acc[c] = make_float2((float)c + threadIdx.x * 1e-9f, 1.f + c); sc[c] = 0.999f + 1e-4f * c + threadIdx.x * 1e-9f; }
    float2 q0 = p0, q1 = p1, q2 = p2, q3 = p3;
    if (MODE == 0) { // force the pairs into vector registers
        q0.x += threadIdx.x * 1e-12f; q1.x += threadIdx.x * 1e-12f; q2.x += threadIdx.x * 1e-12f; q3.x += threadIdx.x * 1e-12f;
    }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 2; ++u) {
#pragma unroll
            for (int c = 0; c < 8; ++c) acc[c] = __ffma2_rn(q0, make_float2(sc[c], sc[c]), acc[c]);
#pragma unroll
            for (int c = 0; c < 8; ++c) acc[c] = __ffma2_rn(q1, make_float2(sc[c], sc[c]), acc[c]);
#pragma unroll
            for (int c = 0; c < 8; ++c) acc[c] = __ffma2_rn(q2, make_float2(sc[c], sc[c]), acc[c]);
#pragma unroll
            for (int c = 0; c < 8; ++c) acc[c] = __ffma2_rn(q3, make_float2(sc[c], sc[c]), acc[c]);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < 8; ++c) s += acc[c].x + acc[c].y;
    if (s == 123.456f) sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ------------------------------------------------------------------ host
static uint64_t sm64(uint64_t &s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static float u01(uint64_t &s) { return (float)(sm64(s) >> 40) * (1.0f / 16777216.0f); }

struct Bound { double c[3]; double r; };

template <int VARIANT, int R, int THREADS>
static void run_filter(const char *name, const std::vector<Bound> &bounds, const float4 *d_ra, const float4 *d_rb, int n_rays,
                       const std::vector<float4> &h_ra, const std::vector<float4> &h_rb, int n_sm) {
    const int nb = (int)bounds.size();
    std::vector<float4> tab(nb);
    double cmax = 0;
    for (auto &b : bounds) cmax = fmax(cmax, sqrt(b.c[0] * b.c[0] + b.c[1] * b.c[1] + b.c[2] * b.c[2]));
    auto up = [](double x) { float f = (float)x; if ((double)f < x) f = nextafterf(f, INFINITY); return f; };
    for (int j = 0; j < nb; ++j) {
        const Bound &b = bounds[j];
        float cx = (float)b.c[0], cy = (float)b.c[1], cz = (float)b.c[2];
        double dx = b.c[0] - cx, dy = b.c[1] - cy, dz = b.c[2] - cz;
        double r = b.r + sqrt(dx * dx + dy * dy + dz * dz);
        if (VARIANT == 0) tab[j] = make_float4(cx, cy, cz, up(r * r * 1.001));
        else if (VARIANT == 1) tab[j] = make_float4(cx, cy, cz, -up(r * r * 1.001 - ((double)cx * cx + (double)cy * cy + (double)cz * cz)));
    }
    if (VARIANT >= 2) {
        for (int j = 0; j < nb; j += 2) {
            float4 t[2];
            for (int k = 0; k < 2; ++k) {
                const Bound &b = bounds[j + k];
                float cx = (float)b.c[0], cy = (float)b.c[1], cz = (float)b.c[2];
                double dx = b.c[0] - cx, dy = b.c[1] - cy, dz = b.c[2] - cz;
                double r = b.r + sqrt(dx * dx + dy * dy + dz * dz);
                float w = VARIANT == 2 ? up(r * r * 1.001) : -up(r * r * 1.001 - ((double)cx * cx + (double)cy * cy + (double)cz * cz));
                t[k] = make_float4(cx, cy, cz, w);
            }
            tab[j] = make_float4(t[0].x, t[1].x, t[0].y, t[1].y);
            tab[j + 1] = make_float4(t[0].z, t[1].z, t[0].w, t[1].w);
        }
    }
    float4 *d_tab;
    int *d_work;
    unsigned long long *d_pass;
    CK(cudaMalloc(&d_tab, nb * sizeof(float4)));
    CK(cudaMemcpy(d_tab, tab.data(), nb * sizeof(float4), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_work, sizeof(int)));
    CK(cudaMalloc(&d_pass, sizeof(unsigned long long)));
    size_t smem = (size_t)nb * 16 + (THREADS / 32) * QCAP * 4 + (THREADS / 32) * 4 + 64;
    auto kern = filter_kernel<VARIANT, R, THREADS>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, smem));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, kern));
    int grid = n_sm * occ;

    // ---- correctness on a small prefix: no false rejects vs double precision
    const int n_chk = 2048;
    unsigned char *d_bits;
    CK(cudaMalloc(&d_bits, (size_t)n_chk * nb));
    CK(cudaMemset(d_bits, 0, (size_t)n_chk * nb));
    CK(cudaMemset(d_work, 0, sizeof(int)));
    CK(cudaMemset(d_pass, 0, sizeof(unsigned long long)));
    kern<<<grid, THREADS, smem>>>(d_tab, nb, d_ra, d_rb, n_chk, (float)cmax, d_work, d_pass, d_bits);
    CK(cudaDeviceSynchronize());
    std::vector<unsigned char> bits((size_t)n_chk * nb);
    CK(cudaMemcpy(bits.data(), d_bits, bits.size(), cudaMemcpyDeviceToHost));
    long long true_hits = 0, filt = 0, false_rej = 0;
    for (int i = 0; i < n_chk; ++i) {
        double o[3] = {h_ra[i].x, h_ra[i].y, h_ra[i].z}, d[3] = {h_ra[i].w, h_rb[i].x, h_rb[i].y};
        double dn = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        for (int j = 0; j < nb; ++j) {
            const Bound &b = bounds[j];
            double L[3] = {b.c[0] - o[0], b.c[1] - o[1], b.c[2] - o[2]};
            double bb = (L[0] * d[0] + L[1] * d[1] + L[2] * d[2]) / dn;
            double dist2 = L[0] * L[0] + L[1] * L[1] + L[2] * L[2] - bb * bb;
            bool t = dist2 <= b.r * b.r;
            bool f = bits[(size_t)i * nb + j] != 0;
            true_hits += t; filt += f;
            if (t && !f) false_rej++;
        }
    }
    // ---- timing
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    unsigned long long passes = 0;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaMemset(d_work, 0, sizeof(int)));
        CK(cudaMemset(d_pass, 0, sizeof(unsigned long long)));
        cudaEventRecord(e0);
        kern<<<grid, THREADS, smem>>>(d_tab, nb, d_ra, d_rb, n_rays, (float)cmax, d_work, d_pass, nullptr);
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
        CK(cudaMemcpy(&passes, d_pass, sizeof(passes), cudaMemcpyDeviceToHost));
    }
    double pairs = (double)n_rays * nb;
    printf("%-28s R=%d thr=%d regs=%d occ=%d grid=%d smem=%zu | %.3f ms  %.3f Tpairs/s | passes/ray %.2f | check: true %lld filt %lld (x%.2f) FALSE_REJECTS %lld\n",
           name, R, THREADS, fa.numRegs, occ, grid, smem, best, pairs / best / 1e9, (double)passes / n_rays, true_hits, filt,
           true_hits ? (double)filt / true_hits : 0.0, false_rej);
    cudaFree(d_tab); cudaFree(d_work); cudaFree(d_pass); cudaFree(d_bits);
}

template <int R, int THREADS, int U, int PREFETCH>
static void run_b2w(const std::vector<Bound> &bounds, const float4 *d_ra, const float4 *d_rb, int n_rays, int n_sm) {
    const int nb = (int)bounds.size();
    std::vector<float4> tab(nb);
    double cmax = 0;
    for (auto &b : bounds) cmax = fmax(cmax, sqrt(b.c[0] * b.c[0] + b.c[1] * b.c[1] + b.c[2] * b.c[2]));
    auto up = [](double x) { float f = (float)x; if ((double)f < x) f = nextafterf(f, INFINITY); return f; };
    for (int j = 0; j < nb; j += 2) {
        float4 t[2];
        for (int k = 0; k < 2; ++k) {
            const Bound &b = bounds[j + k];
            float cx = (float)b.c[0], cy = (float)b.c[1], cz = (float)b.c[2];
            double dx = b.c[0] - cx, dy = b.c[1] - cy, dz = b.c[2] - cz;
            double r = b.r + sqrt(dx * dx + dy * dy + dz * dz);
            t[k] = make_float4(cx, cy, cz, -up(r * r * 1.001 - ((double)cx * cx + (double)cy * cy + (double)cz * cz)));
        }
        tab[j] = make_float4(t[0].x, t[1].x, t[0].y, t[1].y);
        tab[j + 1] = make_float4(t[0].z, t[1].z, t[0].w, t[1].w);
    }
    float4 *d_tab; int *d_work; unsigned long long *d_pass;
    CK(cudaMalloc(&d_tab, nb * sizeof(float4)));
    CK(cudaMemcpy(d_tab, tab.data(), nb * sizeof(float4), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_work, sizeof(int)));
    CK(cudaMalloc(&d_pass, sizeof(unsigned long long)));
    size_t smem = (size_t)nb * 16 + (THREADS / 32) * QCAP * 4 + (THREADS / 32) * 4 + 64;
    auto kern = filter_b2w_kernel<R, THREADS, U, PREFETCH>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, kern));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    unsigned long long passes = 0;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaMemset(d_work, 0, sizeof(int)));
        CK(cudaMemset(d_pass, 0, sizeof(unsigned long long)));
        cudaEventRecord(e0);
        kern<<<n_sm, THREADS, smem>>>(d_tab, nb, d_ra, d_rb, n_rays, (float)cmax, d_work, d_pass, nullptr);
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
        CK(cudaMemcpy(&passes, d_pass, sizeof(passes), cudaMemcpyDeviceToHost));
    }
    double pairs = (double)n_rays * nb;
    printf("B2w R=%d thr=%d U=%d pf=%d regs=%d spill=%zu | %.3f ms  %.3f Tpairs/s | passes/ray %.2f\n", R, THREADS, U, PREFETCH,
           fa.numRegs, (size_t)fa.localSizeBytes, best, pairs / best / 1e9, (double)passes / n_rays);
    cudaFree(d_tab); cudaFree(d_work); cudaFree(d_pass);
}

template <int R, int THREADS, int U, int STATIC = 0>
static void run_b2c(const std::vector<Bound> &all_bounds, const float4 *d_ra, const float4 *d_rb, int n_rays, int n_sm) {
    std::vector<Bound> bounds(all_bounds.begin(), all_bounds.begin() + CTAB_SLOTS);
    const int nb = (int)bounds.size();
    std::vector<float4> tab(nb);
    double cmax = 0;
    for (auto &b : all_bounds) cmax = fmax(cmax, sqrt(b.c[0] * b.c[0] + b.c[1] * b.c[1] + b.c[2] * b.c[2]));
    auto up = [](double x) { float f = (float)x; if ((double)f < x) f = nextafterf(f, INFINITY); return f; };
    for (int j = 0; j < nb; j += 2) {
        float4 t[2];
        for (int k = 0; k < 2; ++k) {
            const Bound &b = bounds[j + k];
            float cx = (float)b.c[0], cy = (float)b.c[1], cz = (float)b.c[2];
            double dx = b.c[0] - cx, dy = b.c[1] - cy, dz = b.c[2] - cz;
            double r = b.r + sqrt(dx * dx + dy * dy + dz * dz);
            t[k] = make_float4(cx, cy, cz, -up(r * r * 1.001 - ((double)cx * cx + (double)cy * cy + (double)cz * cz)));
        }
        tab[j] = make_float4(t[0].x, t[1].x, t[0].y, t[1].y);
        tab[j + 1] = make_float4(t[0].z, t[1].z, t[0].w, t[1].w);
    }
    CK(cudaMemcpyToSymbol(c_tab, tab.data(), nb * sizeof(float4)));
    int *d_work; unsigned long long *d_pass;
    CK(cudaMalloc(&d_work, sizeof(int)));
    CK(cudaMalloc(&d_pass, sizeof(unsigned long long)));
    auto kern = STATIC ? filter_b2s_kernel<R, THREADS, U> : filter_b2c_kernel<R, THREADS, U>;
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, kern));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, 0));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    unsigned long long passes = 0;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaMemset(d_work, 0, sizeof(int)));
        CK(cudaMemset(d_pass, 0, sizeof(unsigned long long)));
        cudaEventRecord(e0);
        kern<<<n_sm * occ, THREADS>>>(nb, d_ra, d_rb, n_rays, (float)cmax, d_work, d_pass);
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
        CK(cudaMemcpy(&passes, d_pass, sizeof(passes), cudaMemcpyDeviceToHost));
    }
    double pairs = (double)n_rays * nb;
    printf(STATIC ? "B2s(const,static) R=%d thr=%d U=%d regs=%d" : "B2c(const) R=%d thr=%d U=%d regs=%d occ=%d spill=%zu | %.3f ms  %.3f Tpairs/s | passes/ray %.2f (of %d bounds)\n", R, THREADS, U,
           fa.numRegs, occ, (size_t)fa.localSizeBytes, best, pairs / best / 1e9, (double)passes / n_rays, nb);
    cudaFree(d_work); cudaFree(d_pass);
}

int main(int argc, char **argv) {
    int n_rays = argc > 1 ? atoi(argv[1]) : (1 << 21);
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int n_sm = prop.multiProcessorCount;
    printf("device %s, %d SMs\n", prop.name, n_sm);
    float *sink;
    CK(cudaMalloc(&sink, n_sm * 8 * 256 * sizeof(float)));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const char *pname[4] = {"FFMA x16 chains", "FFMA2 x16 chains", "8 FFMA : 1 FSETP", "4 FFMA2 : 2 FSETP"};
    for (int mode = 0; mode < 4; ++mode) {
        float best = 1e30f;
        const int iters = 4096;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            if (mode == 0) probe_kernel<0><<<n_sm * 8, 256>>>(iters, sink, 0.999999f, 1e-7f);
            if (mode == 1) probe_kernel<1><<<n_sm * 8, 256>>>(iters, sink, 0.999999f, 1e-7f);
            if (mode == 2) probe_kernel<2><<<n_sm * 8, 256>>>(iters, sink, 0.999999f, 1e-7f);
            if (mode == 3) probe_kernel<3><<<n_sm * 8, 256>>>(iters, sink, 0.999999f, 1e-7f);
            cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1));
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep > 0 && ms < best) best = ms;
        }
        double lane_fma = (double)n_sm * 8 * 256 * iters * 4.0 * 16 * ((mode == 1 || mode == 3) ? 2 : 1);
        printf("probe %-20s %.3f ms  %.2f TFLOP/s (fma lanes x2)\n", pname[mode], best, 2.0 * lane_fma / best / 1e9);
    }

    // synthetic m10k-like table: 5000 spheres r in [0.03,0.1], 5000 triangle bounds r ~ 1.5 * [0.05,0.15]
    uint64_t s = 10000;
    std::vector<Bound> bounds;
    for (int i = 0; i < 10240; ++i) {
        Bound b;
        b.c[0] = -3.5 + 7.0 * u01(s); b.c[1] = -1.8 + 4.8 * u01(s); b.c[2] = -2.0 + 8.0 * u01(s);
        b.r = i < 5120 ? 0.03 + 0.07 * u01(s) : 1.5 * (0.05 + 0.10 * u01(s));
        bounds.push_back(b);
    }
    // rays: 1/5 primary from the camera, the rest from random points in the box, random directions
    std::vector<float4> ra(n_rays), rb(n_rays);
    for (int i = 0; i < n_rays; ++i) {
        float o[3], d[3];
        if (i % 5 == 0) { o[0] = 0; o[1] = 0; o[2] = -8; d[0] = -3 + 6 * u01(s); d[1] = -3 + 6 * u01(s); d[2] = 8; }
        else {
            o[0] = -3.5f + 7 * u01(s); o[1] = -1.8f + 4.8f * u01(s); o[2] = -2 + 8 * u01(s);
            d[0] = -1 + 2 * u01(s); d[1] = -1 + 2 * u01(s); d[2] = -1 + 2 * u01(s);
        }
        float n = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        if (i % 7 != 0) { d[0] /= n; d[1] /= n; d[2] /= n; } // some rays stay un-normalised (refract_ray does not renormalise)
        ra[i] = make_float4(o[0], o[1], o[2], d[0]);
        rb[i] = make_float4(d[1], d[2], 0, 0);
    }
    float4 *d_ra, *d_rb;
    CK(cudaMalloc(&d_ra, n_rays * sizeof(float4)));
    CK(cudaMalloc(&d_rb, n_rays * sizeof(float4)));
    CK(cudaMemcpy(d_ra, ra.data(), n_rays * sizeof(float4), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_rb, rb.data(), n_rays * sizeof(float4), cudaMemcpyHostToDevice));

    run_filter<0, 2, 512>("A  proj", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<0, 4, 512>("A  proj", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<0, 4, 1024>("A  proj", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<0, 6, 512>("A  proj", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<1, 2, 512>("B  quad", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<1, 4, 512>("B  quad", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<1, 4, 1024>("B  quad", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<1, 6, 512>("B  quad", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<2, 2, 512>("A2 proj packed", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<2, 4, 512>("A2 proj packed", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<3, 2, 512>("B2 quad packed", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<3, 4, 512>("B2 quad packed", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<3, 2, 1024>("B2 quad packed", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<3, 6, 512>("B2 quad packed", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<2, 6, 512>("A2 proj packed", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<3, 4, 256>("B2 quad packed", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    run_filter<1, 4, 256>("B  quad", bounds, d_ra, d_rb, n_rays, ra, rb, n_sm);
    for (int mode = 0; mode < 2; ++mode) {
        float best = 1e30f;
        const int iters = 4096;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            if (mode == 0) rf_probe_kernel<0><<<n_sm * 8, 256>>>(iters, sink, make_float2(0.999f, 0.998f), make_float2(1.001f, 1.002f), make_float2(0.9995f, 1.0005f), make_float2(1.f, 0.9999f));
            else rf_probe_kernel<1><<<n_sm * 8, 256>>>(iters, sink, make_float2(0.999f, 0.998f), make_float2(1.001f, 1.002f), make_float2(0.9995f, 1.0005f), make_float2(1.f, 0.9999f));
            cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1));
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep > 0 && ms < best) best = ms;
        }
        double lane_fma = (double)n_sm * 8 * 256 * iters * 2.0 * 32 * 2;
        printf("rf probe FFMA2 pair-operand from %s: %.3f ms %.2f TFLOP/s\n", mode == 0 ? "vector regs" : "kernel params (uniform)", best, 2.0 * lane_fma / best / 1e9);
    }
    printf("---- constant-memory table\n");
    run_b2c<2, 1024, 2>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<4, 512, 2>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<4, 256, 2>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<4, 768, 2>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<4, 512, 4>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<6, 512, 2>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<8, 512, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<8, 256, 2>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<4, 512, 2, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<4, 256, 2, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<4, 768, 2, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<6, 512, 2, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<8, 512, 1, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<8, 256, 2, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2c<2, 1024, 2, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    printf("---- per-warp batches\n");
    run_b2w<2, 1024, 2, 0>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<2, 1024, 4, 0>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<2, 1024, 2, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<2, 1024, 4, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<3, 1024, 2, 0>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<3, 1024, 2, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<3, 768, 2, 0>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<4, 768, 2, 0>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<4, 768, 2, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<4, 768, 4, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<4, 512, 2, 0>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<4, 512, 2, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<4, 512, 4, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<6, 512, 2, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<6, 512, 1, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    run_b2w<8, 512, 1, 1>(bounds, d_ra, d_rb, n_rays, n_sm);
    return 0;
}
