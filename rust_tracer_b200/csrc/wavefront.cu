// wavefront.cu -- everything of the render path that is not the nearest-hit sweep:
//   raygen   Camera::get_ray                       (reference src/render.rs:178-185)
//   spawn    get_light_energy's shadow rays, reflect_ray, refract_ray
//                                                  (src/render.rs:105-125,142-153; scene/mod.rs:189-193)
//   shade    trace_ray's colour sum / render_ray_tree, deepest level first
//                                                  (src/render.rs:57-100; src/render_tree.rs:214-255)
//   resolve  Color::as_u8 (+ the 4-spp average)    (src/scene/color.rs:43-46)
//
// trace_ray's recursion becomes a breadth-first ray tree held in flat pools: level k of every
// pixel's tree is one contiguous range, swept in one launch.  Colours flow back up through
// per-node child slots in the reference's own summation order (deterministic, no atomics
// on colours), so the only arithmetic that differs from the reference is powf/atan2f/acosf.
#include "kernels.h"
#include "shade.cuh"

namespace rt {

#ifndef RT_SHADE_MIN_BLOCKS
#define RT_SHADE_MIN_BLOCKS 8 // resident blocks per SM the shade kernel is compiled for: 8 = 64 registers, 40 B of spills, 50 % occupancy instead of 37.5 %
                              // (c4: spawn + shade + node order 4.70 -> 4.37 ms, profiles/r02_ab_wavefront_occupancy_c4.log); 0 = the compiler's choice
#endif
#ifndef RT_SPAWN_MIN_BLOCKS
#define RT_SPAWN_MIN_BLOCKS 0 // likewise for the spawn kernel (blocks of 512 threads); 3 = 40 registers measured no faster
#endif
#if RT_SHADE_MIN_BLOCKS > 0
#define RT_SHADE_BOUNDS __launch_bounds__(WF_BLOCK, RT_SHADE_MIN_BLOCKS)
#else
#define RT_SHADE_BOUNDS __launch_bounds__(WF_BLOCK)
#endif
#if RT_SPAWN_MIN_BLOCKS > 0
#define RT_SPAWN_BOUNDS __launch_bounds__(SPAWN_BLOCK, RT_SPAWN_MIN_BLOCKS)
#else
#define RT_SPAWN_BOUNDS __launch_bounds__(SPAWN_BLOCK)
#endif
constexpr int WF_BLOCK = 128;

RT_DEV unsigned long long splitmix64(unsigned long long &state) {
    unsigned long long z = (state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
RT_DEV float u01(unsigned long long x) { return xmul((float)(x >> 40), 1.0f / 16777216.0f); }

__global__ void __launch_bounds__(WF_BLOCK) raygen_kernel(FrameParams fp, Pools p, long long sample0, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    long long sidx = sample0 + i;
    long long pix = sidx / fp.spp;
    int s = (int)(sidx - pix * fp.spp);
    // Tree ray i carries the sample `sidx` of the call (t_parent = t_root = its index in the band).  Row by row, a warp's
    // 128 consecutive rays are a fan one pixel high and 128 wide; laid out in tiles of 16 x 8 pixels they form a narrow
    // pyramid that touches about a third as many bound groups.  Chunks of 8 rendered rows that lie wholly inside this band
    // are walked tile by tile (the last tile of a chunk may be narrower, the last chunk lower); the rest stays row by row.
    if (fp.primary_tiles) {
        // 16 x 8 pixels = one batch of 128 rays; at 4 samples per pixel (primary_tiles = 2) 8 x 4 pixels are one batch
        const int TW = (fp.primary_tiles == 2 && fp.spp == 4) ? 8 : 16, TH = (fp.primary_tiles == 2 && fp.spp == 4) ? 4 : 8;
        const long long W = fp.x_res, chunk = pix / (W * TH);
        const long long c0 = chunk * W * TH, c1 = min(c0 + W * TH, (long long)fp.n_rows * W); // pixels of the chunk
        if (c0 * fp.spp >= sample0 && c1 * fp.spp <= sample0 + n) {
            const int h = (int)((c1 - c0) / W);             // rows of this chunk
            const long long k = pix - c0;                   // position in the chunk's tile order
            const int t = (int)(k / (TW * h));              // tile
            const int m = (int)(k - (long long)t * TW * h); // position in the tile
            const int tw = (int)min((long long)TW, W - (long long)t * TW);
            pix = c0 + (long long)(m / tw) * W + (long long)t * TW + m % tw;
            sidx = pix * fp.spp + s;
        }
    }
    int v = fp.rows[pix / fp.x_res];
    int u = (int)(pix % fp.x_res);
    float x_delta = xdiv(xsub(fp.x_max, fp.x_min), (float)fp.x_res);
    float y_delta = xdiv(xsub(fp.y_max, fp.y_min), (float)fp.y_res);
    float fu = (float)u, fv = (float)v;
    if (fp.spp == 4) { // extension, no reference behaviour; mirrored by the oracle
        unsigned long long key = ((unsigned long long)v * (unsigned long long)fp.x_res + (unsigned long long)u) * 4ull +
                                 (unsigned long long)s;
        unsigned long long st = fp.seed ^ (key * 0xD1B54A32D192ED03ull);
        float xi1 = u01(splitmix64(st));
        float xi2 = u01(splitmix64(st));
        fu = xadd(fu, xmul(xadd((float)(s & 1), xi1), 0.5f));
        fv = xadd(fv, xmul(xadd((float)(s >> 1), xi2), 0.5f));
    }
    float x = xadd(fp.x_min, xmul(fu, x_delta));
    float y = xsub(fp.y_max, xmul(fv, y_delta));
    V3 origin = v3(fp.cam_ox, fp.cam_oy, fp.cam_oz);
    V3 dir = vnorm(vsub(v3(x, y, 0.f), origin));
    p.t_a[i] = make_float4(origin.x, origin.y, origin.z, dir.x);
    p.t_b[i] = make_float4(dir.y, dir.z, 0.f, 0.f);
    p.t_parent[i] = (int)(sidx - sample0);
    p.t_root[i] = (int)(sidx - sample0);
}

RT_DEV void load_tree_ray(const Pools &p, int i, V3 &o, V3 &d) {
    float4 a = p.t_a[i], b = p.t_b[i];
    o = v3(a.x, a.y, a.z);
    d = v3(a.w, b.x, b.y);
}

// ---- node order: counting sort of a level's tree rays by the cell of their hit point ----------
RT_DEV unsigned int spread3_dev(unsigned int x) { // <= 10 bits -> every third bit
    x &= 0x3ffu;
    x = (x | (x << 16)) & 0x030000ffu;
    x = (x | (x << 8)) & 0x0300f00fu;
    x = (x | (x << 4)) & 0x030c30c3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

#ifndef RT_NODE_HILBERT
#define RT_NODE_HILBERT 1 // cells inside the box are numbered along a Hilbert curve (no jumps between consecutive cells) rather than in Morton order
#endif
// Index of cell (x, y, z), `bits` bits per axis, along the 3-D Hilbert curve (Skilling, "Programming the Hilbert curve", 2004:
// axes -> transpose, then the bits of the three words interleaved).
RT_DEV unsigned int hilbert3_dev(unsigned int x, unsigned int y, unsigned int z, int bits) {
    unsigned int X[3] = {x, y, z};
    const unsigned int M = 1u << (bits - 1);
    for (unsigned int Q = M; Q > 1u; Q >>= 1) { // inverse undo
        const unsigned int P = Q - 1u;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            if (X[i] & Q) X[0] ^= P;
            else {
                const unsigned int t = (X[0] ^ X[i]) & P;
                X[0] ^= t;
                X[i] ^= t;
            }
        }
    }
    X[1] ^= X[0]; // Gray encode
    X[2] ^= X[1];
    unsigned int t = 0;
    for (unsigned int Q = M; Q > 1u; Q >>= 1)
        if (X[2] & Q) t ^= Q - 1u;
    X[0] ^= t; X[1] ^= t; X[2] ^= t;
    return (spread3_dev(X[0]) << 2) | (spread3_dev(X[1]) << 1) | spread3_dev(X[2]);
}

__global__ void __launch_bounds__(256) node_bin_kernel(DevScene sc, Pools p, int begin, int end) {
    int i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < end;
    if (!live) i = end - 1; // keep the warp whole for the match below
    unsigned long long key = p.t_key[i];
    const int bits = p.node_bits;
    unsigned int bin = (unsigned int)node_bins(bits); // miss
    if (key != KEY_MISS) {
        V3 o, d;
        float4 a = p.t_a[i], b = p.t_b[i];
        o = v3(a.x, a.y, a.z);
        d = v3(a.w, b.x, b.y);
        float t = key_t(key); // only a sort key: any point near the hit will do
        // u in [-1, 1]^3 inside the box of the bounded shapes: `bits` bits per axis, Morton
        // order.  Hit points outside (the far reaches of a plane) get coarser cells of u / (1 + |u|)
        // and come FIRST: their rays carry large margins and are the expensive ones, so the
        // sweeps start them early instead of finishing on them.
        const float ux = fmaf(fmaf(t, d.x, o.x) - sc.bin_lo[0], sc.bin_scale[0], -1.f);
        const float uy = fmaf(fmaf(t, d.y, o.y) - sc.bin_lo[1], sc.bin_scale[1], -1.f);
        const float uz = fmaf(fmaf(t, d.z, o.z) - sc.bin_lo[2], sc.bin_scale[2], -1.f);
        const bool inside = fabsf(ux) <= 1.f && fabsf(uy) <= 1.f && fabsf(uz) <= 1.f; // false for NaN
        if (inside) {
            const float half = (float)(1 << (bits - 1)), top = 2.f * half - 1.f;
            auto cell = [=](float u) { return (unsigned int)fminf(fmaxf((u + 1.f) * half, 0.f), top); };
#if RT_NODE_HILBERT
            bin = NODE_FAR_BINS + hilbert3_dev(cell(ux), cell(uy), cell(uz), bits);
#else
            bin = NODE_FAR_BINS + (spread3_dev(cell(ux)) | (spread3_dev(cell(uy)) << 1) | (spread3_dev(cell(uz)) << 2));
#endif
        } else {
            auto cell = [](float u) {
                float c = u / (1.f + fabsf(u)); // (-1, 1); NaN -> cell 0 below
                float q = (c + 1.f) * (float)(1 << (NODE_FAR_BITS - 1));
                return !(q >= 0.f) ? 0u : (unsigned int)fminf(q, (float)((1 << NODE_FAR_BITS) - 1));
            };
            bin = spread3_dev(cell(ux)) | (spread3_dev(cell(uy)) << 1) | (spread3_dev(cell(uz)) << 2);
        }
    }
    if (live) p.t_sbase[i] = (int)bin; // parked here until spawn_kernel overwrites it
    // neighbouring rays mostly share a cell: one atomic per distinct cell of the warp
    const unsigned int same = __match_any_sync(0xffffffffu, live ? bin : 0xffffffffu);
    if (live && (threadIdx.x & 31) == __ffs(same) - 1) atomicAdd(&p.bins[bin], (unsigned int)__popc(same));
}

// Exclusive scan of bins[0 .. n_bins] (n_bins = node_bins(p.node_bits); up to 2 M bins) in two launches:
//   node_scan_chunks_kernel  block b scans its chunk of SCAN_CHUNK bins in place (exclusive within the chunk)
//                            and leaves the chunk's total in bin_sums[b];
//   node_scan_top_kernel     one block scans the <= 1024 chunk totals in place, and reads off the two prefixes
//                            the level needs: hits outside the box (counters[4]) and all hits (-> counters[1],
//                            the end of the level's shadow rays).
// The prefix of bin j is then bins[j] + bin_sums[j / SCAN_CHUNK]; node_scatter_kernel adds the two when it
// claims places, so there is no third pass over the bins.
constexpr int SCAN_CHUNK = 4096, SCAN_THREADS = 512, SCAN_PER_THREAD = SCAN_CHUNK / SCAN_THREADS;

__global__ void __launch_bounds__(SCAN_THREADS) node_scan_chunks_kernel(Pools p) {
    __shared__ unsigned int warp_sum[SCAN_THREADS / 32];
    const int n = node_bins(p.node_bits) + 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int base = blockIdx.x * SCAN_CHUNK + tid * SCAN_PER_THREAD; // a thread owns SCAN_PER_THREAD consecutive bins (two 16-byte loads)
    unsigned int v[SCAN_PER_THREAD];
    if (base + SCAN_PER_THREAD <= n) {
        const uint4 a = *(const uint4 *)(p.bins + base), b = *(const uint4 *)(p.bins + base + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_PER_THREAD; ++k) v[k] = base + k < n ? p.bins[base + k] : 0u;
    }
    unsigned int mine = 0;
#pragma unroll
    for (int k = 0; k < SCAN_PER_THREAD; ++k) { const unsigned int x = v[k]; v[k] = mine; mine += x; } // exclusive within the thread
    unsigned int t = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int y = __shfl_up_sync(0xffffffffu, t, o);
        if (lane >= o) t += y;
    }
    if (lane == 31) warp_sum[warp] = t;
    __syncthreads();
    if (warp == 0) {
        unsigned int w = lane < SCAN_THREADS / 32 ? warp_sum[lane] : 0u, z = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int y = __shfl_up_sync(0xffffffffu, z, o);
            if (lane >= o) z += y;
        }
        if (lane < SCAN_THREADS / 32) warp_sum[lane] = z - w;
        if (lane == SCAN_THREADS / 32 - 1) p.bin_sums[blockIdx.x] = z; // the chunk's total
    }
    __syncthreads();
    const unsigned int before = warp_sum[warp] + t - mine; // bins of the chunk before this thread's
    if (base + SCAN_PER_THREAD <= n) {
        *(uint4 *)(p.bins + base) = make_uint4(before + v[0], before + v[1], before + v[2], before + v[3]);
        *(uint4 *)(p.bins + base + 4) = make_uint4(before + v[4], before + v[5], before + v[6], before + v[7]);
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_PER_THREAD; ++k)
            if (base + k < n) p.bins[base + k] = before + v[k];
    }
}

__global__ void __launch_bounds__(1024) node_scan_top_kernel(Pools p, int n_chunks, int s_begin, int n_point_lights) {
    __shared__ unsigned int warp_sum[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int v = tid < n_chunks ? p.bin_sums[tid] : 0u; // n_chunks <= 1024
    unsigned int t = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int y = __shfl_up_sync(0xffffffffu, t, o);
        if (lane >= o) t += y;
    }
    if (lane == 31) warp_sum[warp] = t;
    __syncthreads();
    if (warp == 0) {
        unsigned int w = warp_sum[lane], z = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int y = __shfl_up_sync(0xffffffffu, z, o);
            if (lane >= o) z += y;
        }
        warp_sum[lane] = z - w;
    }
    __syncthreads();
    const unsigned int excl = warp_sum[warp] + t - v;
    if (tid < n_chunks) p.bin_sums[tid] = excl;
    const int n_bins = node_bins(p.node_bits);
    if (tid == NODE_FAR_BINS / SCAN_CHUNK) { // prefix of bin NODE_FAR_BINS = number of hits outside the box
        p.counters[4] = excl + p.bins[NODE_FAR_BINS];
        p.counters[5] = 0u;
    }
    if (tid == n_bins / SCAN_CHUNK) { // prefix of the last bin (the misses) = number of hits
        const unsigned int hits = excl + p.bins[n_bins];
        unsigned long long s_end = (unsigned long long)s_begin + (unsigned long long)hits * (unsigned long long)n_point_lights;
        if (s_end > (unsigned long long)p.s_cap) p.counters[3] = 1u;
        else p.counters[1] = (unsigned int)s_end;
    }
}

__global__ void __launch_bounds__(256) node_scatter_kernel(Pools p, int begin, int end) {
    int i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < end;
    const int lane = threadIdx.x & 31;
    const unsigned int bin = live ? (unsigned int)p.t_sbase[i] : 0xffffffffu;
    const unsigned int same = __match_any_sync(0xffffffffu, bin);
    const int leader = __ffs(same) - 1;
    unsigned int base = 0;
    if (live && lane == leader) base = atomicAdd(&p.bins[bin], (unsigned int)__popc(same)) + p.bin_sums[bin / SCAN_CHUNK];
    base = __shfl_sync(0xffffffffu, base, leader);
    RT_CHECK(!live || (int)(base + __popc(same & ((1u << lane) - 1u))) < end - begin);
    if (live) p.t_order[begin + base + __popc(same & ((1u << lane) - 1u))] = i;
}

// Child rays of one block are allocated together: its reflected rays first, then its refracted
// rays.  With the nodes in cell order (node_*_kernel) a run of up to SPAWN_BLOCK consecutive tree
// rays of the next level then starts at neighbouring points AND leaves the same way -- away from
// the surface or into it -- which is what the sweep's group culling feeds on.
#ifndef RT_SPAWN_BLOCK
#define RT_SPAWN_BLOCK 512 // threads per block of the spawn kernel (A/B builds)
#endif
constexpr int SPAWN_BLOCK = RT_SPAWN_BLOCK;

__global__ void RT_SPAWN_BOUNDS
spawn_kernel(DevScene sc, FrameParams fp, Pools p, int begin, int end, int spawn_children, int ordered, int s_begin) {
    const int rank = blockIdx.x * blockDim.x + threadIdx.x; // node number within the level when ordered
    bool refl = false, refr = false, far_node = false;
    V3 ro0, rd0, ro1, rd1;
    int i = -1;
    do {
        if (begin + rank >= end) break;
        i = ordered ? p.t_order[begin + rank] : begin + rank;
        RT_CHECK(i >= begin && i < end);
        p.t_sbase[i] = -1;
        p.t_crefl[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        p.t_crefr[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        V3 o, d;
        load_tree_ray(p, i, o, d);
        Hit h;
        if (!hit_record(sc, p.t_key[i], o, d, h)) {
            if (p.t_hitid) p.t_hitid[i] = -1;
            break;
        }
        if (p.t_hitid) p.t_hitid[i] = h.id;
        atomicAdd(&p.counters[2], 1u);
        const DevMaterial m = sc.materials[h.material];
        float n1, n2;
        refraction_indices(m, h.entering, n1, n2);

        // shadow rays: p = point + 0.0002 * normal (render.rs:147); one per point light
        if (sc.n_point_lights > 0) {
            // ordered: the level's shadow rays are laid out light by light, [light][node]: node number
            // `rank` owns ray s_begin + light * hits + rank (hits = the level's nodes, from node_scan_kernel
            // through counters[1]), so each light's rays are one contiguous, coalesced run for its sweep.
            // Not ordered: [node][light] blocks handed out by atomicAdd.
            unsigned long long sb, lstride = 1;
            bool fits;
            if (ordered) {
                fits = p.counters[3] == 0u; // node_scan_kernel checked the level's end against s_cap
                lstride = (p.counters[1] - (unsigned int)s_begin) / (unsigned int)sc.n_point_lights;
                sb = (unsigned long long)s_begin + (unsigned long long)rank;
            } else {
                sb = (unsigned long long)atomicAdd(&p.counters[1], (unsigned int)sc.n_point_lights);
                fits = sb + (unsigned int)sc.n_point_lights <= (unsigned long long)p.s_cap;
            }
            if (fits) {
                p.t_sbase[i] = (int)sb;
                V3 sp = vadd(h.point, vscale(h.normal, 0.0002f));
                for (int l = 0; l < sc.n_lights; ++l) {
                    DevLight L = sc.lights[l];
                    if (L.kind != 0) continue;
                    V3 ldir = vnorm(vsub(v3(L.px, L.py, L.pz), sp)); // scene/mod.rs:191
                    RT_CHECK(sb + L.point_index * lstride < (unsigned long long)p.s_cap);
                    p.s_a[sb + L.point_index * lstride] = make_float4(sp.x, sp.y, sp.z, ldir.x);
                    p.s_b[sb + L.point_index * lstride] = make_float4(ldir.y, ldir.z, 0.f, 0.f);
                }
            } else {
                p.counters[3] = 1u;
            }
        }
        if (!spawn_children) break;
        far_node = ordered && (unsigned int)rank < p.counters[4];
        if (m.reflectivity > F32_EPS) { // render.rs:70
            reflect_ray(d, h, ro0, rd0);
            refl = true;
        }
        if (m.refraction_index > F32_EPS) // render.rs:86
            refr = refract_ray(d, h, n1, n2, ro1, rd1);
    } while (false);
    if (!spawn_children) return;

    // block-wide allocation: [reflected rays of the block | refracted rays of the block]
    __shared__ unsigned int warp_refl[SPAWN_BLOCK / 32], warp_refr[SPAWN_BLOCK / 32], warp_far[SPAWN_BLOCK / 32];
    __shared__ unsigned int block_base, block_refl;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned int m_refl = __ballot_sync(0xffffffffu, refl), m_refr = __ballot_sync(0xffffffffu, refr);
    const unsigned int m_far = __ballot_sync(0xffffffffu, far_node && refl), m_far2 = __ballot_sync(0xffffffffu, far_node && refr);
    if (lane == 0) {
        warp_refl[warp] = __popc(m_refl);
        warp_refr[warp] = __popc(m_refr);
        warp_far[warp] = __popc(m_far) + __popc(m_far2);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int n_refl = 0, n_refr = 0, n_far = 0;
        for (int w = 0; w < SPAWN_BLOCK / 32; ++w) {
            const unsigned int a = warp_refl[w], b = warp_refr[w];
            warp_refl[w] = n_refl; // exclusive prefixes
            warp_refr[w] = n_refr;
            n_refl += a; n_refr += b; n_far += warp_far[w];
        }
        block_refl = n_refl;
        block_base = (n_refl + n_refr) ? atomicAdd(&p.counters[0], n_refl + n_refr) : 0u;
        if (n_far) atomicAdd(&p.counters[5], n_far);
    }
    __syncthreads();
    const unsigned int below = (1u << lane) - 1u;
    if (refl) {
        const unsigned int c = block_base + warp_refl[warp] + __popc(m_refl & below);
        if (c < (unsigned int)p.t_cap) {
            p.t_a[c] = make_float4(ro0.x, ro0.y, ro0.z, rd0.x);
            p.t_b[c] = make_float4(rd0.y, rd0.z, 0.f, 0.f);
            p.t_parent[c] = (i << 1) | 0;
            p.t_root[c] = p.t_root[i];
        } else {
            p.counters[3] = 1u;
        }
    }
    if (refr) {
        const unsigned int c = block_base + block_refl + warp_refr[warp] + __popc(m_refr & below);
        if (c < (unsigned int)p.t_cap) {
            p.t_a[c] = make_float4(ro1.x, ro1.y, ro1.z, rd1.x);
            p.t_b[c] = make_float4(rd1.y, rd1.z, 0.f, 0.f);
            p.t_parent[c] = (i << 1) | 1;
            p.t_root[c] = p.t_root[i];
        } else {
            p.counters[3] = 1u;
        }
    }
}

__global__ void RT_SHADE_BOUNDS
shade_kernel(DevScene sc, FrameParams fp, Pools p, int begin, int end, int level, int light_stride, float4 *sample_rgb,
             const unsigned char *sample_mask) {
    int i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= end) return;
    if (sample_mask && !sample_mask[p.t_root[i]]) return; // render_forest_filter: tree untouched
    V3 o, d;
    load_tree_ray(p, i, o, d);
    Hit h;
    V3 colour = v3(0.f, 0.f, 0.f);
    bool is_node = hit_record(sc, p.t_key[i], o, d, h);
    if (is_node) {
        const DevMaterial m = sc.materials[h.material];
        if (material_reads_tex(m)) hit_tex(h);
        float n1, n2;
        refraction_indices(m, h.entering, n1, n2);
        V3 ambient = cmul(eval_colorfn(m.ambient, h.tu, h.tv), v3(sc.amb_r, sc.amb_g, sc.amb_b));

        // lights (render.rs:59-68 / render_tree.rs:223-232): Sum starts from BLACK
        V3 lights = v3(0.f, 0.f, 0.f);
        int sb = p.t_sbase[i];
        V3 sp = vadd(h.point, vscale(h.normal, 0.0002f));
        for (int l = 0; l < sc.n_lights; ++l) {
            DevLight L = sc.lights[l];
            V3 ldir, energy = v3(L.cr, L.cg, L.cb);
            if (L.kind == 0) { // PointLight::get_energy (scene/mod.rs:189-206)
                V3 lpos = v3(L.px, L.py, L.pz);
                ldir = vnorm(vsub(lpos, sp));
                RT_CHECK(sb >= 0 && sb + (long long)L.point_index * light_stride < p.s_cap);
                unsigned long long sk = p.s_key[sb + (long long)L.point_index * light_stride];
                if (sk != KEY_MISS) {
                    V3 hp = ray_at(sp, ldir, key_t(sk));
                    if (vlen2(vsub(hp, sp)) < vlen2(vsub(lpos, sp))) energy = v3(0.f, 0.f, 0.f);
                }
            } else { // AmbientLight::get_energy (scene/mod.rs:236-238)
                ldir = v3(0.f, 0.f, 0.f);
            }
            float fresnel = fresnel_reflection(ldir, h.normal, n1, n2);
            lights = vadd(lights, cscale(fresnel, reflected_energy(m, energy, ldir, h)));
        }

        float4 cr = p.t_crefl[i], ct = p.t_crefr[i];
        V3 reflected = v3(0.f, 0.f, 0.f), refracted = v3(0.f, 0.f, 0.f);
        if (!fp.forest_shading) {
            if (m.reflectivity > F32_EPS) { // render.rs:70-84
                V3 ro, rd;
                reflect_ray(d, h, ro, rd);
                float fresnel = fresnel_reflection(rd, h.normal, n1, n2);
                reflected = cscale(fresnel, reflected_energy(m, v3(cr.x, cr.y, cr.z), rd, h));
            }
            if (m.refraction_index > F32_EPS) { // render.rs:86-98
                V3 ro, rd, inner = v3(0.f, 0.f, 0.f);
                if (refract_ray(d, h, n1, n2, ro, rd)) {
                    float fresnel = xsub(1.f, fresnel_reflection(rd, vneg(h.normal), n1, n2));
                    inner = cscale(fresnel, v3(ct.x, ct.y, ct.z));
                }
                refracted = cmul(eval_colorfn(m.diffuse, h.tu, h.tv), inner);
            }
        } else {
            // render_tree.rs:234-249: a missing child evaluates to (BLACK, (0,0,0)); an
            // existing one returns its -eye_dir = norm(its ray direction)
            V3 dir = v3(0.f, 0.f, 0.f);
            if (cr.w != 0.f) {
                V3 ro, rd;
                reflect_ray(d, h, ro, rd);
                dir = vnorm(rd); // -(child eye_dir) = norm(child ray direction)
            }
            float fresnel = fresnel_reflection(dir, h.normal, n1, n2);
            reflected = cscale(fresnel, reflected_energy(m, v3(cr.x, cr.y, cr.z), h.eye, h));
            dir = v3(0.f, 0.f, 0.f);
            if (ct.w != 0.f) {
                // the child exists, so refract_ray succeeded at build time; n1 / n2 cannot have changed since
                // (rt_scene_update_materials refuses that while a forest is alive)
                V3 ro, rd = v3(0.f, 0.f, 0.f);
                refract_ray(d, h, n1, n2, ro, rd);
                dir = vnorm(rd); // -(child eye_dir) = norm(child ray direction)
            }
            float ft = xsub(1.f, fresnel_reflection(dir, vneg(h.normal), n1, n2));
            refracted = cscale(ft, v3(ct.x, ct.y, ct.z));
        }
        colour = vadd(vadd(vadd(ambient, lights), reflected), refracted); // render.rs:100
    }
    RT_CHECK(level == 0 || ((p.t_parent[i] >> 1) >= 0 && (p.t_parent[i] >> 1) < begin));
    if (level == 0) {
        sample_rgb[p.t_parent[i]] = make_float4(colour.x, colour.y, colour.z, is_node ? 1.f : 0.f);
    } else if (is_node) {
        int par = p.t_parent[i];
        float4 out = make_float4(colour.x, colour.y, colour.z, 1.f);
        if (par & 1) p.t_crefr[par >> 1] = out;
        else p.t_crefl[par >> 1] = out;
    }
}

__global__ void __launch_bounds__(256)
resolve_kernel(const float4 *sample_rgb, long long pixel0, int n_pixels, int spp, float *rgb_f32,
               unsigned char *rgb_u8, const unsigned char *pixel_mask, const int *frame_rows, int width) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pixels) return;
    if (pixel_mask && !pixel_mask[i]) return;
    V3 c;
    if (spp == 1) {
        float4 s = sample_rgb[i];
        c = v3(s.x, s.y, s.z);
    } else {
        V3 sum = v3(0.f, 0.f, 0.f);
        for (int s = 0; s < spp; ++s) {
            float4 q = sample_rgb[(long long)i * spp + s];
            sum = vadd(sum, v3(q.x, q.y, q.z));
        }
        c = cscale(0.25f, sum);
    }
    long long o = pixel0 + i; // pixel number among the rendered rows
    if (frame_rows) o = (long long)frame_rows[o / width] * width + o % width; // its place in the whole frame
    o *= 3;
    if (rgb_f32) { rgb_f32[o] = c.x; rgb_f32[o + 1] = c.y; rgb_f32[o + 2] = c.z; }
    if (rgb_u8) { rgb_u8[o] = as_u8(c.x); rgb_u8[o + 1] = as_u8(c.y); rgb_u8[o + 2] = as_u8(c.z); }
}

// One slot per masked pixel: a warp counts its masked lanes and takes that many slots with one atomic.
__global__ void __launch_bounds__(256)
resolve_compact_kernel(const float4 *sample_rgb, int n_pixels, const unsigned char *pixel_mask, int *out_idx, unsigned int *count,
                       float *rgb_f32, unsigned char *rgb_u8) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool on = i < n_pixels && pixel_mask[i];
    const unsigned int ball = __ballot_sync(0xffffffffu, on);
    if (!ball) return;
    const int lane = threadIdx.x & 31;
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(count, (unsigned int)__popc(ball));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (!on) return;
    const size_t k = base + __popc(ball & ((1u << lane) - 1u));
    const float4 s = sample_rgb[i];
    out_idx[k] = i;
    if (rgb_f32) { rgb_f32[3 * k] = s.x; rgb_f32[3 * k + 1] = s.y; rgb_f32[3 * k + 2] = s.z; }
    if (rgb_u8) { rgb_u8[3 * k] = as_u8(s.x); rgb_u8[3 * k + 1] = as_u8(s.y); rgb_u8[3 * k + 2] = as_u8(s.z); }
}

__global__ void forest_sizes_kernel(Pools p, int n, int *sizes) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && p.t_hitid[i] >= 0) atomicAdd(&sizes[p.t_root[i]], 1);
}

__global__ void forest_mark_kernel(Pools p, int n, const int *ids, int n_ids, unsigned char *mask) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int id = p.t_hitid[i];
    if (id < 0) return;
    for (int k = 0; k < n_ids; ++k)
        if (ids[k] == id) { mask[p.t_root[i]] = 1; return; }
}

static unsigned int nblk(long long n, int b) { return (unsigned int)((n + b - 1) / b); }

void launch_raygen(const FrameParams &fp, const Pools &p, long long sample0, int n, cudaStream_t stream) {
    if (n > 0) raygen_kernel<<<nblk(n, WF_BLOCK), WF_BLOCK, 0, stream>>>(fp, p, sample0, n);
}
void launch_spawn(const DevScene &sc, const FrameParams &fp, const Pools &p, int begin, int end, int spawn_children,
                  int ordered, int s_begin, cudaStream_t stream) {
    if (end <= begin) return;
    if (ordered) {
        cudaMemsetAsync(p.bins, 0, (size_t)(node_bins(p.node_bits) + 2) * sizeof(unsigned int), stream);
        node_bin_kernel<<<nblk(end - begin, 256), 256, 0, stream>>>(sc, p, begin, end);
        const int n_chunks = (node_bins(p.node_bits) + 1 + SCAN_CHUNK - 1) / SCAN_CHUNK;
        node_scan_chunks_kernel<<<n_chunks, SCAN_THREADS, 0, stream>>>(p);
        node_scan_top_kernel<<<1, 1024, 0, stream>>>(p, n_chunks, s_begin, sc.n_point_lights);
        node_scatter_kernel<<<nblk(end - begin, 256), 256, 0, stream>>>(p, begin, end);
    }
    spawn_kernel<<<nblk(end - begin, SPAWN_BLOCK), SPAWN_BLOCK, 0, stream>>>(sc, fp, p, begin, end, spawn_children, ordered, s_begin);
}
void launch_shade(const DevScene &sc, const FrameParams &fp, const Pools &p, int begin, int end, int level, int light_stride,
                  float4 *sample_rgb, const unsigned char *sample_mask, cudaStream_t stream) {
    if (end > begin)
        shade_kernel<<<nblk(end - begin, WF_BLOCK), WF_BLOCK, 0, stream>>>(sc, fp, p, begin, end, level, light_stride, sample_rgb,
                                                                          sample_mask);
}
void launch_resolve(const float4 *sample_rgb, long long pixel0, int n_pixels, int spp, float *rgb_f32,
                    unsigned char *rgb_u8, const unsigned char *pixel_mask, const int *frame_rows, int width,
                    cudaStream_t stream) {
    if (n_pixels > 0)
        resolve_kernel<<<nblk(n_pixels, 256), 256, 0, stream>>>(sample_rgb, pixel0, n_pixels, spp, rgb_f32, rgb_u8,
                                                                pixel_mask, frame_rows, width);
}

void launch_resolve_compact(const float4 *sample_rgb, int n_pixels, const unsigned char *pixel_mask, int *out_idx,
                            unsigned int *count, float *rgb_f32, unsigned char *rgb_u8, cudaStream_t stream) {
    if (n_pixels > 0)
        resolve_compact_kernel<<<nblk(n_pixels, 256), 256, 0, stream>>>(sample_rgb, n_pixels, pixel_mask, out_idx, count, rgb_f32, rgb_u8);
}

void launch_forest_sizes(const Pools &p, int n, int *sizes, cudaStream_t stream) {
    if (n > 0) forest_sizes_kernel<<<nblk(n, 256), 256, 0, stream>>>(p, n, sizes);
}
void launch_forest_mark(const Pools &p, int n, const int *ids, int n_ids, unsigned char *mask, cudaStream_t stream) {
    if (n > 0) forest_mark_kernel<<<nblk(n, 256), 256, 0, stream>>>(p, n, ids, n_ids, mask);
}

} // namespace rt
