// api_render.cu -- C ABI: Basic render, RayForest build/eval, FP32 peak (include/rt_b200.h).
// Host orchestration of the wavefront: per band of samples, per ray-tree level:
//   sweep(children of level k  +  shadow rays of level k-1)  ->  spawn(level k)
// then shade from the deepest level up and resolve to RGB.
#include <algorithm>
#include <chrono>
#include <climits>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

#include "api_internal.h"

namespace rt {

// tuning().trace: wall-clock laps of the host side of a call
struct HostLap {
    const char *call;
    bool on;
    std::chrono::steady_clock::time_point t;
    explicit HostLap(const char *c) : call(c), on(tuning().trace != 0), t(std::chrono::steady_clock::now()) {}
    void operator()(const char *what) {
        if (!on) return;
        auto n = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[rt_b200] %s: %-20s %.3f ms\n", call, what, std::chrono::duration<double, std::milli>(n - t).count());
        t = n;
    }
};

// ---- workspace --------------------------------------------------------------------------

constexpr int BATCH_COUNTERS_MIN = 1024; // one work counter per sweep launch of a band: (depth + 1) x (1 + lights) x table chunks

template <typename T> static int dev_alloc(T **p, size_t n, const char *what) {
    *p = nullptr;
    if (n == 0) return RT_OK;
    cudaError_t e = cudaMalloc((void **)p, n * sizeof(T));
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(e == cudaErrorMemoryAllocation ? RT_ERR_NOMEM : RT_ERR_CUDA, "cudaMalloc(%s, %zu bytes): %s", what,
                    n * sizeof(T), cudaGetErrorString(e));
    }
    return RT_OK;
}

void Workspace::release_pools() {
    Pools &p = pools;
    cudaFree(p.t_a); cudaFree(p.t_b); cudaFree(p.t_key); cudaFree(p.t_parent); cudaFree(p.t_sbase);
    cudaFree(p.t_crefl); cudaFree(p.t_crefr); cudaFree(p.t_root); cudaFree(p.t_hitid); cudaFree(p.t_order);
    cudaFree(p.s_a); cudaFree(p.s_b); cudaFree(p.s_key);
    unsigned int *c = p.counters, *b = p.bins, *bs = p.bin_sums;
    const int bits = p.node_bits;
    std::memset(&p, 0, sizeof(p));
    p.counters = c;
    p.bins = b;
    p.bin_sums = bs;
    p.node_bits = bits;
}

void Workspace::release() {
    release_pools();
    cudaFree(pools.counters); pools.counters = nullptr;
    cudaFree(pools.bins); pools.bins = nullptr;
    cudaFree(pools.bin_sums); pools.bin_sums = nullptr;
    cudaFree(sample_rgb); sample_rgb = nullptr; sample_cap = 0;
    cudaFree(rows); rows = nullptr; rows_cap = 0;
    cudaFree(mask); mask = nullptr; mask_cap = 0;
    cudaFree(idx); idx = nullptr; idx_cap = 0;
    if (h_counters) cudaFreeHost(h_counters);
    h_counters = nullptr;
    cudaFree(exact_counter); exact_counter = nullptr;
    cudaFree(batch_counters); batch_counters = nullptr; batch_cap = 0;
    cudaFree(out_f32); cudaFree(out_u8);
    out_f32 = out_u8 = nullptr; out_f32_cap = out_u8_cap = 0;
    cudaFree(tables); tables = nullptr; tables_cap = 0;
    for (int i = 0; i < n_aux; ++i) cudaStreamDestroy(aux[i]);
    n_aux = 0;
    if (stream) cudaStreamDestroy(stream);
    stream = nullptr;
    for (int i = 0; i < n_events; ++i) cudaEventDestroy(events[i]);
    n_events = 0;
}

// ---- per-device cache of one released workspace --------------------------------------------
namespace {
constexpr int MAX_CACHED_DEVICES = 64;
std::mutex g_cache_mu;
Workspace g_cache[MAX_CACHED_DEVICES];
bool g_cache_full[MAX_CACHED_DEVICES];
long long ws_weight(const Workspace &w) { return (long long)w.pools.t_cap + w.pools.s_cap + w.sample_cap; }
} // namespace

Workspace workspace_take_cached(int device) {
    Workspace w;
    if (device < 0 || device >= MAX_CACHED_DEVICES) return w;
    std::lock_guard<std::mutex> lock(g_cache_mu);
    if (g_cache_full[device]) {
        w = g_cache[device];
        g_cache[device] = Workspace();
        g_cache_full[device] = false;
    }
    return w;
}

// Caller has made `device` current.
void workspace_give_cached(int device, Workspace &ws) {
    Workspace drop;
    {
        std::lock_guard<std::mutex> lock(g_cache_mu);
        if (device < 0 || device >= MAX_CACHED_DEVICES) {
            drop = ws;
        } else if (!g_cache_full[device]) {
            g_cache[device] = ws;
            g_cache_full[device] = true;
        } else if (ws_weight(g_cache[device]) < ws_weight(ws)) {
            drop = g_cache[device];
            g_cache[device] = ws;
        } else {
            drop = ws;
        }
    }
    ws = Workspace();
    drop.release();
}

int Workspace::ensure_pools(long long t_cap, long long s_cap) {
    if (t_cap > INT_MAX / 2 || s_cap > INT_MAX) return fail(RT_ERR_NOMEM, "ray pool of %lld tree rays exceeds the index range; lower the band size", t_cap);
    Pools &p = pools;
    if (p.t_cap >= t_cap && p.s_cap >= s_cap && p.t_a) return RT_OK;
    release_pools();
    int rc;
    size_t t = (size_t)t_cap, s = (size_t)std::max<long long>(s_cap, 1);
    if ((rc = dev_alloc(&p.t_a, t, "t_a")) || (rc = dev_alloc(&p.t_b, t, "t_b")) || (rc = dev_alloc(&p.t_key, t, "t_key")) ||
        (rc = dev_alloc(&p.t_parent, t, "t_parent")) || (rc = dev_alloc(&p.t_sbase, t, "t_sbase")) ||
        (rc = dev_alloc(&p.t_crefl, t, "t_crefl")) || (rc = dev_alloc(&p.t_crefr, t, "t_crefr")) ||
        (rc = dev_alloc(&p.t_root, t, "t_root")) || (rc = dev_alloc(&p.t_hitid, t, "t_hitid")) ||
        (rc = dev_alloc(&p.t_order, t, "t_order")) ||
        (rc = dev_alloc(&p.s_a, s, "s_a")) || (rc = dev_alloc(&p.s_b, s, "s_b")) || (rc = dev_alloc(&p.s_key, s, "s_key"))) {
        release_pools();
        return rc;
    }
    p.t_cap = (int)t_cap;
    p.s_cap = (int)s_cap;
    return RT_OK;
}

int Workspace::ensure_samples(long long n) {
    if (sample_cap >= n && sample_rgb) return RT_OK;
    cudaFree(sample_rgb);
    sample_cap = 0;
    int rc = dev_alloc(&sample_rgb, (size_t)n, "sample_rgb");
    if (rc == RT_OK) sample_cap = n;
    return rc;
}

int Workspace::ensure_rows(int n) {
    if (rows_cap >= n && rows) return RT_OK;
    cudaFree(rows);
    rows_cap = 0;
    int rc = dev_alloc(&rows, (size_t)std::max(n, 1), "rows");
    if (rc == RT_OK) rows_cap = std::max(n, 1);
    return rc;
}

int Workspace::ensure_mask(long long n) {
    if (mask_cap >= n && mask) return RT_OK;
    cudaFree(mask);
    mask_cap = 0;
    int rc = dev_alloc(&mask, (size_t)std::max<long long>(n, 1), "mask");
    if (rc == RT_OK) mask_cap = std::max<long long>(n, 1);
    return rc;
}

int Workspace::ensure_idx(long long n) {
    if (idx_cap >= n && idx) return RT_OK;
    cudaFree(idx);
    idx = nullptr; idx_cap = 0;
    int rc = dev_alloc(&idx, (size_t)std::max<long long>(n, 1) + 1, "idx"); // + the counter behind the list
    if (rc == RT_OK) idx_cap = std::max<long long>(n, 1);
    return rc;
}

int Workspace::ensure_batch_counters(int n) {
    n = std::max(n, BATCH_COUNTERS_MIN);
    if (batch_counters && batch_cap >= n) return RT_OK;
    cudaFree(batch_counters);
    batch_counters = nullptr; batch_cap = 0;
    int rc = dev_alloc(&batch_counters, (size_t)n, "batch_counters");
    if (rc == RT_OK) batch_cap = n;
    return rc;
}

int Workspace::ensure_misc() {
    int rc;
    if (!pools.counters && (rc = dev_alloc(&pools.counters, N_COUNTERS, "counters"))) return rc;
    if (!pools.bins && (rc = dev_alloc(&pools.bins, (size_t)node_bins(NODE_BIN_BITS_MAX) + 2, "bins"))) return rc;
    if (!pools.bin_sums && (rc = dev_alloc(&pools.bin_sums, NODE_BIN_CHUNKS_MAX, "bin_sums"))) return rc;
    pools.node_bits = tuning().node_bits == 0 ? 5 : std::min(std::max(tuning().node_bits, 3), NODE_BIN_BITS_MAX); // 0: by level size (cast_band)
    if (!exact_counter && (rc = dev_alloc(&exact_counter, N_SWEEP_COUNTERS, "exact_counter"))) return rc;
    if (!h_counters) {
        cudaError_t e = cudaMallocHost((void **)&h_counters, N_COUNTERS * sizeof(unsigned int) + N_SWEEP_COUNTERS * sizeof(unsigned long long));
        if (e != cudaSuccess) return cuda_fail(e, "cudaMallocHost");
    }
    return RT_OK;
}

// ---- the wavefront ------------------------------------------------------------------------

// Make sure the scene has min(want, 8) auxiliary streams; false if they cannot be had.
static bool ws_aux_streams(Workspace &ws, int want) {
    want = std::min(std::max(want, 1), (int)Workspace::MAX_AUX);
    while (ws.n_aux < want) {
        cudaStream_t s = nullptr;
        if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) {
            cudaGetLastError();
            return ws.n_aux > 0;
        }
        ws.aux[ws.n_aux++] = s;
    }
    return true;
}

// Events of one call, out of the workspace's store (created on first use, kept for the next call: every
// call ends with a synchronisation of its stream, so none is in flight when it is handed out again).
struct EventPool {
    Workspace &ws;
    int used = 0;
    std::vector<cudaEvent_t> extra; // beyond the store: created and destroyed by this call
    explicit EventPool(Workspace &w) : ws(w) {}
    cudaEvent_t get() {
        if (used < Workspace::MAX_EVENTS) {
            if (used == ws.n_events) {
                cudaEvent_t e = nullptr;
                cudaEventCreate(&e);
                ws.events[ws.n_events++] = e;
            }
            return ws.events[used++];
        }
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        extra.push_back(e);
        return e;
    }
    ~EventPool() {
        for (cudaEvent_t e : extra) cudaEventDestroy(e);
    }
};

struct BandLevels {
    std::vector<int> tb, te; // tree-ray range of level k
    std::vector<int> sb, se; // shadow-ray range of level k
    bool light_major = false; // shadow rays of a level laid out [light][node] (ordered spawn), else [node][light]
    unsigned long long nodes = 0;
};

struct Accum {
    unsigned long long tree_rays = 0, shadow_rays = 0, nodes = 0, exact = 0;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> sweep_spans;
    std::vector<long long> sweep_rays; // rays of each span (RT_B200_TRACE)
    std::vector<unsigned long long> sweep_counts; // device counters after each span: exact, general evals, origin evals (trace only)
    double filter_fma = 0.0;           // FMAs executed by the sweeps' filters
    unsigned long long slot_evals[MAX_SEGS] = {0, 0, 0, 0}; // (ray, bound slot) pairs the filters evaluated one by one, per table segment
    unsigned long long all_evals = 0;  // spheres of every level (slots, groups, quads) the filters evaluated
    unsigned long long exact_sphere = 0, exact_triangle = 0; // exact tests by kind (the rest of `exact` are cubes)
    bool audit = false;                // RT_FLAG_AUDIT_ALL_PAIRS: every pair went through the exact test
};

// Device time during which at least one sweep was running: the union of the spans (they overlap when a level's
// shadow sweep runs beside the next level's tree sweep), each placed on the time line of `origin`.
static float sweep_union_ms(const Accum &acc, cudaEvent_t origin, std::vector<float> *each = nullptr) {
    std::vector<std::pair<float, float>> iv;
    for (const auto &sp : acc.sweep_spans) {
        float a = 0.f, b = 0.f;
        if (cudaEventElapsedTime(&a, origin, sp.first) != cudaSuccess || cudaEventElapsedTime(&b, origin, sp.second) != cudaSuccess) {
            cudaGetLastError();
            continue;
        }
        if (each) each->push_back(b - a);
        iv.push_back({a, std::max(a, b)});
    }
    std::sort(iv.begin(), iv.end());
    float total = 0.f, end = -1e30f;
    for (const auto &x : iv) {
        if (x.first > end) { total += x.second - x.first; end = x.second; }
        else if (x.second > end) { total += x.second - end; end = x.second; }
    }
    return total;
}

// exact-test count and filter counts of the sweeps, from the device counters read back into
// ws.h_counters + N_COUNTERS: [0] exact tests, [1] spheres evaluated by the general filter, [2] by the
// origin filter (bounds, group and quad spheres alike), [3 + g] of those, evaluations of bound slots of
// segment g, [7], [8] exact tests of spheres / triangles
static void read_sweep_counters(const Workspace &ws, Accum &acc) {
    unsigned long long c[N_SWEEP_COUNTERS];
    std::memcpy(c, ws.h_counters + N_COUNTERS, sizeof(c));
    acc.exact = c[0];
    acc.filter_fma = (double)c[1] * FILTER_FMA_GENERAL + (double)c[2] * FILTER_FMA_ORIGIN;
    acc.all_evals = c[1] + c[2];
    for (int g = 0; g < MAX_SEGS; ++g) acc.slot_evals[g] = c[3 + g];
    acc.exact_sphere = c[3 + MAX_SEGS];
    acc.exact_triangle = c[4 + MAX_SEGS];
}

#define CK(call, what)                                        \
    do {                                                      \
        cudaError_t e__ = (call);                             \
        if (e__ != cudaSuccess) return cuda_fail(e__, what);  \
    } while (0)

// Cast every ray of samples [sample0, sample0 + n): fills the pools and `lv`.
// Returns RT_OK, an error, or +1 when a pool overflowed (caller grows it and retries).
static int cast_band(rt_scene *scene, Workspace &ws, const FrameParams &fp, long long sample0, int n, int flags,
                     BandLevels &lv, Accum &acc, EventPool &events) {
    const DevScene &sc = scene->dev;
    cudaStream_t st = scene->stream;
    Pools &p = ws.pools;
    const bool audit = (flags & RT_FLAG_AUDIT_ALL_PAIRS) != 0;
    const bool trace = tuning().trace != 0;
    lv = BandLevels();
    lv.light_major = !audit && !scene->light_tabs.empty() && scene->cam.bound; // = by_origin below: spawn runs ordered

    unsigned int init[N_COUNTERS] = {(unsigned int)n, 0u, 0u, 0u, 0u, 0u};
    CK(cudaMemcpyAsync(p.counters, init, sizeof(init), cudaMemcpyHostToDevice, st), "counters init");
    { // a counter per sweep launch: every level sweeps its tree rays and each light's shadow rays, chunk by chunk
        const long long need = (long long)(fp.depth + 2) * (1 + std::max(sc.n_point_lights, 0)) * sweep_table_chunks(sc);
        if (need > (1 << 24)) return fail(RT_ERR_UNSUPPORTED, "depth %d x %d lights x %d table chunks: too many sweep launches per band", fp.depth, sc.n_point_lights, sweep_table_chunks(sc));
        int rc = ws.ensure_batch_counters((int)need);
        if (rc != RT_OK) return rc;
    }
    CK(cudaMemsetAsync(ws.batch_counters, 0, (size_t)ws.batch_cap * sizeof(unsigned int), st), "batch counters");
    ws.batch_next = 0;
    cudaError_t sweep_err = cudaSuccess;
    launch_raygen(fp, p, sample0, n, st);
    scene->launches += 1;

    // `on`: the stream to launch on; `timed`: bracket this sweep with its own pair of events
    auto sweep_on = [&](RaySeg a, RaySeg b, OriginTable ot, cudaStream_t on, bool timed) {
        if (a.n + b.n == 0) return;
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (timed) {
            e0 = events.get(); e1 = events.get();
            cudaEventRecord(e0, on);
        }
        if (audit) launch_sweep_all_pairs(sc, a, b, ws.exact_counter, on);
        else {
            int used = 0;
            cudaError_t e = launch_sweep(sc, a, b, ot, scene->sm_count, ws.batch_counters + ws.batch_next,
                                         ws.batch_cap - ws.batch_next, &used, ws.exact_counter, on);
            ws.batch_next += used;
            scene->launches += used > 0 ? used - 1 : 0;
            if (e != cudaSuccess && sweep_err == cudaSuccess) sweep_err = e;
        }
        if (timed) {
            cudaEventRecord(e1, on);
            acc.sweep_spans.push_back({e0, e1});
            acc.sweep_rays.push_back((long long)a.n + b.n);
        }
        if (trace && timed) { // per-sweep counters: costs a synchronisation per sweep, tracing only
            unsigned long long c[N_SWEEP_COUNTERS] = {};
            cudaMemcpyAsync(c, ws.exact_counter, sizeof(c), cudaMemcpyDeviceToHost, st);
            cudaStreamSynchronize(st);
            acc.sweep_counts.insert(acc.sweep_counts.end(), c, c + 3);
        }
        scene->launches += 1;
        if (audit) acc.exact += (unsigned long long)(a.n + b.n) * (unsigned long long)sc.n_bounded;
    };
    auto sweep = [&](RaySeg a, RaySeg b, OriginTable ot) { sweep_on(a, b, ot, st, true); };
    auto tree_seg = [&](int b, int e) { return RaySeg{p.t_a + b, p.t_b + b, p.t_key + b, e - b}; };
    auto shadow_seg = [&](int b, int e) { return RaySeg{p.s_a + b, p.s_b + b, p.s_key + b, e - b}; };
    const RaySeg none{nullptr, nullptr, nullptr, 0};
    const OriginTable general{};
    // Origin tables: primary rays against the camera's table, the shadow rays of each light
    // (one column of the [node][light] pool) against that light's; the rest as one general sweep.
    const bool by_origin = !audit && !scene->light_tabs.empty() && scene->cam.bound;
    const int npl = std::max(sc.n_point_lights, 1);
    const bool merge_lights = tuning().merge_lights != 0;
    const int heavy_allow = tuning().heavy_allow;
    // heavy_tree / heavy_nodes: how many of the level's first tree rays / of the previous level's
    // first nodes start far outside the scene (counters[4], [5]); their sweeps get finer work items
    int aux_dirty = 0; // side streams that carry sweeps not yet joined into the main stream (pipelined levels)
    auto join_aux = [&]() {
        for (int i = 0; i < aux_dirty; ++i) {
            cudaEvent_t done = events.get();
            cudaEventRecord(done, ws.aux[i]);
            cudaStreamWaitEvent(st, done, 0);
        }
        aux_dirty = 0;
    };
    // whatever way this function is left (a CK() of a failed call included), no sweep stays behind on a side stream:
    // the caller may grow or release the pools next
    struct JoinOnExit {
        decltype(join_aux) &join;
        ~JoinOnExit() { join(); }
    } join_on_exit{join_aux};
    auto sweep_level = [&](int level, int tb, int te, int sb, int se, int heavy_tree, int heavy_nodes) {
        if (!by_origin) {
            sweep(tree_seg(tb, te), shadow_seg(sb, se), general);
            return;
        }
        RaySeg tree = tree_seg(tb, te);
        tree.heavy = std::min(heavy_tree, tree.n);
        if (se > sb) {
            cudaError_t e = cudaMemsetAsync(p.s_key + sb, 0xFF, sizeof(unsigned long long) * (size_t)(se - sb), st);
            if (e != cudaSuccess && sweep_err == cudaSuccess) sweep_err = e;
        }
        // The sweeps of one level (the tree rays and each light's shadow rays) do not depend on
        // each other: launched on separate streams, the next one's CTAs take over the SMs the
        // previous one's have left, instead of waiting for its last work item.  One timed span
        // per level then; tracing keeps them serial and timed one by one.
        const bool fork = tuning().trace != 1 && ws_aux_streams(ws, 1 + sc.n_point_lights);
        // Pipelined: nothing but the final shading waits for a level's shadow sweep, so it is not joined here -- it keeps
        // running on its side stream beside the spawn kernels and the next level's tree sweep (which fill the SMs its last
        // work items leave idle), and cast_band joins the side streams once, at its end.  The tree sweep stays on the main stream.
        const bool pipe = fork && !trace && tuning().level_pipeline != 0;
        const int n_aux = ws.n_aux;
        cudaEvent_t e0 = nullptr, e1 = nullptr, forked = nullptr;
        long long rays = 0;
        if (fork) {
            forked = events.get();
            if (!pipe) {
                e0 = events.get(); e1 = events.get();
                cudaEventRecord(e0, st);
            }
            cudaEventRecord(forked, st);
            for (int i = 0; i < n_aux; ++i) cudaStreamWaitEvent(ws.aux[i], forked, 0);
        }
        int next_aux = 0;
        auto stream_for = [&]() { return fork ? ws.aux[next_aux++ % n_aux] : st; };
        // tree rays first: measured against shadows first (c4: 51.4 vs 52.0 ms)
        sweep_on(tree, none, level == 0 ? scene->cam : general, pipe ? st : stream_for(), !fork || pipe);
        rays += tree.n;
        if (se > sb) {
            auto light_seg = [&](int l) {
                const int per_light = (se - sb) / npl; // [light][node]: light l's rays are one contiguous run
                RaySeg sh{p.s_a + sb + l * per_light, p.s_b + sb + l * per_light, p.s_key + sb + l * per_light, per_light, 1, 1};
                sh.heavy = std::min(heavy_nodes, sh.n);
                return sh;
            };
            // every light's shadow rays in one launch (a light has 1 / L of the CTAs and L times the
            // ray batches per warp: small levels need not cut the table into slices to fill the GPU)
            bool merged = false;
            if (fork && merge_lights && sc.n_point_lights >= 2) {
                std::vector<RaySeg> segs;
                for (int l = 0; l < sc.n_point_lights; ++l) segs.push_back(light_seg(l));
                int used = 0;
                cudaStream_t on = stream_for();
                cudaEvent_t j0 = nullptr, j1 = nullptr;
                if (pipe) { // its own span, on its own stream
                    j0 = events.get(); j1 = events.get();
                    cudaEventRecord(j0, on);
                }
                cudaError_t e = launch_sweep_jobs(sc, segs.data(), scene->light_tabs.data(), sc.n_point_lights, scene->sm_count,
                                                  ws.batch_counters + ws.batch_next, ws.batch_cap - ws.batch_next, &used,
                                                  ws.exact_counter, on);
                if (e == cudaSuccess) {
                    merged = true;
                    ws.batch_next += used;
                    scene->launches += 1;
                    long long shadow_rays = 0;
                    for (const RaySeg &sh : segs) shadow_rays += sh.n;
                    rays += shadow_rays;
                    if (pipe) {
                        cudaEventRecord(j1, on);
                        acc.sweep_spans.push_back({j0, j1});
                        acc.sweep_rays.push_back(shadow_rays);
                    }
                } else if (e != cudaErrorNotSupported && sweep_err == cudaSuccess) sweep_err = e;
                else --next_aux; // nothing was launched on that stream
            }
            for (int l = 0; !merged && l < sc.n_point_lights; ++l) {
                RaySeg sh = light_seg(l);
                sweep_on(sh, none, scene->light_tabs[l], stream_for(), !fork || pipe);
                rays += sh.n;
            }
        }
        if (pipe) {
            aux_dirty = std::max(aux_dirty, std::min(next_aux, n_aux));
        } else if (fork) {
            for (int i = 0; i < std::min(next_aux, n_aux); ++i) {
                cudaEvent_t done = events.get();
                cudaEventRecord(done, ws.aux[i]);
                cudaStreamWaitEvent(st, done, 0);
            }
            cudaEventRecord(e1, st);
            if (rays > 0) {
                acc.sweep_spans.push_back({e0, e1});
                acc.sweep_rays.push_back(rays);
                if (trace) { // trace 2: the level's counters
                    unsigned long long c[N_SWEEP_COUNTERS] = {};
                    cudaMemcpyAsync(c, ws.exact_counter, sizeof(c), cudaMemcpyDeviceToHost, st);
                    cudaStreamSynchronize(st);
                    acc.sweep_counts.insert(acc.sweep_counts.end(), c, c + 3);
                }
            }
        }
    };

    int tb = 0, te = n, sb = 0, se = 0; // current tree level, previous level's shadows
    int heavy_tree = 0, heavy_nodes = 0;
    int k = 0;
    for (; k < fp.depth; ++k) {
        sweep_level(k, tb, te, sb, se, heavy_tree, heavy_nodes);
        // cells of the node order: the more nodes a level has, the finer (a warp's 128 consecutive nodes then lie
        // closer together and its beam touches fewer groups); measured on c4: 7 bits at 8 M nodes, 5 below 250 k
        if (tuning().node_bits == 0) p.node_bits = (te - tb) >= (2 << 20) ? 7 : ((te - tb) >= (1 << 18) ? 6 : 5);
        launch_spawn(sc, fp, p, tb, te, (k + 1 < fp.depth) ? 1 : 0, by_origin ? 1 : 0, se, st);
        if (te > tb) scene->launches += by_origin ? 5 : 1;
        CK(cudaMemcpyAsync(ws.h_counters, p.counters, N_COUNTERS * sizeof(unsigned int), cudaMemcpyDeviceToHost, st), "counters readback");
        CK(cudaStreamSynchronize(st), "wavefront level");
        if (ws.h_counters[3]) { // pool overflow: the caller grows the pools and casts the band again
            join_aux();
            CK(cudaStreamSynchronize(st), "wavefront level");
            return 1;
        }
        lv.tb.push_back(tb); lv.te.push_back(te);
        int new_te = (int)ws.h_counters[0], new_se = (int)ws.h_counters[1];
        lv.sb.push_back(se); lv.se.push_back(new_se);
        sb = se; se = new_se;
        tb = te; te = new_te;
        // far nodes come first in the node order, and their children (allocated as spawn runs
        // through that order) early in the next level: allow for the blocks in flight
        heavy_nodes = by_origin ? (int)ws.h_counters[4] : 0;
        heavy_tree = (by_origin && ws.h_counters[5]) ? (int)std::min<long long>(2ll * ws.h_counters[5] + heavy_allow, te - tb) : 0;
        if (tb == te && sb == se) break; // nothing left to cast
    }
    sweep_level(-1, 0, 0, sb, se, 0, heavy_nodes); // shadows of the last level
    join_aux(); // shading reads every level's shadow keys
    lv.nodes = ws.h_counters[2];
    CK(sweep_err, "sweep launch");
    CK(cudaGetLastError(), "wavefront launch");
    return RT_OK;
}

// Shade levels deepest-first into ws.sample_rgb (band-local sample index).
static int shade_band(rt_scene *scene, const DevScene &sc, Workspace &ws, const FrameParams &fp, const BandLevels &lv,
                      const unsigned char *sample_mask) {
    for (int k = (int)lv.tb.size() - 1; k >= 0; --k) {
        const int npl = std::max(sc.n_point_lights, 1);
        launch_shade(sc, fp, ws.pools, lv.tb[k], lv.te[k], k, lv.light_major ? (lv.se[k] - lv.sb[k]) / npl : 1, ws.sample_rgb,
                     sample_mask, scene->stream);
        if (lv.te[k] > lv.tb[k]) scene->launches += 1;
    }
    CK(cudaGetLastError(), "shade launch");
    return RT_OK;
}

static long long worst_case_tree_rays(long long n, int depth) {
    if (depth >= 31) return LLONG_MAX / 4;
    long long f = (1ll << depth) - 1;
    if (f <= 0) f = 1;
    if (n > LLONG_MAX / 4 / f) return LLONG_MAX / 4;
    return n * f;
}


static void fill_stats(rt_scene *scene, const Accum &acc, unsigned long long pixels, rt_stats *st) {
    std::memset(st, 0, sizeof(*st));
    const HostScene &hs = *scene->host;
    const int *c = hs.counts;
    unsigned long long rays = acc.tree_rays + acc.shadow_rays;
    st->pixels = pixels;
    st->rays = rays;
    st->shadow_rays = acc.shadow_rays;
    st->nodes = acc.nodes;
    st->sphere_tests = rays * (unsigned long long)c[RT_SHAPE_SPHERE];
    st->plane_tests = rays * (unsigned long long)c[RT_SHAPE_PLANE];
    st->triangle_tests = rays * (unsigned long long)c[RT_SHAPE_TRIANGLE];
    st->cube_tests = rays * (unsigned long long)c[RT_SHAPE_CUBE];
    // sphere 54, triangle 30, plane 47 (14 when inv is the identity), cube 33 + 12*30,
    // node 60 + 70 L
    const double plane_flop = 47.0 * (c[RT_SHAPE_PLANE] - hs.planes_identity) + 14.0 * hs.planes_identity;
    double per_ray = 54.0 * c[RT_SHAPE_SPHERE] + 30.0 * c[RT_SHAPE_TRIANGLE] + 393.0 * c[RT_SHAPE_CUBE] + plane_flop;
    st->algorithmic_flops = (double)rays * per_ray + (double)acc.nodes * (60.0 + 70.0 * scene->dev.n_lights);
    st->filter_flops = 2.0 * acc.filter_fma;
    if (acc.audit) { // every pair through the exact test, no filter
        st->exact_sphere_tests = st->sphere_tests;
        st->exact_triangle_tests = st->triangle_tests;
        st->exact_cube_tests = st->cube_tests;
        st->exact_tests = rays * (unsigned long long)scene->dev.n_shapes;
        st->tested_pairs = st->exact_tests;
    } else {
        // pairs whose own bound the filter evaluated, by kind: the segment's evaluations with its pads scaled out
        unsigned long long slot_pairs = 0, slot_all = 0;
        for (int g = 0; g < scene->dev.segs.n; ++g) {
            const int slots = scene->dev.segs.end[g] - scene->dev.segs.begin[g];
            const unsigned long long real = slots ? (unsigned long long)((double)acc.slot_evals[g] * hs.seg_real[g] / slots) : 0ull;
            if (hs.seg_kind[g] == RT_SHAPE_SPHERE) st->bound_evals_sphere += real;
            else if (hs.seg_kind[g] == RT_SHAPE_TRIANGLE) st->bound_evals_triangle += real;
            else st->bound_evals_cube += real;
            slot_pairs += real;
            slot_all += acc.slot_evals[g];
        }
        st->hier_evals = acc.all_evals - slot_all;
        st->exact_sphere_tests = acc.exact_sphere;
        st->exact_triangle_tests = acc.exact_triangle;
        st->exact_cube_tests = acc.exact - acc.exact_sphere - acc.exact_triangle;
        st->exact_tests = acc.exact + st->plane_tests;
        // pairs resolved one by one: a bound-slot evaluation or an exact test (every plane)
        st->tested_pairs = slot_pairs + st->plane_tests;
    }
    st->exact_flops = 54.0 * (double)st->exact_sphere_tests + 30.0 * (double)st->exact_triangle_tests +
                      393.0 * (double)st->exact_cube_tests + (double)rays * plane_flop;
    st->executed_flops = st->filter_flops + st->exact_flops;
}

int check_params(const rt_scene *scene, const rt_camera *cam, const rt_render_params *pr) {
    if (!scene || !cam || !pr) return fail(RT_ERR_INVALID, "null argument");
    if (cam->x_res <= 0 || cam->y_res <= 0) return fail(RT_ERR_INVALID, "resolution %dx%d", cam->x_res, cam->y_res);
    if (pr->depth < 0 || pr->depth > 30) return fail(RT_ERR_INVALID, "depth %d not in 0..30", pr->depth);
    if (pr->spp != 1 && pr->spp != 4) return fail(RT_ERR_UNSUPPORTED, "spp must be 1 or 4, got %d", pr->spp);
    if (pr->world < 1 || pr->rank < 0 || pr->rank >= pr->world || pr->tile_rows < 1)
        return fail(RT_ERR_INVALID, "bad row tiling rank %d world %d tile_rows %d", pr->rank, pr->world, pr->tile_rows);
    return RT_OK;
}

static FrameParams frame_params(const rt_camera *cam, const rt_render_params *pr, const int *d_rows, int n_rows) {
    FrameParams fp;
    fp.cam_ox = cam->origin[0]; fp.cam_oy = cam->origin[1]; fp.cam_oz = cam->origin[2];
    fp.x_min = cam->x_min; fp.x_max = cam->x_max; fp.y_min = cam->y_min; fp.y_max = cam->y_max;
    fp.x_res = cam->x_res; fp.y_res = cam->y_res;
    fp.depth = pr->depth; fp.spp = pr->spp; fp.seed = pr->seed;
    fp.rows = d_rows; fp.n_rows = n_rows;
    fp.forest_shading = (pr->flags & RT_FLAG_FOREST_SHADING) ? 1 : 0;
    fp.primary_tiles = tuning().primary_tiles;
    return fp;
}

static int upload_rows(rt_scene *scene, Workspace &ws, const rt_camera *cam, const rt_render_params *pr,
                       std::vector<int> &rows) {
    rows.clear();
    for (int v = 0; v < cam->y_res; ++v)
        if ((v / pr->tile_rows) % pr->world == pr->rank) rows.push_back(v);
    int rc = ws.ensure_rows((int)rows.size());
    if (rc != RT_OK) return rc;
    if (!rows.empty())
        CK(cudaMemcpyAsync(ws.rows, rows.data(), rows.size() * sizeof(int), cudaMemcpyHostToDevice, scene->stream), "rows upload");
    return RT_OK;
}

// Grow-and-retry wrapper around cast_band.
static int cast_band_sized(rt_scene *scene, Workspace &ws, const FrameParams &fp, long long sample0, int n, int flags,
                           BandLevels &lv, Accum &acc, EventPool &events) {
    const long long worst = worst_case_tree_rays(n, std::max(fp.depth, 1));
    const int npl = std::max(scene->dev.n_point_lights, 0);
    long long t_cap = std::max<long long>(ws.pools.t_cap, std::min<long long>(worst, std::max<long long>((long long)n * 3, 1 << 16)));
    const size_t spans0 = acc.sweep_spans.size(), counts0 = acc.sweep_counts.size();
    for (;;) {
        int rc = ws.ensure_pools(t_cap, t_cap * npl);
        if (rc != RT_OK) return rc;
        rc = cast_band(scene, ws, fp, sample0, n, flags, lv, acc, events);
        if (rc <= 0) return rc;
        cudaMemsetAsync(ws.exact_counter, 0, N_SWEEP_COUNTERS * sizeof(unsigned long long), scene->stream); // the band is recast:
        acc.sweep_spans.resize(spans0); acc.sweep_rays.resize(spans0); acc.sweep_counts.resize(counts0); // its aborted sweeps do not count
        if (t_cap >= worst) return fail(RT_ERR_INTERNAL, "ray pool overflow at the worst-case size");
        t_cap = std::min(worst, t_cap * 2);
    }
}

// Device staging buffers of n_ch channels for a host-buffer call (kept with the workspace).
int ensure_staging(Workspace &ws, size_t n_ch, bool want_f32, bool want_u8) {
    int rc;
    if (want_f32 && ws.out_f32_cap < n_ch * sizeof(float)) {
        cudaFree(ws.out_f32);
        ws.out_f32 = nullptr; ws.out_f32_cap = 0;
        float *p;
        if ((rc = dev_alloc(&p, n_ch, "out_f32")) != RT_OK) return rc;
        ws.out_f32 = p; ws.out_f32_cap = n_ch * sizeof(float);
    }
    if (want_u8 && ws.out_u8_cap < n_ch) {
        cudaFree(ws.out_u8);
        ws.out_u8 = nullptr; ws.out_u8_cap = 0;
        unsigned char *p;
        if ((rc = dev_alloc(&p, n_ch, "out_u8")) != RT_OK) return rc;
        ws.out_u8 = p; ws.out_u8_cap = n_ch;
    }
    return RT_OK;
}

int render_impl(rt_scene *scene, const rt_camera *cam, const rt_render_params *pr, Outputs out, rt_stats *stats) {
    int rc = check_params(scene, cam, pr);
    if (rc != RT_OK) return rc;
    CK(cudaSetDevice(scene->device), "cudaSetDevice");
    HostLap lap("render");
    if ((rc = prepare_camera_table(scene, cam->origin)) != RT_OK) return rc;
    lap("camera table");
    Workspace &ws = scene->ws;
    if ((rc = ws.ensure_misc()) != RT_OK) return rc;
    cudaStream_t st = scene->stream;
    std::vector<int> rows;
    if ((rc = upload_rows(scene, ws, cam, pr, rows)) != RT_OK) return rc;
    lap("rows");
    const long long n_pixels = (long long)rows.size() * cam->x_res;
    const long long n_samples = n_pixels * pr->spp;
    FrameParams fp = frame_params(cam, pr, ws.rows, (int)rows.size());

    EventPool events(ws);
    cudaEvent_t ev0 = events.get(), ev1 = events.get();
    scene->launches = 0;
    CK(cudaMemsetAsync(ws.exact_counter, 0, N_SWEEP_COUNTERS * sizeof(unsigned long long), st), "exact counter");
    cudaEventRecord(ev0, st);

    Accum acc;
    long long band = tuning().band_samples;
    if (n_samples > 0) {
        long long nb = (n_samples + band - 1) / band;
        band = (n_samples + nb - 1) / nb;
        band = (band + pr->spp - 1) / pr->spp * pr->spp; // whole pixels per band
    }
    if ((rc = ws.ensure_samples(std::min(band, std::max<long long>(n_samples, 1)))) != RT_OK) return rc;
    for (long long s0 = 0; s0 < n_samples; s0 += band) {
        int n = (int)std::min(band, n_samples - s0);
        if (fp.depth == 0) {
            CK(cudaMemsetAsync(ws.sample_rgb, 0, (size_t)n * sizeof(float4), st), "black frame"); // render.rs:43-45
        } else {
            BandLevels lv;
            if ((rc = cast_band_sized(scene, ws, fp, s0, n, pr->flags, lv, acc, events)) != RT_OK) return rc;
            acc.tree_rays += (unsigned long long)(lv.te.empty() ? 0 : lv.te.back());
            acc.shadow_rays += (unsigned long long)(lv.se.empty() ? 0 : lv.se.back());
            acc.nodes += lv.nodes;
            if ((rc = shade_band(scene, scene->dev, ws, fp, lv, nullptr)) != RT_OK) return rc;
        }
        launch_resolve(ws.sample_rgb, s0 / pr->spp, n / pr->spp, pr->spp, out.d_f32, out.d_u8, nullptr,
                       (pr->flags & RT_FLAG_FRAME_ROWS) ? ws.rows : nullptr, cam->x_res, st);
        if (n > 0) scene->launches += 1;
    }
    cudaEventRecord(ev1, st);
    CK(cudaMemcpyAsync(ws.h_counters + N_COUNTERS, ws.exact_counter, N_SWEEP_COUNTERS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st), "exact readback");
    CK(cudaStreamSynchronize(st), "render");
    CK(cudaGetLastError(), "render");
    lap("device work");
    acc.audit = (pr->flags & RT_FLAG_AUDIT_ALL_PAIRS) != 0;
    if (!acc.audit) read_sweep_counters(ws, acc);
    if (stats) {
        fill_stats(scene, acc, (unsigned long long)n_pixels, stats);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ev0, ev1);
        const bool trace = tuning().trace != 0;
        std::vector<float> each;
        const float sweep = sweep_union_ms(acc, ev0, &each);
        for (size_t i = 0; i < acc.sweep_spans.size() && i < each.size(); ++i) {
            const float t = each[i];
            if (trace)
            {
                double ex = 0, ev = 0;
                if (acc.sweep_counts.size() >= 3 * (i + 1)) {
                    const unsigned long long *c = &acc.sweep_counts[3 * i], zero[3] = {0, 0, 0}, *q = i ? c - 3 : zero;
                    ex = (double)(c[0] - q[0]); ev = (double)(c[1] - q[1]) + (double)(c[2] - q[2]);
                    if (c[0] < q[0]) { ex = (double)c[0]; ev = (double)c[1] + (double)c[2]; } // counters were reset (band recast)
                }
                std::fprintf(stderr, "[rt_b200] sweep %zu: %lld rays %.3f ms (%.2f Tpairs/s) evals/ray %.0f exact/ray %.1f\n", i,
                             acc.sweep_rays[i], t, (double)acc.sweep_rays[i] * scene->dev.n_bounded / (t * 1e-3) / 1e12,
                             ev / (double)acc.sweep_rays[i], ex / (double)acc.sweep_rays[i]);
            }
        }
        if (trace) std::fprintf(stderr, "[rt_b200] frame %.3f ms, sweeps %.3f ms\n", ms, sweep);
        stats->total_ms = ms;
        stats->sweep_ms = sweep;
        stats->shade_ms = ms - sweep;
        stats->kernel_launches = scene->launches;
    }
    return RT_OK;
}

} // namespace rt

using namespace rt;

extern "C" {

int rt_render_device(rt_scene *scene, const rt_camera *camera, const rt_render_params *params, void *d_rgb_f32,
                     void *d_rgb_u8, rt_stats *stats) {
    Outputs out;
    out.d_f32 = (float *)d_rgb_f32;
    out.d_u8 = (unsigned char *)d_rgb_u8;
    return render_impl(scene, camera, params, out, stats);
}

int rt_render(rt_scene *scene, const rt_camera *camera, const rt_render_params *params, float *rgb_f32,
              uint8_t *rgb_u8, rt_stats *stats) {
    int rc = check_params(scene, camera, params);
    if (rc != RT_OK) return rc;
    if (params->flags & RT_FLAG_FRAME_ROWS) return fail(RT_ERR_INVALID, "RT_FLAG_FRAME_ROWS is for rt_render_device");
    CK(cudaSetDevice(scene->device), "cudaSetDevice");
    Workspace &ws = scene->ws;
    int32_t n_rows = rt_tile_rows_for_rank(camera->y_res, params->rank, params->world, params->tile_rows, nullptr, 0);
    size_t n_ch = (size_t)n_rows * (size_t)camera->x_res * 3;
    if ((rc = ensure_staging(scene->ws, n_ch, rgb_f32 != nullptr, rgb_u8 != nullptr)) != RT_OK) return rc;
    Outputs out;
    out.d_f32 = rgb_f32 ? (float *)ws.out_f32 : nullptr;
    out.d_u8 = rgb_u8 ? (unsigned char *)ws.out_u8 : nullptr;
    if ((rc = render_impl(scene, camera, params, out, stats)) != RT_OK) return rc;
    HostLap lap("render");
    if (rgb_f32 && n_ch) CK(cudaMemcpy(rgb_f32, ws.out_f32, n_ch * sizeof(float), cudaMemcpyDeviceToHost), "D2H rgb_f32");
    if (rgb_u8 && n_ch) CK(cudaMemcpy(rgb_u8, ws.out_u8, n_ch, cudaMemcpyDeviceToHost), "D2H rgb_u8");
    lap("copy to the host");
    return RT_OK;
}

// ---- RayForest ------------------------------------------------------------------------------

int rt_forest_build(rt_scene *scene, const rt_camera *camera, const rt_render_params *params, rt_forest **out,
                    rt_stats *stats) {
    if (!out) return fail(RT_ERR_INVALID, "null argument");
    *out = nullptr;
    int rc = check_params(scene, camera, params);
    if (rc != RT_OK) return rc;
    if (params->spp != 1) return fail(RT_ERR_UNSUPPORTED, "the ray forest keeps one tree per pixel: spp must be 1");
    if (params->depth < 1) return fail(RT_ERR_INVALID, "forest depth must be >= 1");
    CK(cudaSetDevice(scene->device), "cudaSetDevice");
    if ((rc = prepare_camera_table(scene, camera->origin)) != RT_OK) return rc;
    rt_forest *f = new (std::nothrow) rt_forest();
    if (!f) return fail(RT_ERR_NOMEM, "out of host memory");
    f->scene = scene;
    scene->forests += 1; // the forest uses the scene's tables and stream until it is destroyed
    f->ws = workspace_take_cached(scene->device);
    Workspace &ws = f->ws;
    std::vector<int> rows;
    if ((rc = ws.ensure_misc()) != RT_OK || (rc = upload_rows(scene, ws, camera, params, rows)) != RT_OK) {
        rt_forest_destroy(f);
        return rc;
    }
    long long n = (long long)rows.size() * camera->x_res;
    if (n > (1 << 28)) {
        rt_forest_destroy(f);
        return fail(RT_ERR_NOMEM, "forest of %lld trees does not fit one device-resident pool", n);
    }
    rt_render_params pr = *params;
    pr.flags |= RT_FLAG_FOREST_SHADING;
    f->fp = frame_params(camera, &pr, ws.rows, (int)rows.size());
    f->n_samples = (int)n;
    f->width = camera->x_res;
    f->n_rows = (int)rows.size();
    EventPool events(ws);
    cudaEvent_t ev0 = events.get(), ev1 = events.get();
    cudaStream_t st = scene->stream;
    cudaMemsetAsync(ws.exact_counter, 0, N_SWEEP_COUNTERS * sizeof(unsigned long long), st);
    cudaEventRecord(ev0, st);
    Accum acc;
    BandLevels lv;
    if (n > 0 && (rc = cast_band_sized(scene, ws, f->fp, 0, (int)n, pr.flags, lv, acc, events)) != RT_OK) {
        rt_forest_destroy(f);
        return rc;
    }
    cudaEventRecord(ev1, st);
    cudaMemcpyAsync(ws.h_counters + N_COUNTERS, ws.exact_counter, N_SWEEP_COUNTERS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
        rt_forest_destroy(f);
        return cuda_fail(e, "forest build");
    }
    f->tb = lv.tb;
    f->te = lv.te;
    f->sb = lv.sb;
    f->se = lv.se;
    f->light_major = lv.light_major;
    f->total_tree_rays = lv.te.empty() ? 0 : lv.te.back();
    if (stats) {
        acc.tree_rays = (unsigned long long)f->total_tree_rays;
        acc.shadow_rays = lv.se.empty() ? 0 : (unsigned long long)lv.se.back();
        acc.nodes = lv.nodes;
        acc.audit = (pr.flags & RT_FLAG_AUDIT_ALL_PAIRS) != 0;
        if (!acc.audit) read_sweep_counters(ws, acc);
        fill_stats(scene, acc, (unsigned long long)n, stats);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ev0, ev1);
        const float sweep = sweep_union_ms(acc, ev0);
        stats->total_ms = ms; stats->sweep_ms = sweep; stats->shade_ms = ms - sweep;
    }
    *out = f;
    return RT_OK;
}

void rt_forest_destroy(rt_forest *forest) {
    if (!forest) return;
    rt_scene *scene = forest->scene;
    if (scene) {
        cudaSetDevice(scene->device);
        workspace_give_cached(scene->device, forest->ws);
        scene->forests -= 1;
        if (scene->forests == 0 && scene->destroy_pending) scene_free(scene); // rt_scene_destroy came first
    } else {
        forest->ws.release();
    }
    delete forest;
}

int rt_trim(int32_t device) {
    if (device < 0) {
        drop_scene_cache();
        drop_worker_pool();
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
    for (int d = 0; d < ndev; ++d) {
        if (device >= 0 && d != device) continue;
        Workspace w = workspace_take_cached(d);
        CK(cudaSetDevice(d), "cudaSetDevice");
        w.release();
    }
    return RT_OK;
}

// shared by rt_forest_eval / rt_forest_eval_filtered
static int forest_eval_impl(rt_forest *f, const float *ambient, const int32_t *ids, int32_t n_ids, float *rgb_f32,
                            uint8_t *rgb_u8, uint64_t *n_evaluated, rt_stats *stats) {
    if (!f) return fail(RT_ERR_INVALID, "null forest");
    if (ids == nullptr && n_ids != 0) return fail(RT_ERR_INVALID, "null id list");
    rt_scene *scene = f->scene;
    CK(cudaSetDevice(scene->device), "cudaSetDevice");
    Workspace &ws = f->ws;
    cudaStream_t st = scene->stream;
    const int n = f->n_samples;
    const bool filtered = ids != nullptr;
    int rc;
    if ((rc = ws.ensure_samples(std::max(n, 1))) != RT_OK) return rc;
    DevScene sc = scene->dev; // current materials; ambient as passed by the caller
    if (ambient) { sc.amb_r = ambient[0]; sc.amb_g = ambient[1]; sc.amb_b = ambient[2]; }
    EventPool events(ws);
    cudaEvent_t ev0 = events.get(), ev1 = events.get();
    cudaEventRecord(ev0, st);
    const unsigned char *mask = nullptr;
    uint64_t evaluated = (uint64_t)n;
    struct Temps { // freed on every way out
        int *d_ids = nullptr;
        ~Temps() { cudaFree(d_ids); }
    } tmp;
    int *&d_ids = tmp.d_ids;
    if (filtered) {
        if ((rc = ws.ensure_mask(n)) != RT_OK) return rc;
        CK(cudaMemsetAsync(ws.mask, 0, (size_t)std::max(n, 1), st), "mask clear");
        if (n_ids > 0) {
            CK(cudaMalloc((void **)&d_ids, sizeof(int) * n_ids), "cudaMalloc(ids)");
            CK(cudaMemcpyAsync(d_ids, ids, sizeof(int) * n_ids, cudaMemcpyHostToDevice, st), "ids upload");
            launch_forest_mark(ws.pools, f->total_tree_rays, d_ids, n_ids, ws.mask, st);
        }
        mask = ws.mask;
    }
    BandLevels lv;
    lv.tb = f->tb;
    lv.te = f->te;
    lv.sb = f->sb;
    lv.se = f->se;
    lv.light_major = f->light_major;
    if ((rc = shade_band(scene, sc, ws, f->fp, lv, mask)) != RT_OK) return rc;
    const size_t n_ch = (size_t)n * 3;
    // device staging kept with the forest's workspace: no allocation per call
    if ((rc = ensure_staging(ws, std::max<size_t>(n_ch, 1), rgb_f32 != nullptr, rgb_u8 != nullptr)) != RT_OK) return rc;
    float *d_f32 = rgb_f32 ? (float *)ws.out_f32 : nullptr;
    unsigned char *d_u8 = rgb_u8 ? (unsigned char *)ws.out_u8 : nullptr;
    if (!filtered) {
        launch_resolve(ws.sample_rgb, 0, n, 1, d_f32, d_u8, nullptr, nullptr, 0, st);
        cudaEventRecord(ev1, st);
        if (d_f32 && n_ch) CK(cudaMemcpyAsync(rgb_f32, d_f32, n_ch * sizeof(float), cudaMemcpyDeviceToHost, st), "D2H");
        if (d_u8 && n_ch) CK(cudaMemcpyAsync(rgb_u8, d_u8, n_ch, cudaMemcpyDeviceToHost, st), "D2H");
        CK(cudaStreamSynchronize(st), "forest eval");
    } else {
        // render_forest_filter leaves the pixels of untouched trees as they are (render_tree.rs:129-145): only
        // the re-shaded pixels travel -- compacted on the device as (pixel number, colour), scattered into the
        // caller's buffers here.  Nothing of the caller's image goes to the device.
        if ((rc = ws.ensure_idx(n)) != RT_OK) return rc;
        unsigned int *d_count = (unsigned int *)(ws.idx + ws.idx_cap);
        CK(cudaMemsetAsync(d_count, 0, sizeof(unsigned int), st), "count clear");
        launch_resolve_compact(ws.sample_rgb, n, mask, ws.idx, d_count, d_f32, d_u8, st);
        cudaEventRecord(ev1, st);
        unsigned int count = 0;
        CK(cudaMemcpyAsync(&count, d_count, sizeof(count), cudaMemcpyDeviceToHost, st), "D2H count");
        CK(cudaStreamSynchronize(st), "forest eval");
        evaluated = count;
        if (count > 0 && (rgb_f32 || rgb_u8)) {
            std::vector<int> idx(count);
            std::vector<float> cf;
            std::vector<unsigned char> cu;
            CK(cudaMemcpyAsync(idx.data(), ws.idx, sizeof(int) * count, cudaMemcpyDeviceToHost, st), "D2H idx");
            if (rgb_f32) {
                cf.resize((size_t)count * 3);
                CK(cudaMemcpyAsync(cf.data(), d_f32, sizeof(float) * 3 * count, cudaMemcpyDeviceToHost, st), "D2H");
            }
            if (rgb_u8) {
                cu.resize((size_t)count * 3);
                CK(cudaMemcpyAsync(cu.data(), d_u8, (size_t)3 * count, cudaMemcpyDeviceToHost, st), "D2H");
            }
            CK(cudaStreamSynchronize(st), "forest eval");
            for (unsigned int k = 0; k < count; ++k) {
                const size_t o = (size_t)idx[k] * 3;
                if (rgb_f32) { rgb_f32[o] = cf[3 * k]; rgb_f32[o + 1] = cf[3 * k + 1]; rgb_f32[o + 2] = cf[3 * k + 2]; }
                if (rgb_u8) { rgb_u8[o] = cu[3 * k]; rgb_u8[o + 1] = cu[3 * k + 1]; rgb_u8[o + 2] = cu[3 * k + 2]; }
            }
        }
    }
    CK(cudaGetLastError(), "forest eval");
    if (n_evaluated) *n_evaluated = evaluated;
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->pixels = evaluated;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ev0, ev1);
        stats->total_ms = ms;
        stats->shade_ms = ms;
    }
    return RT_OK;
}

int rt_forest_eval(rt_forest *forest, const float *ambient, float *rgb_f32, uint8_t *rgb_u8, rt_stats *stats) {
    return forest_eval_impl(forest, ambient, nullptr, 0, rgb_f32, rgb_u8, nullptr, stats);
}

int rt_forest_eval_filtered(rt_forest *forest, const float *ambient, const int32_t *mutated_ids, int32_t n_ids,
                            float *rgb_f32_inout, uint8_t *rgb_u8_inout, uint64_t *n_evaluated, rt_stats *stats) {
    static const int32_t none = 0;
    if (n_ids < 0) return fail(RT_ERR_INVALID, "negative id count");
    return forest_eval_impl(forest, ambient, mutated_ids ? mutated_ids : &none, mutated_ids ? n_ids : 0, rgb_f32_inout,
                            rgb_u8_inout, n_evaluated, stats);
}

int rt_forest_size(rt_forest *forest, uint64_t *out) { // RayForest::size (render_tree.rs:62-64)
    if (!forest || !out) return fail(RT_ERR_INVALID, "null argument");
    *out = (uint64_t)forest->n_samples;
    return RT_OK;
}

static int forest_tree_sizes(rt_forest *f, std::vector<int> &sizes) {
    rt_scene *scene = f->scene;
    CK(cudaSetDevice(scene->device), "cudaSetDevice");
    const int n = f->n_samples;
    sizes.assign(n, 0);
    if (n == 0) return RT_OK;
    int *d = nullptr;
    CK(cudaMalloc((void **)&d, sizeof(int) * n), "cudaMalloc(sizes)");
    cudaMemsetAsync(d, 0, sizeof(int) * n, scene->stream);
    launch_forest_sizes(f->ws.pools, f->total_tree_rays, d, scene->stream);
    cudaError_t e = cudaMemcpyAsync(sizes.data(), d, sizeof(int) * n, cudaMemcpyDeviceToHost, scene->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(scene->stream);
    cudaFree(d);
    if (e != cudaSuccess) return cuda_fail(e, "forest sizes");
    return RT_OK;
}

int rt_forest_get_stats(rt_forest *forest, rt_forest_stats *out) { // RayForest::stats (render_tree.rs:73-93)
    if (!forest || !out) return fail(RT_ERR_INVALID, "null argument");
    std::vector<int> sizes;
    int rc = forest_tree_sizes(forest, sizes);
    if (rc != RT_OK) return rc;
    if (sizes.empty()) return fail(RT_ERR_INVALID, "empty forest"); // the reference unwraps max() of an empty list
    std::sort(sizes.begin(), sizes.end());
    size_t n = sizes.size();
    uint64_t total = 0;
    for (int s : sizes) total += (uint64_t)s;
    out->num_trees = n;
    out->num_intersections = total;
    out->smallest_tree = (uint64_t)sizes.front();
    out->largest_tree = (uint64_t)sizes.back();
    out->median = (uint64_t)sizes[n / 2];
    out->p90 = (uint64_t)sizes[(size_t)(0.9f * (float)n)];
    out->p95 = (uint64_t)sizes[(size_t)(0.95f * (float)n)];
    out->p99 = (uint64_t)sizes[(size_t)(0.99f * (float)n)];
    return RT_OK;
}

int rt_forest_trees_with(rt_forest *forest, int32_t shape_id, uint64_t *out) { // render_tree.rs:68-70
    if (!forest || !out) return fail(RT_ERR_INVALID, "null argument");
    rt_scene *scene = forest->scene;
    CK(cudaSetDevice(scene->device), "cudaSetDevice");
    Workspace &ws = forest->ws;
    const int n = forest->n_samples;
    int rc = ws.ensure_mask(n);
    if (rc != RT_OK) return rc;
    int *d_id = nullptr;
    CK(cudaMalloc((void **)&d_id, sizeof(int)), "cudaMalloc(id)");
    cudaMemcpyAsync(d_id, &shape_id, sizeof(int), cudaMemcpyHostToDevice, scene->stream);
    cudaMemsetAsync(ws.mask, 0, (size_t)std::max(n, 1), scene->stream);
    launch_forest_mark(ws.pools, forest->total_tree_rays, d_id, 1, ws.mask, scene->stream);
    std::vector<unsigned char> h(std::max(n, 1));
    cudaError_t e = cudaMemcpyAsync(h.data(), ws.mask, (size_t)std::max(n, 1), cudaMemcpyDeviceToHost, scene->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(scene->stream);
    cudaFree(d_id);
    if (e != cudaSuccess) return cuda_fail(e, "trees_with");
    uint64_t c = 0;
    for (int i = 0; i < n; ++i) c += h[i] ? 1 : 0;
    *out = c;
    return RT_OK;
}

int rt_measure_fp32_peak(int32_t device, double *tflops_measured, int32_t *sm_count, int32_t *sm_clock_khz) {
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
    if (device < 0 || device >= ndev) return fail(RT_ERR_CUDA, "CUDA device %d not present", device);
    CK(cudaSetDevice(device), "cudaSetDevice");
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties");
    int clock_khz = 0;
    cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, device);
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (sm_clock_khz) *sm_clock_khz = clock_khz;
    const int blocks = prop.multiProcessorCount * 8;
    float *sink = nullptr;
    CK(cudaMalloc((void **)&sink, (size_t)blocks * 256 * sizeof(float)), "cudaMalloc(sink)");
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) { // first reps warm up clocks; keep the best
        cudaEventRecord(e0, 0);
        double flops = launch_fp32_peak(blocks, 4096, sink, 0);
        cudaEventRecord(e1, 0);
        if ((e = cudaEventSynchronize(e1)) != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms > 0.f) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    if (e != cudaSuccess) return cuda_fail(e, "fp32 peak kernel");
    if (tflops_measured) *tflops_measured = best;
    return RT_OK;
}

} // extern "C"
