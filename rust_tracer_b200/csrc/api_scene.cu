// api_scene.cu -- C ABI: library, scene upload, material update (include/rt_b200.h).
// Turns the reference-shaped scene description (array of shapes in add order) into the
// structure-of-arrays device tables of scene_dev.cuh, and builds the conservative
// bounding spheres the sweep filters with.
#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "api_internal.h"
#include "../../include/rt_b200_debug.h"

namespace rt {

static thread_local char g_err[512] = "";

Tuning &tuning() {
    static Tuning t;
    return t;
}

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int cuda_fail(cudaError_t e, const char *what) {
    return fail(RT_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}

// ---- host-side data parallelism ----------------------------------------------------------
// The scene build walks a few thousand to a few hundred thousand records several times
// (slot order, group spheres, one re-centred table per light and per camera position); with a
// 50 ms frame behind it that is worth a handful of threads (tuning.h: host_threads).
constexpr int MAX_HOST_CHUNKS = 8;

static int host_threads() {
    static const int hw = (int)std::thread::hardware_concurrency();
    int want = tuning().host_threads;
    if (hw > 0) want = std::min(want, hw);
    return std::max(1, std::min(want, MAX_HOST_CHUNKS));
}

// fn(chunk, begin, end) over [0, n) in at most host_threads() chunks of at least min_chunk.
template <typename F> static void parallel_chunks(size_t n, size_t min_chunk, F &&fn) {
    size_t nt = std::min<size_t>((size_t)host_threads(), n / std::max<size_t>(min_chunk, 1));
    if (nt <= 1) {
        fn(0, (size_t)0, n);
        return;
    }
    const size_t step = (n + nt - 1) / nt;
    std::vector<std::thread> th;
    size_t started = 1;
    try {
        for (size_t t = 1; t < nt; ++t) {
            th.emplace_back([&fn, t, step, n] { fn((int)t, std::min(n, t * step), std::min(n, (t + 1) * step)); });
            ++started;
        }
    } catch (...) { // no more threads to be had: the rest runs here
    }
    fn(0, (size_t)0, std::min(n, step));
    for (auto &x : th) x.join();
    for (size_t t = started; t < nt; ++t) fn((int)t, std::min(n, t * step), std::min(n, (t + 1) * step));
}

// ---- bounding spheres (host, double precision) ------------------------------------------

static bool inv3(const double a[9], double out[9]) {
    double c00 = a[4] * a[8] - a[5] * a[7], c01 = a[5] * a[6] - a[3] * a[8], c02 = a[3] * a[7] - a[4] * a[6];
    double det = a[0] * c00 + a[1] * c01 + a[2] * c02;
    if (!(std::fabs(det) > 0.0) || !std::isfinite(det)) return false;
    double id = 1.0 / det;
    out[0] = c00 * id; out[1] = (a[2] * a[7] - a[1] * a[8]) * id; out[2] = (a[1] * a[5] - a[2] * a[4]) * id;
    out[3] = c01 * id; out[4] = (a[0] * a[8] - a[2] * a[6]) * id; out[5] = (a[2] * a[3] - a[0] * a[5]) * id;
    out[6] = c02 * id; out[7] = (a[1] * a[6] - a[0] * a[7]) * id; out[8] = (a[0] * a[4] - a[1] * a[3]) * id;
    return true;
}

// largest singular value of a 3x3 matrix = sqrt(largest eigenvalue of M^T M)
static double sigma_max3(const double m[9]) {
    double s[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) s[i][j] = m[0 * 3 + i] * m[0 * 3 + j] + m[1 * 3 + i] * m[1 * 3 + j] + m[2 * 3 + i] * m[2 * 3 + j];
    double p1 = s[0][1] * s[0][1] + s[0][2] * s[0][2] + s[1][2] * s[1][2];
    double q = (s[0][0] + s[1][1] + s[2][2]) / 3.0;
    double lam;
    if (p1 <= 1e-300) {
        lam = std::max(s[0][0], std::max(s[1][1], s[2][2]));
    } else {
        double p2 = (s[0][0] - q) * (s[0][0] - q) + (s[1][1] - q) * (s[1][1] - q) + (s[2][2] - q) * (s[2][2] - q) + 2.0 * p1;
        double p = std::sqrt(p2 / 6.0);
        double b[3][3];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) b[i][j] = (s[i][j] - (i == j ? q : 0.0)) / p;
        double detb = b[0][0] * (b[1][1] * b[2][2] - b[1][2] * b[2][1]) - b[0][1] * (b[1][0] * b[2][2] - b[1][2] * b[2][0]) +
                      b[0][2] * (b[1][0] * b[2][1] - b[1][1] * b[2][0]);
        double r = std::min(1.0, std::max(-1.0, detb / 2.0));
        lam = q + 2.0 * p * std::cos(std::acos(r) / 3.0);
    }
    return std::sqrt(std::max(lam, 0.0));
}

// Object->world map of a shape whose exact test moves the ray by inv (3x4): centre and
// the largest stretch of a unit vector.
// `aniso` = largest / smallest stretch (1 for a similarity transform).
static bool world_extent(const float *inv16, double centre[3], double &stretch, double &aniso) {
    double a[9] = {inv16[0], inv16[1], inv16[2], inv16[4], inv16[5], inv16[6], inv16[8], inv16[9], inv16[10]};
    double t[3] = {inv16[3], inv16[7], inv16[11]};
    double f[9];
    if (!inv3(a, f)) return false;
    for (int i = 0; i < 3; ++i) centre[i] = -(f[i * 3 + 0] * t[0] + f[i * 3 + 1] * t[1] + f[i * 3 + 2] * t[2]);
    stretch = sigma_max3(f);
    double inv_min = sigma_max3(a); // 1 / smallest stretch of the forward map
    aniso = stretch * inv_min;
    return std::isfinite(stretch) && std::isfinite(aniso) && std::isfinite(centre[0]) && std::isfinite(centre[1]) &&
           std::isfinite(centre[2]);
}

static float round_up_f32(double x) {
    float f = (float)x;
    if ((double)f < x) f = std::nextafter(f, INFINITY);
    return f;
}

// Radius of a triangle's (and a cube's) bound over its circumradius about the centroid: a
// correctness margin ("filter margins"), a constant of the library.
constexpr double TRI_BOUND_SCALE = 1.5;

static float round_down_f32(double x) {
    float f = (float)x;
    if ((double)f > x) f = std::nextafter(f, -INFINITY);
    return f;
}

// One bound slot: f32 centre, W = |C|^2 - R^2 rounded down, and |C| + R (for cmax).
// R = r * scale + absolute slack for the reference's own rounding of the object-space ray
// + the displacement of the centre when it is rounded to f32 ("filter margins").
// A bound too large or too far for f32 squares gets W = -FLT_MAX: every ray passes it.
struct Bound {
    float cx, cy, cz, w;
    double reach;
    double rf = 0.0; // the radius W was built from
};
static Bound bound_from(const double c[3], double rf);
static Bound make_bound(const double c[3], double r, double scale) {
    double cn = std::sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
    return bound_from(c, r * scale + 64.0 * 1.1920929e-07 * (cn + r));
}
// Sphere (c, rf) as a table entry: the radius grows by the displacement of the f32 centre.
static Bound bound_from(const double c[3], double rf) {
    Bound b;
    b.cx = (float)c[0]; b.cy = (float)c[1]; b.cz = (float)c[2];
    double dx = c[0] - b.cx, dy = c[1] - b.cy, dz = c[2] - b.cz;
    rf += std::sqrt(dx * dx + dy * dy + dz * dz);
    double c2 = (double)b.cx * b.cx + (double)b.cy * b.cy + (double)b.cz * b.cz;
    b.reach = std::sqrt(c2) + rf;
    b.rf = rf;
    if (!(b.reach < 1e15)) {
        b.cx = b.cy = b.cz = 0.f;
        b.w = -3.402823466e+38f;
        b.reach = 0.0;
    } else {
        b.w = round_down_f32(c2 - rf * rf * (1.0 + 1e-6));
    }
    return b;
}

// Sphere around the bounds of one group of GROUP_SLOTS slots (pads skipped): centre G = centre
// of the members' box, moved a few times towards the member that reaches furthest.  A group with a member every ray must pass is itself passed by every
// ray; a group of pads by none.
// Margins: a member (C, R) at distance a from G passes a ray when dist^2(C, line) <= R^2 + mu;
// then dist(G, line) <= a + sqrt(R^2 + mu), and (a + sqrt(R^2 + mu))^2 <= (a + R)^2 + (1 + a / R) mu.
// The group radius therefore counts every member as at least a / 16 wide, and the sweep gives
// group spheres 17 times the margin mu of a bound (sweep.cu: GROUP_MARGIN).  Quads are built from
// their 4 group spheres in the same way with a / 2: 3 times the margin of a group (QUAD_MARGIN), and
// supers from their 4 quad spheres likewise: 3 times the margin of a quad (SUPER_MARGIN).
static Bound make_group(const Bound *slots, const int *shape, int n, double floor_div) {
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    int m = 0;
    for (int j = 0; j < n; ++j) {
        if (shape[j] < 0) continue;
        if (slots[j].w == -3.402823466e+38f) return Bound{0.f, 0.f, 0.f, -3.402823466e+38f, 0.0, 0.0};
        const double p[3] = {slots[j].cx, slots[j].cy, slots[j].cz};
        for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p[a] - slots[j].rf); hi[a] = std::max(hi[a], p[a] + slots[j].rf); }
        ++m;
    }
    if (m == 0) return Bound{0.f, 0.f, 0.f, 3.402823466e+38f, 0.0, 0.0};
    double c[3] = {0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1]), 0.5 * (lo[2] + hi[2])};
    for (int it = 0; it < 8 && m > 1; ++it) { // shrink: step towards the furthest-reaching member by half its lead
        double far1 = -1.0, far2 = -1.0, dir[3] = {0.0, 0.0, 0.0}, dist1 = 0.0;
        for (int j = 0; j < n; ++j) {
            if (shape[j] < 0) continue;
            double dx = slots[j].cx - c[0], dy = slots[j].cy - c[1], dz = slots[j].cz - c[2];
            double a = std::sqrt(dx * dx + dy * dy + dz * dz), reach = a + slots[j].rf;
            if (reach > far1) { far2 = far1; far1 = reach; dir[0] = dx; dir[1] = dy; dir[2] = dz; dist1 = a; }
            else if (reach > far2) far2 = reach;
        }
        if (!(dist1 > 1e-12) || far2 < 0.0) break;
        double step = std::min(0.5 * (far1 - far2), dist1) / dist1;
        c[0] += dir[0] * step; c[1] += dir[1] * step; c[2] += dir[2] * step;
    }
    double rg = 0.0;
    for (int j = 0; j < n; ++j) {
        if (shape[j] < 0) continue;
        double dx = slots[j].cx - c[0], dy = slots[j].cy - c[1], dz = slots[j].cz - c[2];
        double a = std::sqrt(dx * dx + dy * dy + dz * dz);
        rg = std::max(rg, a + std::max(slots[j].rf, a / floor_div));
    }
    return bound_from(c, rg * (1.0 + 1e-9));
}

// Margin factor of a segment ("filter margins"): 1.5 x (32 for the filter's own
// rounding + 25 for the reference's object-space ray + 15 aniso^2 for its discriminant).
static float seg_margin(double aniso) { return round_up_f32(1.5 * (57.0 + 15.0 * aniso * aniso)); }
// Its share for the reference's exact test alone (origin tables, sweep.cu: ray_coef_origin).
static float seg_margin_exact(double aniso) { return round_up_f32(1.5 * (25.0 + 15.0 * aniso * aniso)); }

// Below origin_min_slots a sweep is launch-bound and one general launch per level is cheaper.
static bool origin_tables_wanted(int n_slots) { return tuning().origin_tables != 0 && n_slots >= tuning().origin_min_slots; }

void build_origin_table(const HostScene &hs, const float o[3], std::vector<float4> &out, OriginTable *tab, bool toward) {
    struct S { float x, y, z, w; };
    auto one = [&](const HostSlot &h, double &rmax, bool &clip) {
        if (h.kind == 1) return S{0.f, 0.f, 0.f, 3.402823466e+38f};
        if (h.kind == 0) {
            double x = (double)h.cx - o[0], y = (double)h.cy - o[1], z = (double)h.cz - o[2];
            float fx = (float)x, fy = (float)y, fz = (float)z;
            double dx = x - fx, dy = y - fy, dz = z - fz;
            double rf = h.rf + std::sqrt(dx * dx + dy * dy + dz * dz);
            double c2 = (double)fx * fx + (double)fy * fy + (double)fz * fz;
            if (std::sqrt(c2) + rf < 1e15) {
                rmax = std::max(rmax, rf);
                return S{fx, fy, fz, round_down_f32(c2 - rf * rf * (1.0 + 1e-6))};
            }
        }
        clip = false; // a bound every ray must pass: no far clip in this table
        return S{0.f, 0.f, 0.f, -3.402823466e+38f};
    };
    // slots, then groups, quads, supers (each paired); rmax (the far clip of slot tests) is the slots' alone
    const size_t ns = hs.slots.size(), ng = hs.groups.size(), nq = hs.quads.size(), nsu = hs.supers.size();
    out.resize(ns + ng + nq + nsu);
    bool clip = toward;
    auto fill = [&](const std::vector<HostSlot> &v, float4 *dst) { // returns the largest radius
        double rm[MAX_HOST_CHUNKS] = {0.0};
        bool cl[MAX_HOST_CHUNKS];
        for (bool &c : cl) c = true;
        parallel_chunks(v.size() / 2, 1024, [&](int chunk, size_t begin, size_t end) {
            double r = 0.0;
            bool c = true;
            for (size_t q = begin; q < end; ++q) {
                const size_t j = 2 * q;
                S a = one(v[j], r, c), b = one(v[j + 1], r, c);
                dst[j] = make_float4(a.x, b.x, a.y, b.y);
                dst[j + 1] = make_float4(a.z, b.z, a.w, b.w);
            }
            rm[chunk] = r;
            cl[chunk] = c;
        });
        double r = 0.0;
        for (int c = 0; c < MAX_HOST_CHUNKS; ++c) { r = std::max(r, rm[c]); clip = clip && cl[c]; }
        return r;
    };
    const double rmax_slots = fill(hs.slots, out.data());
    fill(hs.groups, out.data() + ns);
    fill(hs.quads, out.data() + ns + ng);
    fill(hs.supers, out.data() + ns + ng + nq);
    double on = std::sqrt((double)o[0] * o[0] + (double)o[1] * o[1] + (double)o[2] * o[2]);
    double r = (on + (double)hs.dev.cmax) * (1.0 + 1e-6);
    tab->reach = std::isfinite(r) && r < 3e38 ? round_up_f32(r) : 3.0e38f;
    tab->ox = o[0]; tab->oy = o[1]; tab->oz = o[2];
    tab->rmax = (clip && rmax_slots > 0.0 && rmax_slots < 1e15) ? round_up_f32(rmax_slots * (1.0 + 1e-6)) : 0.f;
    tab->shadow = toward ? 1 : 0;
}

int prepare_camera_table(rt_scene *scene, const float origin[3]) {
    if (!scene->cam_tab) return RT_OK; // origin tables are off for this scene
    if (scene->cam.bound && std::memcmp(scene->cam_origin, origin, 3 * sizeof(float)) == 0) return RT_OK;
    const HostScene &hs = *scene->host;
    OriginTable t;
    cudaError_t e;
    { // the devices of a multi-GPU scene share the host copy: the first one to ask builds it
        std::lock_guard<std::mutex> lock(hs.cam_mu);
        if (!hs.cam_valid || std::memcmp(hs.cam_origin, origin, 3 * sizeof(float)) != 0) {
            build_origin_table(hs, origin, hs.cam_host, &hs.cam_tab, false);
            std::memcpy(hs.cam_origin, origin, 3 * sizeof(float));
            hs.cam_valid = true;
        }
        t = hs.cam_tab;
        e = cudaMemcpyAsync(scene->cam_tab, hs.cam_host.data(), hs.cam_host.size() * sizeof(float4), cudaMemcpyHostToDevice, scene->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(scene->stream); // cam_host may be rebuilt once the lock is gone
    }
    if (e != cudaSuccess) return cuda_fail(e, "camera table upload");
    t.bound = scene->cam_tab;
    t.group = scene->cam_tab + hs.slots.size();
    t.quad = t.group + hs.groups.size();
    t.super = t.quad + hs.quads.size();
    scene->cam = t;
    std::memcpy(scene->cam_origin, origin, 3 * sizeof(float));
    return RT_OK;
}

// All device tables of a scene live in ONE allocation, filled by one batch of copies: a scene of
// 10k shapes has 13 tables + one per light, and a cudaMalloc + blocking cudaMemcpy each would
// cost more than the copies themselves.
struct Arena {
    struct Item { const void *src; size_t bytes, off; const void **dev; };
    std::vector<Item> items;
    size_t total = 0;
    void add_raw(const void *src, size_t bytes, const void **dev) {
        *dev = nullptr;
        if (bytes == 0) return;
        items.push_back({src, bytes, total, dev});
        total += (bytes + 255) & ~(size_t)255;
    }
    template <typename T> void add(const std::vector<T> &host, const T **dev) {
        add_raw(host.data(), host.size() * sizeof(T), (const void **)dev);
    }
    // src == nullptr: space only.  The block belongs to the scene's workspace, which outlives the scene
    // in the per-device cache: create -> render -> destroy in a loop allocates nothing.
    int commit(rt_scene *s, cudaStream_t st) {
        if (total == 0) return RT_OK;
        Workspace &ws = s->ws;
        if (ws.tables_cap < total) {
            cudaFree(ws.tables);
            ws.tables = nullptr; ws.tables_cap = 0;
            const size_t cap = total + total / 4;
            cudaError_t e = cudaMalloc(&ws.tables, cap);
            if (e != cudaSuccess) {
                cudaGetLastError();
                ws.tables = nullptr;
                return fail(e == cudaErrorMemoryAllocation ? RT_ERR_NOMEM : RT_ERR_CUDA, "cudaMalloc(scene tables, %zu bytes): %s", cap,
                            cudaGetErrorString(e));
            }
            ws.tables_cap = cap;
        }
        void *d = ws.tables;
        cudaError_t e;
        const size_t packed = blob ? packed_bytes() : 0;
        if (blob && packed > 0 && blob->size() == packed) { // the items' bytes lie in `blob` at the same offsets: one copy
            for (const Item &it : items) *it.dev = (const char *)d + it.off;
            e = cudaMemcpyAsync(d, blob->data(), packed, cudaMemcpyHostToDevice, st);
            if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy(scene tables)");
        } else {
            for (const Item &it : items) {
                *it.dev = (const char *)d + it.off;
                if (!it.src) continue;
                e = cudaMemcpyAsync((char *)d + it.off, it.src, it.bytes, cudaMemcpyHostToDevice, st);
                if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy(scene table)");
            }
        }
        if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return cuda_fail(e, "scene table upload");
        return RT_OK;
    }
    // bytes up to the end of the last item that has a source (space-only items come last)
    size_t packed_bytes() const {
        size_t end = 0;
        for (const Item &it : items)
            if (it.src) end = std::max(end, it.off + it.bytes);
        return end;
    }
    // the items' bytes at their offsets, for HostScene::blob
    void pack(std::vector<char> &out) const {
        out.assign(packed_bytes(), 0);
        for (const Item &it : items)
            if (it.src) std::memcpy(out.data() + it.off, it.src, it.bytes);
    }
    const std::vector<char> *blob = nullptr;
};

static int convert_colorfn(const rt_colorfn &in, DevColorFn &out) {
    if (in.kind != RT_COLOR_CONST && in.kind != RT_COLOR_CHECKERBOARD)
        return fail(RT_ERR_UNSUPPORTED, "texture function kind %d is not a device texture (only CONST, CHECKERBOARD)", in.kind);
    out.kind = in.kind;
    out.r = in.rgb[0]; out.g = in.rgb[1]; out.b = in.rgb[2];
    return RT_OK;
}

int convert_materials(const rt_material *in, int n, std::vector<DevMaterial> &out) {
    out.resize(n);
    for (int i = 0; i < n; ++i) {
        int rc;
        if ((rc = convert_colorfn(in[i].ambient, out[i].ambient)) != RT_OK) return rc;
        if ((rc = convert_colorfn(in[i].diffuse, out[i].diffuse)) != RT_OK) return rc;
        if ((rc = convert_colorfn(in[i].specular, out[i].specular)) != RT_OK) return rc;
        if (in[i].kind != RT_MATERIAL_PHONG && in[i].kind != RT_MATERIAL_TEXTURE_PHONG)
            return fail(RT_ERR_INVALID, "material %d: unknown kind %d", i, in[i].kind);
        out[i].power = in[i].power;
        out[i].reflectivity = in[i].reflectivity;
        out[i].refraction_index = in[i].refraction_index;
        out[i].kind = in[i].kind;
    }
    return RT_OK;
}

// Triangle::new's normal (reference src/scene/triangle.rs:25-29), host f32, no FMA: this
// TU is built with -ffp-contract=off (Makefile).
static void tri_normal(const float *a, const float *b, const float *c, float *n) {
    float e0[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]};
    float e1[3] = {c[0] - b[0], c[1] - b[1], c[2] - b[2]};
    float x = e0[1] * e1[2] - e0[2] * e1[1];
    float y = e0[2] * e1[0] - e0[0] * e1[2];
    float z = e0[0] * e1[1] - e0[1] * e1[0];
    float len = std::sqrt(x * x + y * y + z * z);
    n[0] = x / len; n[1] = y / len; n[2] = z / len;
}

static void pack_tri(const float *v0, const float *v1, const float *v2, const float *n, float4 *out) {
    out[0] = make_float4(v0[0], v0[1], v0[2], v1[0]);
    out[1] = make_float4(v1[1], v1[2], v2[0], v2[1]);
    out[2] = make_float4(v2[2], n[0], n[1], n[2]);
}

// Cube::new (reference src/scene/cube.rs:21-77): the 12 triangles in add order
static void cube_tris(std::vector<float4> &out) {
    const float V[8][3] = {{0.5f, 0.5f, -0.5f}, {0.5f, -0.5f, -0.5f}, {-0.5f, -0.5f, -0.5f}, {-0.5f, 0.5f, -0.5f},
                           {0.5f, 0.5f, 0.5f},  {-0.5f, 0.5f, 0.5f},  {-0.5f, -0.5f, 0.5f},  {0.5f, -0.5f, 0.5f}};
    const int T[12][3] = {{1, 2, 3}, {0, 1, 3}, {7, 5, 4}, {5, 7, 6}, {0, 4, 7}, {7, 1, 0},
                          {5, 3, 6}, {6, 3, 2}, {5, 4, 0}, {0, 3, 5}, {1, 7, 6}, {6, 2, 1}};
    out.resize(36);
    for (int k = 0; k < 12; ++k) {
        float n[3];
        tri_normal(V[T[k][0]], V[T[k][1]], V[T[k][2]], n);
        pack_tri(V[T[k][0]], V[T[k][1]], V[T[k][2]], n, &out[3 * k]);
    }
}

// ---- scene build (host only) -----------------------------------------------------------------

static void add_tables(const HostScene &hs, Arena &arena, DevScene &d, const DevMaterial **dm, std::vector<OriginTable> &light_tabs,
                       const float4 **cam_tab);

int build_host_scene(const rt_scene_desc *desc, HostScene &hs) {
    const bool trace = tuning().trace != 0; // phase times of the host-side build
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto t_last = now();
    auto lap = [&](const char *what) {
        if (!trace) return;
        auto t = now();
        std::fprintf(stderr, "[rt_b200] scene build: %-22s %.3f ms\n", what, std::chrono::duration<double, std::milli>(t - t_last).count());
        t_last = t;
    };
    if (desc->n_shapes < 0 || desc->n_materials < 0 || desc->n_lights < 0 || (desc->n_shapes && !desc->shapes) ||
        (desc->n_materials && !desc->materials) || (desc->n_lights && !desc->lights))
        return fail(RT_ERR_INVALID, "malformed scene description");
    if (desc->n_shapes >= (1 << 24)) return fail(RT_ERR_UNSUPPORTED, "more than 2^24 shapes");

    const int n = desc->n_shapes;
    std::vector<int4> &meta = hs.meta;
    std::vector<float4> &geom = hs.geom;
    meta.resize(n);
    geom.resize(3 * (size_t)n);
    std::vector<int> &plane_shape = hs.plane_shape;
    std::vector<float4> &plane_data = hs.plane_data;
    // error class of a bound = segment of the table: 0 similarity-transformed spheres, 1 triangles,
    // 2 other spheres, 3 cubes (one kind of shape per segment: the sweeps count their filter
    // evaluations per segment, rt_stats.bound_evals_*)
    struct B { Bound b; int shape; int cls; double aniso; };
    static const int cls_kind[MAX_SEGS] = {RT_SHAPE_SPHERE, RT_SHAPE_TRIANGLE, RT_SHAPE_SPHERE, RT_SHAPE_CUBE};
    std::vector<B> bounds;
    int *counts = hs.counts, planes_identity = 0, first_nan = -1;

    for (int i = 0; i < n; ++i) {
        const rt_shape &s = desc->shapes[i];
        if (s.material < 0 || s.material >= desc->n_materials)
            return fail(RT_ERR_INVALID, "shape %d: material %d out of range", i, s.material);
        if (s.kind < 0 || s.kind > 3) return fail(RT_ERR_INVALID, "shape %d: unknown kind %d", i, s.kind);
        counts[s.kind]++;
        meta[i] = make_int4(s.kind, s.material, -1, 0);
        const float *iv = s.inv_transform;
        float4 r0 = make_float4(iv[0], iv[1], iv[2], iv[3]), r1 = make_float4(iv[4], iv[5], iv[6], iv[7]),
               r2 = make_float4(iv[8], iv[9], iv[10], iv[11]);
        double c[3], stretch, aniso;
        switch (s.kind) {
        case RT_SHAPE_SPHERE:
        case RT_SHAPE_CUBE:
            geom[3 * i] = r0; geom[3 * i + 1] = r1; geom[3 * i + 2] = r2;
            if (!world_extent(iv, c, stretch, aniso)) return fail(RT_ERR_INVALID, "shape %d: singular inverse transform", i);
            if (s.kind == RT_SHAPE_SPHERE)
                bounds.push_back({make_bound(c, stretch, 1.0 + 1e-4), i, aniso <= 1.001 ? 0 : 2, aniso});
            else bounds.push_back({make_bound(c, stretch * 0.8660254037844387, TRI_BOUND_SCALE), i, 3, aniso});
            if (first_nan < 0) first_nan = i;
            break;
        case RT_SHAPE_TRIANGLE: {
            const float *p = s.p;
            pack_tri(p, p + 3, p + 6, p + 9, &geom[3 * i]);
            for (int k = 0; k < 3; ++k) c[k] = ((double)p[k] + p[3 + k] + p[6 + k]) / 3.0;
            double r = 0.0;
            for (int v = 0; v < 3; ++v) {
                double dx = p[3 * v] - c[0], dy = p[3 * v + 1] - c[1], dz = p[3 * v + 2] - c[2];
                r = std::max(r, std::sqrt(dx * dx + dy * dy + dz * dz));
            }
            if (!std::isfinite(r)) return fail(RT_ERR_INVALID, "shape %d: non-finite triangle", i);
            bounds.push_back({make_bound(c, r, TRI_BOUND_SCALE), i, 1, 1.0});
            if (first_nan < 0) first_nan = i;
            break;
        }
        case RT_SHAPE_PLANE: {
            geom[3 * i] = r0; geom[3 * i + 1] = r1; geom[3 * i + 2] = r2;
            meta[i].z = (int)plane_shape.size();
            plane_shape.push_back(i);
            const float *p = s.p, *f = s.transform;
            plane_data.push_back(make_float4(p[0], p[1], p[2], 0.f));
            plane_data.push_back(make_float4(p[3], p[4], p[5], 0.f));
            plane_data.push_back(make_float4(p[6], p[7], p[8], 0.f));
            plane_data.push_back(make_float4(p[9], p[10], p[11], 0.f));
            plane_data.push_back(make_float4(f[0], f[1], f[2], 0.f));
            plane_data.push_back(make_float4(f[4], f[5], f[6], 0.f));
            plane_data.push_back(make_float4(f[8], f[9], f[10], 0.f));
            static const float I[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
            if (std::memcmp(iv, I, sizeof(I)) == 0) planes_identity++;
            break;
        }
        }
    }
    hs.planes_identity = planes_identity;
    // segments by error class, each padded to SEG_PAD slots; within a class the bounds are put in
    // the leaf order of a median-split tree whose leaves hold exactly GROUP_SLOTS bounds, so that
    // the slots of a group are neighbours in space; slot pairs interleaved for FFMA2
    lap("shape records, bounds");
    std::stable_sort(bounds.begin(), bounds.end(), [](const B &a, const B &b) {
        bool fa = a.b.w != -3.402823466e+38f, fb = b.b.w != -3.402823466e+38f; // beyond-range bounds last in their class
        return a.cls != b.cls ? a.cls < b.cls : (fa && !fb);
    });
    {
        auto coord = [](const B &x, int ax) { return ax == 0 ? x.b.cx : (ax == 1 ? x.b.cy : x.b.cz); };
        std::vector<std::pair<int, int>> work; // [begin, end) ranges still to split
        for (size_t i = 0; i < bounds.size();) {
            size_t j = i;
            while (j < bounds.size() && bounds[j].cls == bounds[i].cls && bounds[j].b.w != -3.402823466e+38f) ++j;
            if (j > i) work.push_back({(int)i, (int)j});
            while (j < bounds.size() && bounds[j].cls == bounds[i].cls) ++j;
            i = j;
        }
        // one median split of [lo, hi) at a whole number of groups; false when it is a single leaf
        auto split = [&](int lo, int hi, int &mid) {
            const int n = hi - lo;
            if (n <= GROUP_SLOTS) return false;
            float mn[3] = {3.4e38f, 3.4e38f, 3.4e38f}, mx[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
            for (int i = lo; i < hi; ++i)
                for (int ax = 0; ax < 3; ++ax) { mn[ax] = std::min(mn[ax], coord(bounds[i], ax)); mx[ax] = std::max(mx[ax], coord(bounds[i], ax)); }
            int ax = 0;
            if (mx[1] - mn[1] > mx[ax] - mn[ax]) ax = 1;
            if (mx[2] - mn[2] > mx[ax] - mn[ax]) ax = 2;
            const int groups = (n + GROUP_SLOTS - 1) / GROUP_SLOTS, left = (groups / 2) * GROUP_SLOTS;
            std::nth_element(bounds.begin() + lo, bounds.begin() + lo + left, bounds.begin() + hi,
                             [&](const B &a, const B &b) { return coord(a, ax) < coord(b, ax); });
            mid = lo + left;
            return true;
        };
        // the top of the tree here, until there is a subtree per host thread; the subtrees (disjoint ranges) in parallel
        const size_t want = (size_t)host_threads();
        std::vector<std::pair<int, int>> roots;
        size_t head = 0;
        while (head < work.size() && (work.size() - head) + roots.size() < want) { // breadth first: subtrees of like size
            auto [lo, hi] = work[head++];
            int mid;
            if (hi - lo >= 4096 && split(lo, hi, mid)) {
                work.push_back({lo, mid});
                work.push_back({mid, hi});
            } else {
                roots.push_back({lo, hi});
            }
        }
        roots.insert(roots.end(), work.begin() + head, work.end());
        std::sort(roots.begin(), roots.end(), [](const std::pair<int, int> &a, const std::pair<int, int> &b) { return a.second - a.first > b.second - b.first; });
        parallel_chunks(roots.size(), 1, [&](int, size_t begin, size_t end) {
            std::vector<std::pair<int, int>> stack(roots.begin() + begin, roots.begin() + end);
            while (!stack.empty()) {
                auto [lo, hi] = stack.back();
                stack.pop_back();
                int mid;
                if (!split(lo, hi, mid)) continue;
                stack.push_back({lo, mid});
                stack.push_back({mid, hi});
            }
        });
    }
    lap("slot order");
    const int nb = (int)bounds.size();
    SweepSegs segs;
    std::memset(&segs, 0, sizeof(segs));
    std::vector<Bound> slots;
    std::vector<int> &bound_shape = hs.bound_shape;
    double cmax = 0.0;
    const Bound pad = {0.f, 0.f, 0.f, 3.402823466e+38f, 0.0}; // passes no ray
    for (int j = 0; j < nb;) {
        int cls = bounds[j].cls;
        double aniso = 1.0;
        int g = segs.n++;
        segs.begin[g] = (int)slots.size();
        hs.seg_kind[g] = cls_kind[cls];
        for (; j < nb && bounds[j].cls == cls; ++j) {
            slots.push_back(bounds[j].b);
            bound_shape.push_back(bounds[j].shape);
            aniso = std::max(aniso, bounds[j].aniso);
            cmax = std::max(cmax, bounds[j].b.reach);
            hs.seg_real[g]++;
        }
        while (slots.size() % SEG_PAD) { slots.push_back(pad); bound_shape.push_back(-1); }
        segs.end[g] = (int)slots.size();
        segs.k[g] = seg_margin(aniso);
        segs.kx[g] = seg_margin_exact(aniso);
    }
    for (int g = segs.n; g < MAX_SEGS; ++g) segs.begin[g] = segs.end[g] = INT_MAX; // unused: no slot lies beyond them (sweep.cu: seg_of)
    const int n_slots = (int)slots.size();
    const int n_groups = n_slots / GROUP_SLOTS; // even: SEG_PAD is two groups
    std::vector<Bound> groups(n_groups);
    {
        double reach[MAX_HOST_CHUNKS] = {0.0};
        parallel_chunks((size_t)n_groups, 64, [&](int chunk, size_t begin, size_t end) {
            double r = 0.0;
            for (size_t g = begin; g < end; ++g) {
                groups[g] = make_group(&slots[g * GROUP_SLOTS], &bound_shape[g * GROUP_SLOTS], GROUP_SLOTS, 16.0);
                r = std::max(r, groups[g].reach);
            }
            reach[chunk] = r;
        });
        for (double r : reach) cmax = std::max(cmax, r); // a group sphere may reach beyond its members
    }
    const int per_quad = QUAD_SLOTS / GROUP_SLOTS, n_quads = n_groups / per_quad; // even: SEG_PAD is two quads
    std::vector<Bound> quads(n_quads);
    {
        std::vector<int> live(n_groups);
        for (int g = 0; g < n_groups; ++g) live[g] = groups[g].w == 3.402823466e+38f ? -1 : 0; // a group of pads
        for (int q = 0; q < n_quads; ++q) {
            quads[q] = make_group(&groups[(size_t)q * per_quad], &live[(size_t)q * per_quad], per_quad, 2.0);
            cmax = std::max(cmax, quads[q].reach);
        }
    }
    const int per_super = SUPER_SLOTS / QUAD_SLOTS, n_supers = n_quads / per_super; // even: SEG_PAD is two supers
    std::vector<Bound> supers(n_supers);
    {
        std::vector<int> live(n_quads);
        for (int q = 0; q < n_quads; ++q) live[q] = quads[q].w == 3.402823466e+38f ? -1 : 0; // a quad of pads
        for (int q = 0; q < n_supers; ++q) {
            supers[q] = make_group(&quads[(size_t)q * per_super], &live[(size_t)q * per_super], per_super, 2.0);
            cmax = std::max(cmax, supers[q].reach);
        }
    }
    lap("groups, quads, supers");
    auto paired = [](const std::vector<Bound> &v) {
        std::vector<float4> out(v.size());
        for (size_t j = 0; j + 1 < v.size(); j += 2) {
            const Bound &a = v[j], &b = v[j + 1];
            out[j] = make_float4(a.cx, b.cx, a.cy, b.cy);
            out[j + 1] = make_float4(a.cz, b.cz, a.w, b.w);
        }
        return out;
    };
    hs.bound = paired(slots); hs.group = paired(groups); hs.quad = paired(quads); hs.super = paired(supers);
    hs.slot_meta.resize(n_slots);
    hs.slot_geom.assign(3 * (size_t)n_slots, make_float4(0.f, 0.f, 0.f, 0.f));
    for (int j = 0; j < n_slots; ++j) {
        const int sh = bound_shape[j];
        hs.slot_meta[j] = make_int2(sh >= 0 ? meta[sh].x : -1, sh);
        if (sh >= 0) for (int k = 0; k < 3; ++k) hs.slot_geom[3 * (size_t)j + k] = geom[3 * (size_t)sh + k];
    }

    int rc = convert_materials(desc->materials, desc->n_materials, hs.materials);
    if (rc != RT_OK) return rc;
    hs.n_materials = desc->n_materials;
    std::vector<DevLight> &lights = hs.lights;
    lights.resize(desc->n_lights);
    int n_point = 0;
    for (int l = 0; l < desc->n_lights; ++l) {
        const rt_light &L = desc->lights[l];
        if (L.kind != RT_LIGHT_POINT && L.kind != RT_LIGHT_AMBIENT) return fail(RT_ERR_INVALID, "light %d: unknown kind", l);
        lights[l].px = L.pos[0]; lights[l].py = L.pos[1]; lights[l].pz = L.pos[2];
        lights[l].kind = L.kind;
        lights[l].cr = L.color[0]; lights[l].cg = L.color[1]; lights[l].cb = L.color[2];
        lights[l].point_index = (L.kind == RT_LIGHT_POINT) ? n_point++ : -1;
    }
    cube_tris(hs.cube_tris);
    lap("slot tables, materials");

    DevScene &d = hs.dev;
    std::memset(&d, 0, sizeof(d));
    d.n_shapes = n; d.n_bounded = nb; d.n_planes = (int)plane_shape.size();
    d.n_slots = n_slots; d.segs = segs; d.cmax = round_up_f32(cmax * (1.0 + 1e-6));
    { // largest sphere of each level of the hierarchy, for the sweeps' distance clip (origin tables move a centre by its f32 rounding only)
        {
            double r = 0.0;
            bool all_pass = false;
            for (const Bound &b : slots) {
                if (b.w == -3.402823466e+38f) all_pass = true;
                else if (b.w != 3.402823466e+38f) r = std::max(r, b.rf);
            }
            r = r * (1.0 + 1e-5) + 1e-5 * cmax;
            d.slot_rmax = (all_pass || !(r < 1e15)) ? 3.0e38f : round_up_f32(r);
        }
        const std::vector<Bound> *levels[3] = {&groups, &quads, &supers};
        for (int l = 0; l < 3; ++l) {
            double r = 0.0;
            bool all_pass = false;
            for (const Bound &b : *levels[l]) {
                if (b.w == -3.402823466e+38f) all_pass = true;
                else if (b.w != 3.402823466e+38f) r = std::max(r, b.rf);
            }
            r = r * (1.0 + 1e-5) + 1e-5 * cmax;
            d.level_rmax[l] = (all_pass || !(r < 1e15)) ? 3.0e38f : round_up_f32(r);
        }
    }
    for (int g = 0; g < segs.n; ++g) { // runs of segments with the same margins
        SweepSegs &r = d.runs;
        if (r.n > 0 && r.k[r.n - 1] == segs.k[g] && r.kx[r.n - 1] == segs.kx[g] && r.end[r.n - 1] == segs.begin[g]) {
            r.end[r.n - 1] = segs.end[g];
        } else {
            r.begin[r.n] = segs.begin[g]; r.end[r.n] = segs.end[g]; r.k[r.n] = segs.k[g]; r.kx[r.n] = segs.kx[g];
            r.n += 1;
        }
    }
    { // node cells: the box around the bounded shapes, Pools::node_bits bits per axis (hit points outside are clamped)
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        for (const Bound &b : slots) {
            if (b.w == 3.402823466e+38f || b.w == -3.402823466e+38f) continue;
            const double c[3] = {b.cx, b.cy, b.cz};
            for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], c[a] - b.rf); hi[a] = std::max(hi[a], c[a] + b.rf); }
        }
        for (int a = 0; a < 3; ++a) {
            double ext = (hi[a] - lo[a]) * 1.5; // half as large again: the floor under and around the shapes is inside
            bool ok = ext > 0.0 && std::isfinite(ext) && ext < 1e30;
            d.bin_lo[a] = ok ? (float)(lo[a] - ext / 6.0) : 0.f;
            d.bin_scale[a] = ok ? (float)(2.0 / ext) : 0.f; // (x - lo) * scale - 1 in [-1, 1] inside the box
        }
    }
    d.n_lights = desc->n_lights; d.n_point_lights = n_point; d.first_nan_shape = first_nan;
    d.amb_r = desc->ambient[0]; d.amb_g = desc->ambient[1]; d.amb_b = desc->ambient[2];

    // the bound table seen from every point light (the one seen from the camera is built at render time)
    hs.origin = origin_tables_wanted(n_slots);
    if (hs.origin) {
        auto host_slot = [](const Bound &b) {
            return HostSlot{b.cx, b.cy, b.cz, b.rf, b.w == 3.402823466e+38f ? 1 : (b.w == -3.402823466e+38f ? 2 : 0)};
        };
        hs.slots.reserve(slots.size()); hs.groups.reserve(groups.size()); hs.quads.reserve(quads.size());
        for (const Bound &b : slots) hs.slots.push_back(host_slot(b));
        for (const Bound &b : groups) hs.groups.push_back(host_slot(b));
        for (const Bound &b : quads) hs.quads.push_back(host_slot(b));
        for (const Bound &b : supers) hs.supers.push_back(host_slot(b));
        hs.light_tabs.resize(n_point);
        hs.light_host.resize(n_point);
        for (int l = 0; l < desc->n_lights; ++l) {
            if (lights[l].point_index < 0) continue;
            const float o[3] = {lights[l].px, lights[l].py, lights[l].pz};
            build_origin_table(hs, o, hs.light_host[lights[l].point_index], &hs.light_tabs[lights[l].point_index], true);
        }
    }
    lap("light tables");
    { // the same tables once more, packed the way upload_scene lays them out on a device (HostScene::blob)
        Arena arena;
        DevScene dd = hs.dev;
        const DevMaterial *dm = nullptr;
        const float4 *cam = nullptr;
        std::vector<OriginTable> lt = hs.light_tabs;
        add_tables(hs, arena, dd, &dm, lt, &cam);
        try {
            arena.pack(hs.blob);
        } catch (...) {
            hs.blob.clear(); // no memory for the copy: upload_scene sends the tables one by one
        }
    }
    lap("upload blob");
    return RT_OK;
}

// ---- the last built scene, kept ----------------------------------------------------------------
// A host that renders frame after frame hands the library the same description every time (the
// reference builds its Scene once per process, src/main.rs:33-34, and the one-shot calls of this ABI
// rebuild it per frame).  The tables of the last description built are therefore kept, together with
// a copy of that description: an identical one (same counts, same bytes) gets them back for the price
// of a memcmp (~0.2 ms for 10k shapes against ~2.5 ms of building).  rt_trim(-1) drops the entry.
namespace {
struct SceneCacheEntry {
    std::vector<rt_shape> shapes;
    std::vector<rt_material> materials;
    std::vector<rt_light> lights;
    float ambient[3] = {0.f, 0.f, 0.f};
    int origin_tables = 0, origin_min_slots = 0; // the knobs the build depends on
    std::shared_ptr<const HostScene> host;
};
std::mutex g_scene_cache_mu;
SceneCacheEntry g_scene_cache;

bool same_description(const SceneCacheEntry &c, const rt_scene_desc *d) {
    if (!c.host || (int)c.shapes.size() != d->n_shapes || (int)c.materials.size() != d->n_materials ||
        (int)c.lights.size() != d->n_lights)
        return false;
    if (c.origin_tables != tuning().origin_tables || c.origin_min_slots != tuning().origin_min_slots) return false;
    if (std::memcmp(c.ambient, d->ambient, sizeof(c.ambient)) != 0) return false;
    return (d->n_shapes == 0 || std::memcmp(c.shapes.data(), d->shapes, sizeof(rt_shape) * (size_t)d->n_shapes) == 0) &&
           (d->n_materials == 0 || std::memcmp(c.materials.data(), d->materials, sizeof(rt_material) * (size_t)d->n_materials) == 0) &&
           (d->n_lights == 0 || std::memcmp(c.lights.data(), d->lights, sizeof(rt_light) * (size_t)d->n_lights) == 0);
}
} // namespace

void drop_scene_cache() {
    std::lock_guard<std::mutex> lock(g_scene_cache_mu);
    g_scene_cache = SceneCacheEntry();
}

int host_scene_for(const rt_scene_desc *desc, std::shared_ptr<const HostScene> *out) {
    if (desc->n_shapes < 0 || desc->n_materials < 0 || desc->n_lights < 0 || (desc->n_shapes && !desc->shapes) ||
        (desc->n_materials && !desc->materials) || (desc->n_lights && !desc->lights))
        return fail(RT_ERR_INVALID, "malformed scene description");
    const bool use_cache = tuning().scene_cache != 0;
    if (use_cache) {
        std::lock_guard<std::mutex> lock(g_scene_cache_mu);
        if (same_description(g_scene_cache, desc)) {
            *out = g_scene_cache.host;
            return RT_OK;
        }
    }
    std::shared_ptr<HostScene> hs;
    try {
        hs = std::make_shared<HostScene>();
    } catch (...) {
        return fail(RT_ERR_NOMEM, "out of host memory");
    }
    int rc = build_host_scene(desc, *hs);
    if (rc != RT_OK) return rc;
    *out = hs;
    if (use_cache) {
        try {
            SceneCacheEntry e;
            e.shapes.assign(desc->shapes, desc->shapes + desc->n_shapes);
            e.materials.assign(desc->materials, desc->materials + desc->n_materials);
            e.lights.assign(desc->lights, desc->lights + desc->n_lights);
            std::memcpy(e.ambient, desc->ambient, sizeof(e.ambient));
            e.origin_tables = tuning().origin_tables;
            e.origin_min_slots = tuning().origin_min_slots;
            e.host = hs;
            std::lock_guard<std::mutex> lock(g_scene_cache_mu);
            g_scene_cache = std::move(e);
        } catch (...) { // no memory for the copy: nothing is kept
        }
    }
    return RT_OK;
}

// ---- scene upload (one device) ------------------------------------------------------------------

// The tables of a scene in upload order; `light_tabs` (a copy of hs.light_tabs) and *cam_tab receive their device addresses.
static void add_tables(const HostScene &hs, Arena &arena, DevScene &d, const DevMaterial **dm, std::vector<OriginTable> &light_tabs,
                       const float4 **cam_tab) {
    arena.add(hs.bound, &d.bound); arena.add(hs.group, &d.group); arena.add(hs.quad, &d.quad); arena.add(hs.super, &d.super);
    arena.add(hs.bound_shape, &d.bound_shape); arena.add(hs.slot_meta, &d.slot_meta); arena.add(hs.slot_geom, &d.slot_geom);
    arena.add(hs.meta, &d.shape_meta); arena.add(hs.geom, &d.geom);
    arena.add(hs.plane_shape, &d.plane_shape); arena.add(hs.plane_data, &d.plane_data);
    arena.add(hs.cube_tris, &d.cube_tris); arena.add(hs.materials, dm); arena.add(hs.lights, &d.lights);
    if (hs.origin) {
        const int n = hs.dev.n_slots;
        for (size_t l = 0; l < light_tabs.size(); ++l) arena.add(hs.light_host[l], &light_tabs[l].bound);
        arena.add_raw(nullptr, (size_t)(n + n / GROUP_SLOTS + n / QUAD_SLOTS + n / SUPER_SLOTS) * sizeof(float4), (const void **)cam_tab); // space only, last
    }
}

HostScene::~HostScene() {
    if (blob_pinned) {
        cudaHostUnregister((void *)blob.data());
        cudaGetLastError(); // (at process exit the runtime may be gone already)
    }
}

void scene_free(rt_scene *scene) {
    if (!scene) return;
    cudaSetDevice(scene->device);
    workspace_give_cached(scene->device, scene->ws); // the scene's stream goes with it
    delete scene;
}

int upload_scene(const std::shared_ptr<const HostScene> &host, int device, rt_scene **out) {
    *out = nullptr;
    const HostScene &hs = *host;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount (no CPU fallback exists)");
    if (device < 0 || device >= ndev) return fail(RT_ERR_CUDA, "CUDA device %d not present (%d devices)", device, ndev);
    if ((e = cudaSetDevice(device)) != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    rt_scene *s = new (std::nothrow) rt_scene();
    if (!s) return fail(RT_ERR_NOMEM, "out of host memory");
    s->device = device;
    s->host = host;
    s->n_materials = hs.n_materials;
    s->host_materials = hs.materials;
    s->dev = hs.dev;
    DevScene &d = s->dev;
    cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, device);
    const DevMaterial *dm = nullptr;
    Arena arena;
    const float4 *cam_tab = nullptr;
    const int n_groups = d.n_slots / GROUP_SLOTS, n_quads = d.n_slots / QUAD_SLOTS, n_supers = d.n_slots / SUPER_SLOTS;
    s->light_tabs = hs.light_tabs;
    add_tables(hs, arena, d, &dm, s->light_tabs, &cam_tab);
    if (!hs.blob.empty()) { // page-lock the packed tables once; without it the copy is staged like any pageable one
        std::lock_guard<std::mutex> lock(hs.blob_mu);
        if (!hs.blob_pin_tried) {
            hs.blob_pin_tried = true;
            if (cudaHostRegister((void *)hs.blob.data(), hs.blob.size(), cudaHostRegisterPortable) == cudaSuccess) hs.blob_pinned = true;
            else cudaGetLastError();
        }
        arena.blob = &hs.blob;
    }
    s->ws = workspace_take_cached(device);
    if (!s->ws.stream && (e = cudaStreamCreateWithFlags(&s->ws.stream, cudaStreamNonBlocking)) != cudaSuccess) {
        scene_free(s);
        return cuda_fail(e, "cudaStreamCreate");
    }
    s->stream = s->own_stream = s->ws.stream;
    int rc = arena.commit(s, s->stream);
    if (rc != RT_OK) {
        scene_free(s);
        return rc;
    }
    d.materials = dm;
    s->d_materials = const_cast<DevMaterial *>(dm);
    for (OriginTable &t : s->light_tabs) {
        t.group = t.bound + d.n_slots;
        t.quad = t.group + n_groups;
        t.super = t.quad + n_quads;
    }
    s->cam_tab = const_cast<float4 *>(cam_tab);
    *out = s;
    return RT_OK;
}

} // namespace rt

using namespace rt;

extern "C" {

int rt_abi_version(void) { return RT_B200_ABI_VERSION; }
const char *rt_last_error(void) { return g_err; }

int rt_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
    return n;
}

void rt_camera_default(int32_t x_res, int32_t y_res, rt_camera *out) {
    out->origin[0] = 0.f; out->origin[1] = 0.f; out->origin[2] = -8.f;
    out->x_min = -3.f; out->x_max = 3.f; out->y_min = -3.f; out->y_max = 3.f;
    out->x_res = x_res; out->y_res = y_res;
}

int32_t rt_tile_rows_for_rank(int32_t height, int32_t rank, int32_t world, int32_t tile_rows, int32_t *rows,
                              int32_t cap) {
    if (height < 0 || world < 1 || rank < 0 || rank >= world || tile_rows < 1) return fail(RT_ERR_INVALID, "bad tiling");
    int32_t n = 0;
    for (int32_t v = 0; v < height; ++v)
        if ((v / tile_rows) % world == rank) {
            if (rows && n < cap) rows[n] = v;
            ++n;
        }
    return n;
}

int rt_scene_create(const rt_scene_desc *desc, int32_t device, rt_scene **out) {
    if (!desc || !out) return fail(RT_ERR_INVALID, "null argument");
    *out = nullptr;
    const bool trace = tuning().trace != 0;
    auto t0 = std::chrono::steady_clock::now();
    std::shared_ptr<const HostScene> hs;
    int rc = host_scene_for(desc, &hs); // a malformed description is reported even where there is no device
    if (rc != RT_OK) return rc;
    auto t1 = std::chrono::steady_clock::now();
    rc = upload_scene(hs, device, out);
    if (trace)
        std::fprintf(stderr, "[rt_b200] scene create: tables %.3f ms, upload %.3f ms\n", std::chrono::duration<double, std::milli>(t1 - t0).count(),
                     std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count());
    return rc;
}

int rt_scene_set_stream(rt_scene *scene, void *stream) {
    if (!scene) return fail(RT_ERR_INVALID, "null scene");
    scene->stream = stream ? (cudaStream_t)stream : scene->own_stream;
    return RT_OK;
}

int rt_scene_update_materials(rt_scene *scene, const rt_material *materials, int32_t n) {
    if (!scene || (!materials && n)) return fail(RT_ERR_INVALID, "null argument");
    if (n != scene->n_materials) return fail(RT_ERR_INVALID, "expected %d materials, got %d", scene->n_materials, n);
    std::vector<DevMaterial> mats;
    int rc = convert_materials(materials, n, mats);
    if (rc != RT_OK) return rc;
    if (scene->forests > 0) {
        // A ray forest fixes which child rays exist and where they go (render_tree.rs:166-212 builds the
        // tree once; render_ray_tree :214-255 only re-reads colours and weights).  Both depend on
        // reflectivity and refraction_index, so those must not change under a live forest; colours and
        // power may (the reference's GUI only ever calls set_diffuse, gui.rs:221-235).
        const std::vector<DevMaterial> &old = scene->host_materials;
        for (int i = 0; i < n; ++i)
            if (mats[i].reflectivity != old[i].reflectivity || mats[i].refraction_index != old[i].refraction_index)
                return fail(RT_ERR_UNSUPPORTED, "material %d: reflectivity / refraction_index cannot change while a ray forest built from this scene exists", i);
    }
    cudaError_t e = cudaSetDevice(scene->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    if (n) {
        e = cudaMemcpy(scene->d_materials, mats.data(), sizeof(DevMaterial) * n, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy(materials)");
    }
    scene->host_materials = mats;
    return RT_OK;
}

void rt_scene_destroy(rt_scene *scene) {
    if (!scene) return;
    if (scene->forests > 0) { // forests built from it still use its tables and stream: the last one frees it
        scene->destroy_pending = true;
        return;
    }
    scene_free(scene);
}

// ---- debug hooks (include/rt_b200_debug.h) ---------------------------------------------------

namespace {
struct Knob {
    const char *name;
    int Tuning::*field;
    long long lo, hi;
};
const Knob KNOBS[] = {
    {"rays_per_thread", &Tuning::rays_per_thread, 0, 8},   {"want_light", &Tuning::want_light, 1, 4096},
    {"want_heavy", &Tuning::want_heavy, 1, 4096},          {"want_heavy_jobs", &Tuning::want_heavy_jobs, 1, 4096},
    {"tail_min", &Tuning::tail_min, 0, 1 << 20},          {"slices", &Tuning::slices, 0, 1 << 20},
    {"tail_pct", &Tuning::tail_pct, 0, 100},               {"tail_mult", &Tuning::tail_mult, 1, 64},
    {"item_max", &Tuning::item_max, 0, 32},                {"table_cap", &Tuning::table_cap, 0, 1 << 24},
    {"merge_lights", &Tuning::merge_lights, 0, 1},         {"heavy_allow", &Tuning::heavy_allow, 0, 1 << 24},
    {"node_bits", &Tuning::node_bits, 0, NODE_BIN_BITS_MAX},
    {"band_samples", &Tuning::band_samples, 1 << 10, 1 << 28}, {"origin_tables", &Tuning::origin_tables, 0, 1},
    {"origin_min_slots", &Tuning::origin_min_slots, 0, 1 << 24}, {"host_threads", &Tuning::host_threads, 1, 64},
    {"shadow_anyhit", &Tuning::shadow_anyhit, 0, 1},       {"trace", &Tuning::trace, 0, 2},
    {"scene_cache", &Tuning::scene_cache, 0, 1},         {"job_rotate", &Tuning::job_rotate, 0, 1},
    {"small_tree_rays", &Tuning::small_tree_rays, 0, 1 << 30}, {"tree_rays_per_thread", &Tuning::tree_rays_per_thread, 1, 4},
    {"small_shadow_rays", &Tuning::small_shadow_rays, 0, 1 << 30},
    {"level_pipeline", &Tuning::level_pipeline, 0, 1},     {"primary_tiles", &Tuning::primary_tiles, 0, 2},
};
} // namespace

// Test hook: the bound hierarchy build_host_scene derives from a description, as the host sees it -- level 0 bound slots,
// 1 groups, 2 quads, 3 supers, five doubles per sphere (cx, cy, cz, radius incl. slack, kind: 0 sphere, 1 pad, 2 passes every
// ray).  Returns the number of spheres of the level (up to `cap` are written), or a negative rt_status.  No GPU.
int rt_debug_scene_spheres(const rt_scene_desc *desc, int32_t level, double *out, int32_t cap) {
    if (!desc || level < 0 || level > 3 || (cap > 0 && !out)) return fail(RT_ERR_INVALID, "bad argument");
    std::shared_ptr<const HostScene> hs;
    int rc = host_scene_for(desc, &hs);
    if (rc != RT_OK) return rc;
    if (!hs->origin) return fail(RT_ERR_UNSUPPORTED, "the scene keeps no host copy of its hierarchy (origin tables are off)");
    const std::vector<HostSlot> &v = level == 0 ? hs->slots : (level == 1 ? hs->groups : (level == 2 ? hs->quads : hs->supers));
    for (size_t i = 0; i < v.size() && (int32_t)i < cap; ++i) {
        out[5 * i] = v[i].cx; out[5 * i + 1] = v[i].cy; out[5 * i + 2] = v[i].cz; out[5 * i + 3] = v[i].rf; out[5 * i + 4] = v[i].kind;
    }
    return (int)v.size();
}
// The table of point light `light` (by DevLight::point_index order) as the sweeps read it, unpacked: four doubles per sphere
// (cx, cy, cz, W) relative to the light -- the slots, then the groups, quads and supers -- and out_origin = the light's
// position, reach, rmax.  Returns the number of spheres (at most `cap` written) or a negative rt_status.
int rt_debug_scene_light_table(const rt_scene_desc *desc, int32_t light, double *out, int32_t cap, double out_origin[5]) {
    if (!desc || light < 0 || (cap > 0 && !out)) return fail(RT_ERR_INVALID, "bad argument");
    std::shared_ptr<const HostScene> hs;
    int rc = host_scene_for(desc, &hs);
    if (rc != RT_OK) return rc;
    if (!hs->origin || light >= (int)hs->light_host.size()) return fail(RT_ERR_INVALID, "no such light table");
    const std::vector<float4> &t = hs->light_host[light];
    for (size_t j = 0; j + 1 < t.size() && (int32_t)(j + 1) < cap; j += 2) { // pairs: (x0, x1, y0, y1), (z0, z1, W0, W1)
        const float4 a = t[j], b = t[j + 1];
        out[4 * j] = a.x; out[4 * j + 1] = a.z; out[4 * j + 2] = b.x; out[4 * j + 3] = b.z;
        out[4 * j + 4] = a.y; out[4 * j + 5] = a.w; out[4 * j + 6] = b.y; out[4 * j + 7] = b.w;
    }
    if (out_origin) {
        const OriginTable &ot = hs->light_tabs[light];
        out_origin[0] = ot.ox; out_origin[1] = ot.oy; out_origin[2] = ot.oz; out_origin[3] = ot.reach; out_origin[4] = ot.rmax;
    }
    return (int)t.size();
}
// ... and the scalars the sweeps take from it: out = {n_slots, cmax, slot_rmax, level_rmax[0..2]}.
int rt_debug_scene_scalars(const rt_scene_desc *desc, double out[6]) {
    if (!desc || !out) return fail(RT_ERR_INVALID, "bad argument");
    std::shared_ptr<const HostScene> hs;
    int rc = host_scene_for(desc, &hs);
    if (rc != RT_OK) return rc;
    const DevScene &d = hs->dev;
    out[0] = d.n_slots; out[1] = d.cmax; out[2] = d.slot_rmax; out[3] = d.level_rmax[0]; out[4] = d.level_rmax[1]; out[5] = d.level_rmax[2];
    return RT_OK;
}

int rt_debug_set(const char *name, int64_t value) {
    if (!name) return fail(RT_ERR_INVALID, "null argument");
    for (const Knob &k : KNOBS)
        if (std::strcmp(k.name, name) == 0) {
            if (value < k.lo || value > k.hi) return fail(RT_ERR_INVALID, "%s = %lld is not in %lld..%lld", name, (long long)value, k.lo, k.hi);
            tuning().*(k.field) = (int)value;
            return RT_OK;
        }
    return fail(RT_ERR_INVALID, "no tuning knob named %s", name);
}

int64_t rt_debug_get(const char *name) {
    if (name)
        for (const Knob &k : KNOBS)
            if (std::strcmp(k.name, name) == 0) return tuning().*(k.field);
    return INT64_MIN;
}

void rt_debug_reset(void) { tuning() = Tuning(); }

} // extern "C"
