// ORACLE -- TEST INFRASTRUCTURE ONLY (see orc_math.hpp).
// RayForest object of src/render_tree.rs:51-145 kept across calls, so that tests can edit
// materials between evaluations the way src/gui.rs:163-235 does.
#include <atomic>
#include <thread>

#include "orc_render.hpp"

using namespace orc;

namespace {
struct Forest {
    int w = 0, h = 0;
    std::vector<RayTree> trees; // row-major [v][u] (the reference indexes [u][v])
};
} // namespace

extern "C" {

// generate_ray_forest (src/render_tree.rs:147-164)
void *orc_forest_build(const rt_scene_desc *desc, const rt_camera *cam, int depth, int threads) {
    Scene scene;
    if (!scene_from_desc(*desc, scene)) return nullptr;
    Forest *f = new Forest();
    f->w = cam->x_res;
    f->h = cam->y_res;
    f->trees.resize((size_t)f->w * f->h);
    if (threads < 1) threads = 1;
    std::atomic<int> next(0);
    auto work = [&]() {
        Counters cnt;
        for (;;) {
            int v = next.fetch_add(1);
            if (v >= f->h) break;
            for (int u = 0; u < f->w; ++u) {
                RayTree &t = f->trees[(size_t)v * f->w + u];
                t.root = build_ray_tree(scene, camera_ray(*cam, u, v, 0.f, 0.f, false), depth, t, cnt);
            }
        }
    };
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) pool.emplace_back(work);
    for (auto &th : pool) th.join();
    return f;
}

void orc_forest_free(void *h) { delete (Forest *)h; }

// render_forest (:121-127) when ids == NULL, render_forest_filter (:129-145) otherwise.
// Materials and ambient are read from `desc` NOW (they may have been edited since build).
// Returns the number of trees evaluated.
long long orc_forest_eval(void *h, const rt_scene_desc *desc, const float *ambient, const int32_t *ids, int n_ids,
                          float *rgb_f32, uint8_t *rgb_u8) {
    Forest *f = (Forest *)h;
    Scene scene;
    if (!scene_from_desc(*desc, scene)) return -1;
    Color amb = ambient ? Color(ambient[0], ambient[1], ambient[2]) : scene.ambient;
    long long evaluated = 0;
    for (int v = 0; v < f->h; ++v)
        for (int u = 0; u < f->w; ++u) {
            const RayTree &t = f->trees[(size_t)v * f->w + u];
            if (ids) {
                bool touched = false;
                for (int k = 0; k < n_ids && !touched; ++k) touched = t.shapes.count(ids[k]) != 0;
                if (!touched) continue;
            }
            ++evaluated;
            Vector3 dir;
            Color c = render_ray_tree(scene, t, t.root, amb, dir);
            size_t o = ((size_t)v * f->w + u) * 3;
            if (rgb_f32) { rgb_f32[o] = c.r; rgb_f32[o + 1] = c.g; rgb_f32[o + 2] = c.b; }
            if (rgb_u8) {
                rgb_u8[o] = as_u8_channel(c.r);
                rgb_u8[o + 1] = as_u8_channel(c.g);
                rgb_u8[o + 2] = as_u8_channel(c.b);
            }
        }
    return evaluated;
}

// RayForest::stats (:73-93): out[8] = num_trees, num_intersections, smallest, largest, median, p90, p95, p99
void orc_forest_stats(void *h, uint64_t *out) {
    Forest *f = (Forest *)h;
    std::vector<uint64_t> sizes;
    for (const RayTree &t : f->trees) sizes.push_back(t.size());
    std::sort(sizes.begin(), sizes.end());
    uint64_t total = 0;
    for (uint64_t s : sizes) total += s;
    size_t n = sizes.size();
    out[0] = n;
    out[1] = total;
    out[2] = sizes.front();
    out[3] = sizes.back();
    out[4] = sizes[n / 2];
    out[5] = sizes[(size_t)(0.9f * (float)n)];
    out[6] = sizes[(size_t)(0.95f * (float)n)];
    out[7] = sizes[(size_t)(0.99f * (float)n)];
}

// RayForest::trees_with (:68-70)
uint64_t orc_forest_trees_with(void *h, int32_t shape_id) {
    Forest *f = (Forest *)h;
    uint64_t n = 0;
    for (const RayTree &t : f->trees) n += t.shapes.count(shape_id) ? 1 : 0;
    return n;
}

// RayTree::size of one pixel (:39-48), for the ray_tree_size style checks
uint64_t orc_forest_tree_size(void *h, int u, int v) {
    Forest *f = (Forest *)h;
    return f->trees[(size_t)v * f->w + u].size();
}

} // extern "C"
