// api_multi.cu -- C ABI: one scene on several GPUs of the box, driven from ONE host call
// (include/rt_b200.h: rt_multi_*, rt_render_multi).
//
// The reference renders a frame with one blocking call, render::render(&camera, &scene,
// &mut buffer, depth) (src/render.rs:31-38, caller src/main.rs:244-261).  rt_multi_render is
// that call for a host that owns 1-8 GPUs: the scene tables are built once on the host and
// uploaded to every device, each device renders the interleaved row tiles
// (v / tile_rows) % n == i of the frame -- pixels are independent, no collective -- and
//   * rt_multi_render: every device copies its own finished rows straight to their place in the
//     caller's host image over its own PCIe link (one strided copy per device, in parallel), or
//   * rt_multi_render_device: every device's resolve kernel stores its rows into ONE frame
//     on the first device over NVLink / NVSwitch peer access (RT_FLAG_FRAME_ROWS).
// One persistent worker thread per device carries that device's CUDA context, stream and
// rt_scene; the calling thread only hands out the job and waits.
#include <condition_variable>
#include <cstring>
#include <functional>
#include <string>
#include <thread>

#include "api_internal.h"

namespace rt {

#define CK(call, what)                                        \
    do {                                                      \
        cudaError_t e__ = (call);                             \
        if (e__ != cudaSuccess) return cuda_fail(e__, what);  \
    } while (0)

// A thread that runs one job at a time; the caller waits for its status.
class Worker {
  public:
    Worker() : th_([this] { loop(); }) {}
    ~Worker() {
        {
            std::lock_guard<std::mutex> lock(mu_);
            quit_ = true;
        }
        cv_.notify_all();
        th_.join();
    }
    void post(std::function<int()> job) {
        {
            std::lock_guard<std::mutex> lock(mu_);
            job_ = std::move(job);
            busy_ = true;
        }
        cv_.notify_all();
    }
    // status of the last job; `error` receives its rt_last_error() text when it failed
    int wait(std::string *error) {
        std::unique_lock<std::mutex> lock(mu_);
        cv_.wait(lock, [this] { return !busy_; });
        if (rc_ != RT_OK && error) *error = err_;
        return rc_;
    }

  private:
    void loop() {
        for (;;) {
            std::function<int()> job;
            {
                std::unique_lock<std::mutex> lock(mu_);
                cv_.wait(lock, [this] { return quit_ || (busy_ && job_); });
                if (quit_) return;
                job = std::move(job_);
                job_ = nullptr;
            }
            int rc;
            std::string err;
            try {
                rc = job();
                if (rc != RT_OK) err = rt_last_error(); // thread-local: carry it to the caller's thread
            } catch (...) {
                rc = RT_ERR_INTERNAL;
                err = "exception in a device worker";
            }
            {
                std::lock_guard<std::mutex> lock(mu_);
                rc_ = rc;
                err_ = err;
                busy_ = false;
            }
            cv_.notify_all();
        }
    }
    std::mutex mu_;
    std::condition_variable cv_;
    std::function<int()> job_;
    bool busy_ = false, quit_ = false;
    int rc_ = RT_OK;
    std::string err_;
    std::thread th_; // last: starts when everything above exists
};

// Idle workers are kept for the next rt_multi_create: starting and joining 8 threads was ~0.4 ms of the 1.3 ms that the
// one-shot call on 8 devices spent outside its devices' render (profiles/r02_e2e_laps_multi_device.log).  A worker is
// not tied to a device (every job selects its own).  rt_trim(-1) joins them.
namespace {
std::mutex g_pool_mu;
std::vector<std::unique_ptr<Worker>> g_pool;
} // namespace

static std::unique_ptr<Worker> worker_take() {
    {
        std::lock_guard<std::mutex> lock(g_pool_mu);
        if (!g_pool.empty()) {
            std::unique_ptr<Worker> w = std::move(g_pool.back());
            g_pool.pop_back();
            return w;
        }
    }
    return std::unique_ptr<Worker>(new Worker());
}
static void worker_give(std::unique_ptr<Worker> w) {
    std::lock_guard<std::mutex> lock(g_pool_mu);
    if (g_pool.size() < 64) g_pool.push_back(std::move(w)); // (else: joined here, by its destructor)
}
void drop_worker_pool() {
    std::vector<std::unique_ptr<Worker>> drop;
    {
        std::lock_guard<std::mutex> lock(g_pool_mu);
        drop.swap(g_pool);
    }
}

} // namespace rt

using namespace rt;

struct rt_multi {
    std::vector<int> devices;
    std::vector<rt_scene *> scenes;               // scenes[i] lives on devices[i], touched by workers[i] only
    std::vector<std::unique_ptr<Worker>> workers;
    std::vector<char> peer_ok;                    // device i may store into memory of devices[0]
    std::shared_ptr<const HostScene> host;
};

namespace {

// Run job(i) on every worker, wait for all; the first failure is reported on the calling thread.
int run_all(rt_multi *m, const std::function<int(int)> &job) {
    const int n = (int)m->devices.size();
    for (int i = 0; i < n; ++i) m->workers[i]->post([&job, i] { return job(i); });
    int rc = RT_OK;
    std::string err;
    for (int i = 0; i < n; ++i) {
        std::string e;
        const int r = m->workers[i]->wait(&e);
        if (r != RT_OK && rc == RT_OK) {
            rc = r;
            err = "device " + std::to_string(m->devices[i]) + ": " + e;
        }
    }
    return rc == RT_OK ? RT_OK : fail(rc, "%s", err.c_str());
}

void add_stats(rt_stats &sum, const rt_stats &s) {
    sum.pixels += s.pixels; sum.rays += s.rays; sum.shadow_rays += s.shadow_rays; sum.nodes += s.nodes;
    sum.sphere_tests += s.sphere_tests; sum.plane_tests += s.plane_tests; sum.triangle_tests += s.triangle_tests;
    sum.cube_tests += s.cube_tests; sum.exact_tests += s.exact_tests; sum.algorithmic_flops += s.algorithmic_flops;
    sum.kernel_launches += s.kernel_launches; sum.filter_flops += s.filter_flops; sum.tested_pairs += s.tested_pairs;
    sum.bound_evals_sphere += s.bound_evals_sphere; sum.bound_evals_triangle += s.bound_evals_triangle;
    sum.bound_evals_cube += s.bound_evals_cube; sum.hier_evals += s.hier_evals;
    sum.exact_sphere_tests += s.exact_sphere_tests; sum.exact_triangle_tests += s.exact_triangle_tests;
    sum.exact_cube_tests += s.exact_cube_tests; sum.exact_flops += s.exact_flops; sum.executed_flops += s.executed_flops;
    // device times: the slowest device is the time of the frame
    sum.sweep_ms = std::max(sum.sweep_ms, s.sweep_ms);
    sum.shade_ms = std::max(sum.shade_ms, s.shade_ms);
    sum.total_ms = std::max(sum.total_ms, s.total_ms);
}

// Rows of device i's compact tile buffer -> their place in a whole host frame: full tiles with one
// strided copy (compact tile k = frame tile k n + i), a ragged last tile with a plain one.
template <typename T>
cudaError_t scatter_rows_to_host(T *host, const T *dev, int i, int n, int tile_rows, int height, int width, cudaStream_t st) {
    const size_t row = (size_t)width * 3 * sizeof(T), tile = row * (size_t)tile_rows;
    const int my_rows = rt_tile_rows_for_rank(height, i, n, tile_rows, nullptr, 0);
    const int full = my_rows / tile_rows, rest = my_rows - full * tile_rows;
    cudaError_t e = cudaSuccess;
    if (full > 0)
        e = cudaMemcpy2DAsync((char *)host + (size_t)i * tile, tile * (size_t)n, dev, tile, tile, (size_t)full, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && rest > 0)
        e = cudaMemcpyAsync((char *)host + ((size_t)full * n + i) * tile, (const char *)dev + (size_t)full * tile, row * (size_t)rest,
                            cudaMemcpyDeviceToHost, st);
    return e;
}

int check_multi(const rt_multi *m, const rt_camera *camera, const rt_render_params *params) {
    if (!m || !camera || !params) return fail(RT_ERR_INVALID, "null argument");
    if (params->rank != 0 || params->world != 1)
        return fail(RT_ERR_INVALID, "rt_multi_render splits the frame over its own devices: rank / world must be 0 / 1");
    if (params->flags & RT_FLAG_FRAME_ROWS) return fail(RT_ERR_INVALID, "RT_FLAG_FRAME_ROWS is set by the library here");
    return check_params(m->scenes[0], camera, params);
}

} // namespace

extern "C" {

int rt_multi_create(const rt_scene_desc *desc, const int32_t *devices, int32_t n_devices, rt_multi **out) {
    if (!desc || !out || (n_devices > 0 && !devices)) return fail(RT_ERR_INVALID, "null argument");
    *out = nullptr;
    if (n_devices < 1 || n_devices > 64) return fail(RT_ERR_INVALID, "%d devices", n_devices);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount (no CPU fallback exists)");
    for (int i = 0; i < n_devices; ++i) {
        if (devices[i] < 0 || devices[i] >= ndev) return fail(RT_ERR_CUDA, "CUDA device %d not present (%d devices)", devices[i], ndev);
        for (int j = 0; j < i; ++j)
            if (devices[j] == devices[i]) return fail(RT_ERR_INVALID, "device %d listed twice", devices[i]);
    }
    std::shared_ptr<const HostScene> hs;
    rt_multi *m = nullptr;
    try {
        m = new rt_multi();
        m->devices.assign(devices, devices + n_devices);
        m->scenes.assign(n_devices, nullptr);
        m->peer_ok.assign(n_devices, 0);
        for (int i = 0; i < n_devices; ++i) m->workers.push_back(worker_take());
    } catch (...) {
        delete m;
        return fail(RT_ERR_NOMEM, "out of host memory or threads");
    }
    int rc = host_scene_for(desc, &hs); // once, for all devices
    if (rc != RT_OK) {
        delete m;
        return rc;
    }
    m->host = hs;
    rc = run_all(m, [m](int i) {
        int rc = upload_scene(m->host, m->devices[i], &m->scenes[i]);
        if (rc != RT_OK) return rc;
        // may this device store into the first one's memory?  (rt_multi_render_device only)
        if (i == 0) m->peer_ok[i] = 1;
        else {
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, m->devices[i], m->devices[0]) == cudaSuccess && can) {
                cudaError_t e = cudaDeviceEnablePeerAccess(m->devices[0], 0);
                if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) m->peer_ok[i] = 1;
            }
            cudaGetLastError();
        }
        return (int)RT_OK;
    });
    if (rc != RT_OK) {
        rt_multi_destroy(m);
        return rc;
    }
    *out = m;
    return RT_OK;
}

void rt_multi_destroy(rt_multi *m) {
    if (!m) return;
    if (!m->workers.empty()) {
        for (size_t i = 0; i < m->workers.size(); ++i)
            m->workers[i]->post([m, i] {
                rt_scene_destroy(m->scenes[i]);
                m->scenes[i] = nullptr;
                return (int)RT_OK;
            });
        for (auto &w : m->workers) w->wait(nullptr);
        for (auto &w : m->workers) worker_give(std::move(w)); // idle now: kept for the next rt_multi_create
        m->workers.clear();
    }
    delete m;
}

int32_t rt_multi_device_count(const rt_multi *m) { return m ? (int32_t)m->devices.size() : fail(RT_ERR_INVALID, "null argument"); }

int rt_multi_update_materials(rt_multi *m, const rt_material *materials, int32_t n) {
    if (!m) return fail(RT_ERR_INVALID, "null argument");
    return run_all(m, [=](int i) { return rt_scene_update_materials(m->scenes[i], materials, n); });
}

int rt_multi_render(rt_multi *m, const rt_camera *camera, const rt_render_params *params, float *rgb_f32, uint8_t *rgb_u8,
                    rt_stats *stats) {
    int rc = check_multi(m, camera, params);
    if (rc != RT_OK) return rc;
    const int n = (int)m->devices.size();
    std::vector<rt_stats> st(n);
    rc = run_all(m, [&](int i) {
        rt_scene *scene = m->scenes[i];
        rt_render_params pr = *params;
        pr.rank = i;
        pr.world = n;
        CK(cudaSetDevice(scene->device), "cudaSetDevice");
        const int32_t my_rows = rt_tile_rows_for_rank(camera->y_res, i, n, pr.tile_rows, nullptr, 0);
        const size_t n_ch = (size_t)my_rows * (size_t)camera->x_res * 3;
        int rc = ensure_staging(scene->ws, n_ch, rgb_f32 != nullptr, rgb_u8 != nullptr);
        if (rc != RT_OK) return rc;
        Outputs out;
        out.d_f32 = rgb_f32 ? (float *)scene->ws.out_f32 : nullptr;
        out.d_u8 = rgb_u8 ? (unsigned char *)scene->ws.out_u8 : nullptr;
        if ((rc = render_impl(scene, camera, &pr, out, &st[i])) != RT_OK) return rc;
        // this device's rows, straight to their place in the caller's image, over this device's own PCIe link
        cudaStream_t s = scene->stream;
        if (rgb_u8) CK(scatter_rows_to_host(rgb_u8, out.d_u8, i, n, pr.tile_rows, camera->y_res, camera->x_res, s), "D2H rgb_u8");
        if (rgb_f32) CK(scatter_rows_to_host(rgb_f32, out.d_f32, i, n, pr.tile_rows, camera->y_res, camera->x_res, s), "D2H rgb_f32");
        CK(cudaStreamSynchronize(s), "D2H");
        return (int)RT_OK;
    });
    if (rc != RT_OK) return rc;
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        for (const rt_stats &s : st) add_stats(*stats, s);
    }
    return RT_OK;
}

int rt_multi_render_device(rt_multi *m, const rt_camera *camera, const rt_render_params *params, void *d_rgb_f32, void *d_rgb_u8,
                           rt_stats *stats) {
    int rc = check_multi(m, camera, params);
    if (rc != RT_OK) return rc;
    const int n = (int)m->devices.size();
    for (int i = 0; i < n; ++i)
        if (!m->peer_ok[i])
            return fail(RT_ERR_UNSUPPORTED, "device %d has no peer access to device %d: use rt_multi_render (host buffers)", m->devices[i],
                        m->devices[0]);
    std::vector<rt_stats> st(n);
    rc = run_all(m, [&](int i) {
        rt_render_params pr = *params;
        pr.rank = i;
        pr.world = n;
        pr.flags |= RT_FLAG_FRAME_ROWS; // every row goes to its own place in the first device's frame, over NVLink
        Outputs out;
        out.d_f32 = (float *)d_rgb_f32;
        out.d_u8 = (unsigned char *)d_rgb_u8;
        return render_impl(m->scenes[i], camera, &pr, out, &st[i]); // returns after this device's stores are complete
    });
    if (rc != RT_OK) return rc;
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        for (const rt_stats &s : st) add_stats(*stats, s);
    }
    return RT_OK;
}

int rt_render_multi(const rt_scene_desc *desc, const int32_t *devices, int32_t n_devices, const rt_camera *camera,
                    const rt_render_params *params, float *rgb_f32, uint8_t *rgb_u8, rt_stats *stats) {
    rt_multi *m = nullptr;
    int rc = rt_multi_create(desc, devices, n_devices, &m);
    if (rc != RT_OK) return rc;
    rc = rt_multi_render(m, camera, params, rgb_f32, rgb_u8, stats);
    if (rc != RT_OK) { // keep the message of the failure across the teardown
        const std::string msg = rt_last_error();
        rt_multi_destroy(m);
        return fail(rc, "%s", msg.c_str());
    }
    rt_multi_destroy(m);
    return RT_OK;
}

} // extern "C"
