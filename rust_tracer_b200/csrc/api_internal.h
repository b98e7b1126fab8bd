// api_internal.h -- host-side state behind the opaque handles of include/rt_b200.h.
#pragma once
#include <memory>
#include <mutex>
#include <vector>

#include "../../include/rt_b200.h"
#include "kernels.h"
#include "tuning.h"

namespace rt {

int fail(int code, const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);
int convert_materials(const rt_material *in, int n, std::vector<DevMaterial> &out);

// Device pools of one wavefront (kernels.h: Pools) plus the per-call scratch.
struct Workspace {
    Pools pools{};
    float4 *sample_rgb = nullptr; // one colour per sample of the current band
    long long sample_cap = 0;
    int *rows = nullptr;          // device copy of the rendered rows
    int rows_cap = 0;
    unsigned char *mask = nullptr; // per-sample "re-shade" flags (filtered forest eval)
    long long mask_cap = 0;
    int *idx = nullptr;            // the re-shaded pixels of a filtered forest eval (+ their count behind the list)
    long long idx_cap = 0;
    unsigned int *h_counters = nullptr; // pinned mirror of pools.counters
    unsigned long long *exact_counter = nullptr;
    unsigned int *batch_counters = nullptr; // one work counter per sweep launch of a band
    int batch_cap = 0, batch_next = 0;
    void *out_f32 = nullptr, *out_u8 = nullptr; // device staging for host-buffer calls
    size_t out_f32_cap = 0, out_u8_cap = 0;
    void *tables = nullptr; // the device tables of the scene that holds this workspace (api_scene.cu: Arena)
    size_t tables_cap = 0;
    // Streams and events travel with the workspace (and its per-device cache) instead of being created and
    // destroyed per scene / per call: with 8 devices driven from 8 threads of one process those driver calls
    // queue up behind each other (rt_multi_destroy took 3 ms, a frame's device span 9.2 instead of 7.3 ms).
    static constexpr int MAX_AUX = 8, MAX_EVENTS = 512;
    cudaStream_t stream = nullptr;          // the scene's own stream
    cudaStream_t aux[MAX_AUX] = {};         // side streams of a level's sweeps
    int n_aux = 0;
    cudaEvent_t events[MAX_EVENTS] = {};    // handed out per call by EventPool (api_render.cu)
    int n_events = 0;

    int ensure_pools(long long t_cap, long long s_cap);
    int ensure_samples(long long n);
    int ensure_rows(int n);
    int ensure_mask(long long n);
    int ensure_idx(long long n);
    int ensure_misc();
    int ensure_batch_counters(int n);
    void release_pools();
    void release();
};

// One released workspace per device is kept instead of freed: its ray pools are gigabytes,
// and cudaMalloc / cudaFree of them would dominate a create -> render -> destroy call.
// rt_trim() frees the cache.
// Host copy of one bound slot (api_scene.cu: make_bound), kept to re-centre the table.
struct HostSlot {
    float cx, cy, cz;
    double rf;  // radius incl. slack
    int kind;   // 0 bound, 1 pad (passes no ray), 2 beyond the filter's range (passes every ray)
};

void drop_worker_pool(); // api_multi.cu: joins the idle worker threads kept between rt_multi objects
Workspace workspace_take_cached(int device);
void workspace_give_cached(int device, Workspace &ws);

} // namespace rt

namespace rt {

// Everything rt_scene_create derives from a scene description on the host, before any device is
// touched: the structure-of-arrays tables of scene_dev.cuh, the bound / group / quad spheres and
// one re-centred table per point light.  Built once and uploaded to as many devices as render
// the scene (rt_multi_create shares one among its devices).
struct HostScene {
    DevScene dev{};  // scalar fields only; the table pointers are set per device by upload_scene
    int n_materials = 0;
    int counts[4] = {0, 0, 0, 0}; // shapes per rt_shape_kind
    int planes_identity = 0;
    // what a bound slot of segment g holds (rt_shape_kind; the sweeps count filter evaluations per segment)
    int seg_kind[MAX_SEGS] = {0, 0, 0, 0};
    int seg_real[MAX_SEGS] = {0, 0, 0, 0}; // bound slots of the segment that hold a shape (the rest are pads)
    std::vector<float4> bound, group, quad, super;
    std::vector<int> bound_shape;
    std::vector<int2> slot_meta;
    std::vector<float4> slot_geom;
    std::vector<int4> meta;
    std::vector<float4> geom;
    std::vector<int> plane_shape;
    std::vector<float4> plane_data, cube_tris;
    std::vector<DevMaterial> materials;
    std::vector<DevLight> lights;
    // origin tables (kernels.h: OriginTable); empty when the table is too small to pay
    bool origin = false;
    std::vector<HostSlot> slots, groups, quads, supers;
    std::vector<std::vector<float4>> light_host; // by DevLight::point_index: slots, groups, quads, supers re-centred on the light
    std::vector<OriginTable> light_tabs;         // their scalar fields (table pointers null)
    // the table of the camera origin of the last render, shared by the devices of a multi-GPU scene
    mutable std::mutex cam_mu;
    mutable bool cam_valid = false;
    mutable float cam_origin[3] = {0.f, 0.f, 0.f};
    mutable std::vector<float4> cam_host;
    mutable OriginTable cam_tab;
    // Everything upload_scene sends to a device, packed in upload order (api_scene.cu: Arena, 256-byte aligned) so that ONE
    // copy does it; page-locked on first upload (cudaHostRegister): one DMA per device instead of ~19 staged copies.
    std::vector<char> blob;
    mutable std::mutex blob_mu;
    mutable bool blob_pin_tried = false, blob_pinned = false;
    HostScene() = default;
    HostScene(const HostScene &) = delete;
    HostScene &operator=(const HostScene &) = delete;
    ~HostScene();
};

int build_host_scene(const rt_scene_desc *desc, HostScene &hs);
// The tables for `desc`: those of the last description built if it is the same one byte for byte, else a new build.
int host_scene_for(const rt_scene_desc *desc, std::shared_ptr<const HostScene> *out);
void drop_scene_cache();
} // namespace rt

struct rt_scene {
    int device = 0;
    rt::DevScene dev{};
    std::shared_ptr<const rt::HostScene> host;
    rt::DevMaterial *d_materials = nullptr;
    std::vector<rt::DevMaterial> host_materials; // what d_materials currently holds
    int n_materials = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;     // the stream work is launched on
    cudaStream_t own_stream = nullptr; // the workspace's stream (ws.stream)
    unsigned int launches = 0;         // kernels launched by the current call
    int forests = 0;                   // live rt_forest handles built from this scene (they use its stream and tables)
    bool destroy_pending = false;      // rt_scene_destroy was called while forests were alive: the last forest frees it
    rt::Workspace ws;
    std::vector<rt::OriginTable> light_tabs; // by DevLight::point_index, device pointers set
    float4 *cam_tab = nullptr;
    rt::OriginTable cam;                     // bound == cam_tab once built
    float cam_origin[3] = {0.f, 0.f, 0.f};
};

namespace rt {
// Upload a built scene to `device`: one allocation, one batch of copies.
int upload_scene(const std::shared_ptr<const HostScene> &hs, int device, rt_scene **out);
// Re-centre the bound, group and quad tables on `o` (out = slots, groups, quads, supers, each in
// the paired layout of DevScene::bound) and fill every field of *tab but the table pointers; toward: the table serves shadow rays travelling towards `o`.
void build_origin_table(const HostScene &hs, const float o[3], std::vector<float4> &out, OriginTable *tab, bool toward);
// Make scene->cam the table of this camera origin (no-op when origin tables are off).
int prepare_camera_table(rt_scene *scene, const float origin[3]);
// Device outputs of a render (either may be null).
struct Outputs {
    float *d_f32 = nullptr;
    unsigned char *d_u8 = nullptr;
};
int check_params(const rt_scene *scene, const rt_camera *cam, const rt_render_params *pr);
// The whole render of one call on the scene's device; returns after the work is complete.
int render_impl(rt_scene *scene, const rt_camera *cam, const rt_render_params *pr, Outputs out, rt_stats *stats);
int ensure_staging(Workspace &ws, size_t n_ch, bool want_f32, bool want_u8);
// Free a scene whose last user is gone (rt_scene_destroy, or the last forest of a destroyed scene).
void scene_free(rt_scene *scene);
} // namespace rt

struct rt_forest {
    rt_scene *scene = nullptr;
    rt::Workspace ws;   // owns the pools of the kept ray trees
    rt::FrameParams fp{};
    std::vector<int> tb, te; // tree-ray range of each level
    std::vector<int> sb, se; // shadow-ray range of each level
    bool light_major = false; // layout of a level's shadow rays (api_render.cu: BandLevels)
    int total_tree_rays = 0;
    int n_samples = 0;       // = trees = pixels (spp is 1 in forest mode)
    int width = 0, n_rows = 0;
};
