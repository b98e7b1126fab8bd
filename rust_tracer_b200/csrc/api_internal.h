// api_internal.h -- host-side state behind the opaque handles of include/rt_b200.h.
#pragma once
#include <vector>

#include "../../include/rt_b200.h"
#include "kernels.h"

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
    unsigned int *h_counters = nullptr; // pinned mirror of pools.counters
    unsigned long long *exact_counter = nullptr;
    unsigned int *batch_counters = nullptr; // one work counter per sweep launch of a band
    int batch_next = 0;
    void *out_f32 = nullptr, *out_u8 = nullptr; // device staging for host-buffer calls
    size_t out_f32_cap = 0, out_u8_cap = 0;
    void *tables = nullptr; // the device tables of the scene that holds this workspace (api_scene.cu: Arena)
    size_t tables_cap = 0;

    int ensure_pools(long long t_cap, long long s_cap);
    int ensure_samples(long long n);
    int ensure_rows(int n);
    int ensure_mask(long long n);
    int ensure_misc();
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

Workspace workspace_take_cached(int device);
void workspace_give_cached(int device, Workspace &ws);

} // namespace rt

struct rt_scene {
    int device = 0;
    rt::DevScene dev{};
    std::vector<void *> allocs;
    rt::DevMaterial *d_materials = nullptr;
    int n_materials = 0;
    int counts[4] = {0, 0, 0, 0}; // shapes per rt_shape_kind
    int planes_identity = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;     // the stream work is launched on
    cudaStream_t own_stream = nullptr; // created with the scene
    std::vector<cudaStream_t> aux;     // the sweeps of one wavefront level run side by side on these
    unsigned int launches = 0;         // kernels launched by the current call
    rt::Workspace ws;
    // origin tables (kernels.h: OriginTable): one per point light, built with the scene, and
    // one for the camera origin of the last render.  Empty when the table is too small to pay.
    std::vector<rt::HostSlot> slots, groups, quads;
    std::vector<rt::OriginTable> light_tabs; // by DevLight::point_index
    float4 *cam_tab = nullptr;
    rt::OriginTable cam;                     // bound == cam_tab once built
    float cam_origin[3] = {0.f, 0.f, 0.f};
};

namespace rt {
// Re-centre the scene's bound table and group table on `o` (out = slots, groups, quads, each in
// the paired layout of DevScene::bound) and fill every field of *tab but the table pointers; toward: the table serves shadow rays travelling towards `o`.
void build_origin_table(const rt_scene *scene, const float o[3], std::vector<float4> &out, OriginTable *tab, bool toward);
// Make scene->cam the table of this camera origin (no-op when origin tables are off).
int prepare_camera_table(rt_scene *scene, const float origin[3]);
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
