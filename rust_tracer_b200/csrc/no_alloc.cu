// no_alloc.cu -- C ABI: rt_forest_trees_with_static (include/rt_b200.h), a variant of
// rt_forest_trees_with (api_render.cu; RayForest::trees_with, src/render_tree.rs:68-70) that
// allocates nothing.  It has no counterpart in the reference: it was added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
// serving loops, harnesses that forbid it).
//
// rt_forest_trees_with marks the trees that hold a shape in a per-tree byte mask from the
// forest's workspace (grown with cudaMalloc on first use), uploads the id through a cudaMalloc'd
// word and counts the mask on the host.  Here the mask and the count live in a file-scope
// __device__ array of fixed size, the id travels as a kernel argument and the count is made on
// the device; the only transfer is the 8-byte result.
//
// The array exists once per GPU that loads the library (RT_STATIC_TREES_MAX + 256 bytes, held
// until the process exits) and is shared by every call: it keeps the last call's contents, so
// each call clears the part it uses first, and two calls that overlap on one GPU (different
// streams or threads) race on it -- callers serialise them per device.
#include <mutex>

#include "api_internal.h"

namespace rt {

constexpr int STATIC_TREES_MAX = RT_STATIC_TREES_MAX;  // include/rt_b200.h
constexpr int STATIC_COUNT_AT = STATIC_TREES_MAX;       // the 8-byte count sits behind the mask, 256-byte aligned
static_assert(STATIC_TREES_MAX % 256 == 0, "the count behind the mask must stay aligned");

// In host code this name is only a stand-in for the symbol, NOT a device address: the host side
// below asks cudaGetSymbolAddress for the address, once per device.
static __device__ __align__(256) unsigned char g_tree_scratch[STATIC_TREES_MAX + 256];

// forest_mark_kernel (wavefront.cu) for one id passed by value
__global__ void forest_mark_one_kernel(Pools p, int n, int id, unsigned char *mask) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (p.t_hitid[i] == id && id >= 0) mask[p.t_root[i]] = 1;
}

// number of marked trees: one atomicAdd per warp that saw any
__global__ void mask_count_kernel(const unsigned char *mask, int n, unsigned long long *count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int b = __ballot_sync(0xffffffffu, i < n && mask[i] != 0);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(count, (unsigned long long)__popc(b));
}

namespace {

constexpr int MAX_DEVICES = 64;
std::mutex scratch_mu;
unsigned char *scratch_of[MAX_DEVICES] = {}; // device address of g_tree_scratch, per device

// `device` must be current
int scratch_for(int device, unsigned char **out) {
    if (device < 0 || device >= MAX_DEVICES) return fail(RT_ERR_INVALID, "device %d", device);
    std::lock_guard<std::mutex> lock(scratch_mu);
    if (!scratch_of[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_tree_scratch);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_tree_scratch)");
        if (!p) return fail(RT_ERR_INTERNAL, "no device address for the static scratch");
        scratch_of[device] = (unsigned char *)p;
    }
    *out = scratch_of[device];
    return RT_OK;
}

} // namespace
} // namespace rt

using namespace rt;

extern "C" int rt_forest_trees_with_static(rt_forest *forest, int32_t shape_id, uint64_t *out) {
    if (!forest || !out) return fail(RT_ERR_INVALID, "null argument");
    const int n = forest->n_samples;
    if (n > STATIC_TREES_MAX)
        return fail(RT_ERR_INVALID, "the forest has %d trees; rt_forest_trees_with_static holds at most %d (RT_STATIC_TREES_MAX): use rt_forest_trees_with",
                    n, STATIC_TREES_MAX);
    rt_scene *scene = forest->scene;
    cudaError_t e = cudaSetDevice(scene->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    unsigned char *scratch = nullptr;
    int rc = scratch_for(scene->device, &scratch);
    if (rc != RT_OK) return rc;
    cudaStream_t st = scene->stream;
    unsigned long long *count = (unsigned long long *)(scratch + STATIC_COUNT_AT);
    // the array still holds the previous call's marks and count
    if (n > 0) e = cudaMemsetAsync(scratch, 0, (size_t)n, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(count, 0, sizeof(*count), st);
    if (e != cudaSuccess) return cuda_fail(e, "scratch clear");
    const int rays = forest->total_tree_rays;
    if (rays > 0 && n > 0) {
        forest_mark_one_kernel<<<(unsigned int)((rays + 255) / 256), 256, 0, st>>>(forest->ws.pools, rays, shape_id, scratch);
        mask_count_kernel<<<(unsigned int)((n + 255) / 256), 256, 0, st>>>(scratch, n, count);
        if ((e = cudaGetLastError()) != cudaSuccess) return cuda_fail(e, "trees_with_static launch");
    }
    unsigned long long h = 0;
    e = cudaMemcpyAsync(&h, count, sizeof(h), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return cuda_fail(e, "trees_with_static");
    *out = (uint64_t)h;
    return RT_OK;
}
