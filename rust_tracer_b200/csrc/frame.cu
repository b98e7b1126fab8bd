// frame.cu -- C ABI: the multi-GPU frame (include/rt_b200.h: rt_frame_*).
//
// The reference writes every pixel into one RenderBuffer (src/render.rs:31-38).  With the frame
// split over GPUs that buffer lives on the GPU that owns the image; the other ranks map it (CUDA
// IPC, peer access over NVLink / NVSwitch) and their resolve kernels store rows straight into
// it.  "Rank r has finished frame #e" and "the owner has consumed frame #e" are 32-bit words
// behind the pixels, written and polled by one-warp kernels in stream order -- no host barrier,
// no collective.
#include <cstdio>
#include <cstring>

#include "api_internal.h"

namespace rt {

#define CK(call, what)                                        \
    do {                                                      \
        cudaError_t e__ = (call);                             \
        if (e__ != cudaSuccess) return cuda_fail(e__, what);  \
    } while (0)

static size_t signal_offset(size_t bytes) { return (bytes + 255) & ~(size_t)255; }

__device__ __forceinline__ unsigned int load_acquire_sys(const unsigned int *p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Everything earlier on the stream (the resolve kernel's stores into the frame) happened before
// this thread; the system-scope release makes it visible to whoever acquires the word.
__global__ void frame_signal_kernel(unsigned int *word, unsigned int value) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(word), "r"(value) : "memory");
}

// One warp: lane l watches words l, l + 32, ... until each has reached `value` (wrap-safe), or
// the time is up.  *status (pinned host memory) = 0 done, 1 timed out.
__global__ void frame_wait_kernel(const unsigned int *words, int n, unsigned int value, unsigned long long timeout_ns,
                                  int *status) {
    const unsigned long long t0 = global_timer_ns();
    bool late = false;
    for (int i = threadIdx.x; i < n && !late; i += 32) {
        while ((int)(load_acquire_sys(words + i) - value) < 0) {
            if (global_timer_ns() - t0 > timeout_ns) {
                late = true;
                break;
            }
            __nanosleep(100);
        }
    }
    const unsigned any_late = __ballot_sync(0xffffffffu, late);
    if (threadIdx.x == 0) *status = any_late ? 1 : 0;
}

struct WaitStatus {
    int *host = nullptr; // pinned, mapped: the wait kernel's verdict
};
static thread_local WaitStatus g_wait;

} // namespace rt

using namespace rt;

extern "C" {

int rt_frame_alloc(int32_t device, size_t bytes, void **d_ptr) {
    if (!d_ptr || bytes == 0) return fail(RT_ERR_INVALID, "null argument");
    *d_ptr = nullptr;
    CK(cudaSetDevice(device), "cudaSetDevice");
    const size_t total = signal_offset(bytes) + RT_FRAME_SIGNAL_WORDS * sizeof(unsigned int);
    cudaError_t e = cudaMalloc(d_ptr, total); // its own allocation: an IPC handle names a whole cudaMalloc block
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(e == cudaErrorMemoryAllocation ? RT_ERR_NOMEM : RT_ERR_CUDA, "cudaMalloc(frame, %zu bytes): %s", total,
                    cudaGetErrorString(e));
    }
    e = cudaMemset((char *)*d_ptr + signal_offset(bytes), 0, RT_FRAME_SIGNAL_WORDS * sizeof(unsigned int));
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        cudaFree(*d_ptr);
        *d_ptr = nullptr;
        return cuda_fail(e, "cudaMemset(frame signals)");
    }
    return RT_OK;
}

int rt_frame_export(int32_t device, void *d_ptr, uint8_t handle[RT_FRAME_HANDLE_BYTES]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_FRAME_HANDLE_BYTES, "handle size");
    if (!d_ptr || !handle) return fail(RT_ERR_INVALID, "null argument");
    CK(cudaSetDevice(device), "cudaSetDevice");
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, d_ptr), "cudaIpcGetMemHandle");
    std::memcpy(handle, &h, sizeof(h));
    return RT_OK;
}

int rt_frame_import(int32_t device, const uint8_t handle[RT_FRAME_HANDLE_BYTES], void **d_ptr) {
    if (!d_ptr || !handle) return fail(RT_ERR_INVALID, "null argument");
    *d_ptr = nullptr;
    CK(cudaSetDevice(device), "cudaSetDevice");
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    cudaError_t e = cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        cudaGetLastError();
        *d_ptr = nullptr;
        return cuda_fail(e, "cudaIpcOpenMemHandle");
    }
    return RT_OK;
}

int rt_frame_release(int32_t device, void *d_ptr, int32_t imported) {
    if (!d_ptr) return RT_OK;
    CK(cudaSetDevice(device), "cudaSetDevice");
    CK(imported ? cudaIpcCloseMemHandle(d_ptr) : cudaFree(d_ptr), "frame release");
    return RT_OK;
}

int rt_frame_signal(int32_t device, void *d_ptr, size_t bytes, int32_t word, uint32_t value, void *stream) {
    if (!d_ptr || bytes == 0 || word < 0 || word >= RT_FRAME_SIGNAL_WORDS) return fail(RT_ERR_INVALID, "bad frame signal word");
    CK(cudaSetDevice(device), "cudaSetDevice");
    unsigned int *words = (unsigned int *)((char *)d_ptr + signal_offset(bytes));
    frame_signal_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(words + word, value);
    CK(cudaGetLastError(), "frame_signal_kernel");
    return RT_OK;
}

int rt_frame_wait(int32_t device, void *d_ptr, size_t bytes, int32_t first_word, int32_t n_words, uint32_t value,
                  uint32_t timeout_ms, void *stream) {
    if (!d_ptr || bytes == 0 || first_word < 0 || n_words < 1 || first_word + n_words > RT_FRAME_SIGNAL_WORDS)
        return fail(RT_ERR_INVALID, "bad frame signal range");
    CK(cudaSetDevice(device), "cudaSetDevice");
    if (!g_wait.host) CK(cudaHostAlloc((void **)&g_wait.host, sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable), "cudaHostAlloc(wait status)");
    int *d_status = nullptr;
    CK(cudaHostGetDevicePointer((void **)&d_status, g_wait.host, 0), "cudaHostGetDevicePointer");
    *g_wait.host = -1;
    const unsigned int *words = (const unsigned int *)((char *)d_ptr + signal_offset(bytes));
    if (timeout_ms == 0 || timeout_ms > 60000u) timeout_ms = 60000u; // never spin on a GPU for longer than a minute
    frame_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(words + first_word, n_words, value, (unsigned long long)timeout_ms * 1000000ull,
                                                          d_status);
    CK(cudaGetLastError(), "frame_wait_kernel");
    CK(cudaStreamSynchronize((cudaStream_t)stream), "frame wait");
    if (*(volatile int *)g_wait.host != 0)
        return fail(RT_ERR_TIMEOUT, "frame signal words [%d, %d) did not reach %u within %u ms", first_word, first_word + n_words, value,
                    timeout_ms);
    return RT_OK;
}

} // extern "C"
