// GroupNorm entry point for callers that must not allocate device memory.
//
// Some users of this library cannot hand it a scratch buffer because they cannot obtain one:
// a serving process that gave its whole device-memory budget to its own arena at start-up and
// must stay inside that plan, or a caller on a latency-critical path where cudaMalloc/cudaFree
// (which may synchronise the device) are off limits.  For them the statistics scratch of
// sw_groupnorm_nhwc lives in a fixed array inside the library image instead.  Nothing here calls
// cudaMalloc, cudaFree or any allocating routine; the price is a hard size limit and one shared
// buffer per GPU.
#include "../include/swarm_b200.h"
#include "host_util.h"

#include <mutex>

// defined in norm.cu: statistics + apply kernels on a caller-chosen scratch pointer
int sw_groupnorm_launch(const void* x, const void* x2, void* y, const void* gamma, const void* beta,
                        int n, int hw, int c1, int c2, int groups, float eps, int silu,
                        float* scratch, cudaStream_t st, const sw_peers* peers);

namespace {

// 1 MiB per GPU that loads the library, held until the process exits.
static __device__ __align__(256) float g_gn_static_scratch[SW_GN_STATIC_MAX_FLOATS];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
float* g_scratch_ptr[kMaxDevices] = {nullptr};

// In host code `g_gn_static_scratch` is only a handle for the runtime, not an address a kernel
// can dereference; the real device address is looked up once per device (with that device
// current) and cached.
int scratch_for_current_device(float** out) {
    int dev = -1;
    SW_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices)
        return swh::fail(-1, "sw_groupnorm_nhwc_static: device index %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (g_scratch_ptr[dev] == nullptr) {
        float* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_gn_static_scratch);
        if (e != cudaSuccess || p == nullptr)
            return swh::fail(int(e ? e : cudaErrorInvalidSymbol),
                             "sw_groupnorm_nhwc_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                             cudaGetErrorString(e));
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return 0;
}

}  // namespace

extern "C" int sw_groupnorm_nhwc_static(const void* x, const void* x2, void* y, const void* gamma,
                                        const void* beta, int n, int hw, int c1, int c2, int groups,
                                        float eps, int silu, void* stream) {
    if (!x || !y) return swh::fail(-1, "sw_groupnorm_nhwc_static: null argument");
    const int64_t need = sw_groupnorm_scratch_floats(n, hw, c1 + c2, groups);
    if (need < 0) return swh::fail(-1, "sw_groupnorm_nhwc_static: bad shape n=%d hw=%d c=%d groups=%d", n, hw,
                                   c1 + c2, groups);
    if (need > int64_t(SW_GN_STATIC_MAX_FLOATS))
        return swh::fail(-1,
                         "sw_groupnorm_nhwc_static: n=%d hw=%d c=%d groups=%d needs %lld scratch floats, the "
                         "built-in buffer holds %d; use sw_groupnorm_nhwc with your own scratch",
                         n, hw, c1 + c2, groups, (long long)need, int(SW_GN_STATIC_MAX_FLOATS));
    float* scratch = nullptr;
    int rc = scratch_for_current_device(&scratch);
    if (rc) return rc;
    // The buffer still holds the previous call's numbers.  Only the per-image ticket counters
    // have to start at zero, and the launch below clears exactly those on `stream` first.
    return sw_groupnorm_launch(x, x2, y, gamma, beta, n, hw, c1, c2, groups, eps, silu, scratch,
                               static_cast<cudaStream_t>(stream), nullptr);
}
