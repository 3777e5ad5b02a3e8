// Small cross-GPU kernels that complete the fused compute+collective paths: epoch-flag signal /
// wait over NVLink peer mappings, and the owner-side reduction of a reduce-scatter whose partial
// tiles were pushed by the GEMM epilogues of all ranks (sw_igemm_rs).
#include "../include/swarm_b200.h"
#include "common.cuh"
#include "host_util.h"

namespace {

using namespace sw;

__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void wait_flag_ge(const int* p, int v) {
    for (unsigned i = 0; ld_acquire_sys(p) < v; ++i) {
        if (i > (1u << 27)) __trap();  // bounded: a lost peer must not hang the box
        __nanosleep(64);
    }
}

struct PeerFlags {
    int world, rank, epoch;
    int* flags[8];
    const int* epoch_base;  // optional device counter: epoch += *epoch_base * epoch_mul
    int epoch_mul;
};

__device__ __forceinline__ int resolve_epoch(const PeerFlags& f) {
    return f.epoch + (f.epoch_base ? __ldcg(f.epoch_base) * f.epoch_mul : 0);
}

// Everything this rank wrote before (on this stream) becomes visible to the peers, then its flag
// slot on every peer is raised to `epoch`.
__global__ void k_peer_signal(const PeerFlags f) {
    __threadfence_system();
    if (threadIdx.x < f.world) st_release_sys(f.flags[threadIdx.x] + f.rank, resolve_epoch(f));
}

__global__ void k_peer_wait(const PeerFlags f) {
    if (threadIdx.x < f.world) wait_flag_ge(f.flags[f.rank] + threadIdx.x, resolve_epoch(f));
}

// Pairwise exchange (halo rows between neighbouring patches, the per-step epsilon all-gather): one
// CTA per transfer copies `nbytes` from local memory into the destination rank's buffer over its
// NVLink mapping (16-byte stores), publishes them (fence + release flag on the destination), then
// waits for the flag the destination raises after ITS transfer to this rank has landed.  Every
// rank runs the same kernel at the same point of its stream, so the waits cannot cycle.
struct Xfers {
    int count;
    int dst[8];
    const uint4* src[8];
    uint4* dst_ptr[8];
    long long n16[8];
};

__global__ void __launch_bounds__(1024) k_peer_exchange(const PeerFlags f, const Xfers x) {
    const int t = blockIdx.x;
    if (t >= x.count) return;
    const uint4* __restrict__ src = x.src[t];
    uint4* __restrict__ dst = x.dst_ptr[t];
    for (long long i = threadIdx.x; i < x.n16[t]; i += blockDim.x) dst[i] = src[i];
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int epoch = resolve_epoch(f);
        st_release_sys(f.flags[x.dst[t]] + f.rank, epoch);
        wait_flag_ge(f.flags[f.rank] + x.dst[t], epoch);
    }
}

// out[r, :] = sum_src staging[src][r][:] (+ bias) (+ residual[r, :]), fp32 accumulation in rank order.
__global__ void k_rs_reduce(const __half* __restrict__ staging, __half* __restrict__ out,
                            const __half* __restrict__ bias, const __half* __restrict__ residual, int m_local,
                            int ncols, int ld_stage, int ldo, int ldr, const PeerFlags f) {
    __shared__ int ready;
    if (threadIdx.x < f.world) wait_flag_ge(f.flags[f.rank] + threadIdx.x, resolve_epoch(f));
    __syncthreads();
    (void)ready;
    const int octs = ncols >> 3;
    const long long total = (long long)m_local * octs;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int o = int(i % octs);
        const long long r = i / octs;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.0f;
        for (int src = 0; src < f.world; ++src) {
            const uint4 v = __ldcg(reinterpret_cast<const uint4*>(
                staging + ((size_t)src * m_local + r) * ld_stage + o * 8));
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t2 = unpack_h2(w[j]);
                acc[2 * j] += t2.x;
                acc[2 * j + 1] += t2.y;
            }
        }
        if (bias) {
            const uint4 v = *reinterpret_cast<const uint4*>(bias + o * 8);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t2 = unpack_h2(w[j]);
                acc[2 * j] += t2.x;
                acc[2 * j + 1] += t2.y;
            }
        }
        if (residual) {
            const uint4 v = *reinterpret_cast<const uint4*>(residual + r * ldr + o * 8);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t2 = unpack_h2(w[j]);
                acc[2 * j] += t2.x;
                acc[2 * j + 1] += t2.y;
            }
        }
        *reinterpret_cast<uint4*>(out + r * ldo + o * 8) =
            make_uint4(pack_h2(acc[0], acc[1]), pack_h2(acc[2], acc[3]), pack_h2(acc[4], acc[5]),
                       pack_h2(acc[6], acc[7]));
    }
}

int fill(PeerFlags* f, const sw_peers* p) {
    if (!p || p->world < 1 || p->world > 8) return swh::fail(-1, "bad peer description");
    memset(f, 0, sizeof(*f));
    f->world = p->world;
    f->rank = p->rank;
    f->epoch = p->epoch;
    f->epoch_base = p->epoch_base;
    f->epoch_mul = p->epoch_mul;
    for (int d = 0; d < p->world; ++d) f->flags[d] = p->flags[d];
    return 0;
}

}  // namespace

extern "C" int sw_peer_signal(const sw_peers* peers, void* stream) {
    PeerFlags f;
    if (int rc = fill(&f, peers)) return rc;
    k_peer_signal<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(f);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int sw_peer_wait(const sw_peers* peers, void* stream) {
    PeerFlags f;
    if (int rc = fill(&f, peers)) return rc;
    k_peer_wait<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(f);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int sw_peer_exchange(const sw_peers* peers, const sw_xfer* xfers, int count, void* stream) {
    PeerFlags f;
    if (int rc = fill(&f, peers)) return rc;
    if (count < 0 || count > 8 || (count > 0 && !xfers)) return swh::fail(-1, "sw_peer_exchange: 0..8 transfers");
    if (count == 0) return 0;
    Xfers x;
    memset(&x, 0, sizeof(x));
    x.count = count;
    for (int i = 0; i < count; ++i) {
        const sw_xfer& t = xfers[i];
        const bool empty = t.nbytes == 0;  // flag-only: a pairwise barrier
        if (t.dst < 0 || t.dst >= peers->world || t.dst == peers->rank || t.nbytes < 0 || t.nbytes % 16 ||
            (!empty && (!t.src || !t.dst_ptr ||
                        (reinterpret_cast<uintptr_t>(t.src) | reinterpret_cast<uintptr_t>(t.dst_ptr)) % 16)))
            return swh::fail(-1, "sw_peer_exchange: transfer %d is malformed (16-byte granularity, dst != rank)", i);
        for (int j = 0; j < i; ++j)
            if (xfers[j].dst == t.dst) return swh::fail(-1, "sw_peer_exchange: one transfer per destination");
        x.dst[i] = t.dst;
        x.src[i] = static_cast<const uint4*>(t.src);
        x.dst_ptr[i] = static_cast<uint4*>(t.dst_ptr);
        x.n16[i] = t.nbytes / 16;
    }
    k_peer_exchange<<<count, 1024, 0, static_cast<cudaStream_t>(stream)>>>(f, x);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int sw_rs_reduce(const void* staging_local, void* out, const void* bias, const void* residual,
                            int m_local, int ncols, int ld_stage, int ldo, int ldr, const sw_peers* peers,
                            void* stream) {
    PeerFlags f;
    if (int rc = fill(&f, peers)) return rc;
    if (ncols % 8 || ld_stage % 8 || ldo % 8 || (residual && ldr % 8))
        return swh::fail(-1, "sw_rs_reduce: columns / leading dimensions must be multiples of 8");
    const long long total = (long long)m_local * (ncols / 8);
    int blocks = int((total + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    k_rs_reduce<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __half*>(staging_local), static_cast<__half*>(out), static_cast<const __half*>(bias),
        static_cast<const __half*>(residual), m_local, ncols, ld_stage, ldo, ldr, f);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}
