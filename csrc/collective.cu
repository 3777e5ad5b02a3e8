// Small cross-GPU kernels that complete the fused compute+collective paths: epoch-flag signal /
// wait over NVLink peer mappings, and the owner-side reduction of a reduce-scatter whose partial
// tiles were pushed by the GEMM epilogues of all ranks (sw_igemm_rs).
#include "../include/swarm_b200.h"
#include "common.cuh"
#include "host_util.h"

#include <stdlib.h>

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
    int sysfence;           // SW_PEER_SYSFENCE=1: every thread runs a system-scope fence before the flag (A/B switch)
};

__device__ __forceinline__ int resolve_epoch(const PeerFlags& f) {
    return f.epoch + (f.epoch_base ? __ldcg(f.epoch_base) * f.epoch_mul : 0);
}

// Everything this rank wrote before (on this stream) becomes visible to the peers, then its flag
// slot on every peer is raised to `epoch`.
__global__ void k_peer_signal(const PeerFlags f) {
    // Everything earlier kernels of this stream wrote (TMA stores included) must be visible to the peers before the
    // flag: one system-scope fence per signalling thread, then the release store.
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
    int slices;             // CTAs per transfer
    unsigned int* tickets;  // [8] arrival counters of this launch's transfers (zero between launches)
    int dst[8];
    const uint4* src[8];
    uint4* dst_ptr[8];
    long long n16[8];
};

// Arrival counters for the sliced transfers.  Launches of one process are stream-ordered per gang, and the last CTA
// of a transfer re-arms its counter, so a small rotating pool is enough (slot = launch sequence number % 64).
__device__ unsigned int g_xfer_tickets[64 * 8];

// blockIdx.x = transfer * slices + slice.  A large transfer (the K/V of a video clip's frames is tens of MB) is
// cut into `slices` contiguous pieces so that it is carried by many SMs' store pipes instead of one; the slice
// that finishes last publishes the whole transfer and waits for the partner's.
__global__ void __launch_bounds__(1024) k_peer_exchange(const PeerFlags f, const Xfers x) {
    const int t = blockIdx.x / x.slices, sl = blockIdx.x - t * x.slices;
    if (t >= x.count) return;
    const uint4* __restrict__ src = x.src[t];
    uint4* __restrict__ dst = x.dst_ptr[t];
    const long long per = (x.n16[t] + x.slices - 1) / x.slices;
    const long long lo = per * sl, hi = lo + per < x.n16[t] ? lo + per : x.n16[t];
    long long i = lo + threadIdx.x;
    for (; i + 3 * 1024 < hi; i += 4 * 1024) {  // four independent 16-byte loads in flight per thread
        const uint4 a = src[i], b = src[i + 1024], c = src[i + 2048], d = src[i + 3072];
        dst[i] = a, dst[i + 1024] = b, dst[i + 2048] = c, dst[i + 3072] = d;
    }
    for (; i < hi; i += 1024) dst[i] = src[i];
    // ordering: every thread's stores -> CTA barrier -> (sliced: device-scope fence + ticket) -> ONE st.release.sys,
    // which is cumulative over everything that happened before it.  No system-scope fence per CTA: on a box with
    // peer mappings each one stalls for microseconds.
    if (f.sysfence)
        __threadfence_system();
    else if (x.slices > 1)
        __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        bool last = true;
        if (x.slices > 1) {
            last = atomicAdd(&x.tickets[t], 1u) == unsigned(x.slices) - 1u;
            if (last) {
                x.tickets[t] = 0u;
                __threadfence();  // (the other slices' stores, observed through the ticket, precede the flag)
            }
        }
        if (last) {
            const int epoch = resolve_epoch(f);
            st_release_sys(f.flags[x.dst[t]] + f.rank, epoch);
            wait_flag_ge(f.flags[f.rank] + x.dst[t], epoch);
        }
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
    static int env = -1;
    if (env < 0) {
        const char* e = getenv("SW_PEER_SYSFENCE");
        env = e ? atoi(e) : 0;
    }
    f->sysfence = env;
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
    long long biggest = 0;
    for (int i = 0; i < count; ++i) biggest = x.n16[i] > biggest ? x.n16[i] : biggest;
    // one slice per 256 KB, at most 16 per transfer and ~one wave of the chip in total
    int slices = int((biggest * 16 + (256 << 10) - 1) / (256 << 10));
    if (slices > 16) slices = 16;
    if (slices * count > 128) slices = 128 / count;
    if (slices < 1) slices = 1;
    {
        static int env = -1;  // SW_XFER_SLICES=n caps the CTAs per transfer (A/B switch; 1 = the single-CTA copy)
        if (env < 0) {
            const char* e = getenv("SW_XFER_SLICES");
            env = e ? atoi(e) : 0;
        }
        if (env > 0 && slices > env) slices = env;
    }
    x.slices = slices;
    static unsigned int* ticket_base[64] = {nullptr};  // per device: the symbol has one instance on each
    static unsigned int seq = 0;
    int dev = 0;
    SW_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return swh::fail(-1, "sw_peer_exchange: device index %d out of range", dev);
    if (!ticket_base[dev])  // (first call on a device: resolve the address outside any stream capture)
        SW_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&ticket_base[dev]), g_xfer_tickets));
    x.tickets = ticket_base[dev] + (seq++ % 64) * 8;
    k_peer_exchange<<<count * slices, 1024, 0, static_cast<cudaStream_t>(stream)>>>(f, x);
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
