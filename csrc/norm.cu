// GroupNorm(+SiLU) and LayerNorm over channels-last fp16 activations (memory-bound; 16-byte
// vector accesses, fp32 statistics).  The tensors of a diffusion UNet fit in the 126 MB L2, so
// the statistics pass and the apply pass both stream from L2, not HBM.
#include "../include/swarm_b200.h"
#include "common.cuh"
#include "host_util.h"

#include <cooperative_groups.h>

#include <stdlib.h>

namespace {

using namespace sw;

// Cross-GPU extension of GroupNorm (sequence / patch parallel: every rank holds H/world rows of
// the feature map, the statistics are global).  The all-reduce of the per-(image, group) sums is
// done by the GroupNorm kernels themselves over NVLink peer mappings: the statistics kernel's
// finishing CTA stores its sums into every peer's slot and raises an epoch flag with
// st.release.sys; the apply kernel polls the flags with ld.acquire.sys and adds the slots in rank
// order (deterministic).  world == 1 disables all of it.
struct GnDist {
    int world, rank, epoch;
    float* peer_stats[8];  // rank d: float [2 (epoch parity)][world][n][groups][2]
    int* peer_flags[8];    // rank d: int [world], slot [src] written by rank src
    const int* epoch_base;  // optional device counter: epoch += *epoch_base * epoch_mul (graph replays)
    int epoch_mul;
    int sysfence;           // SW_PEER_SYSFENCE=1: system-scope fence by every thread before the flag (A/B switch)
};

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
        if (i > (1u << 27)) __trap();  // a lost peer must not hang the box
        __nanosleep(64);
    }
}

__device__ __forceinline__ int gn_epoch(const GnDist& d) {
    return d.epoch + (d.epoch_base ? __ldcg(d.epoch_base) * d.epoch_mul : 0);
}

// ---- GroupNorm statistics, deterministic: grid (chunks, n).  Thread <-> fixed channel octet,
// strided rows.  No floating-point atomics anywhere: per-thread sums go to shared memory, are
// combined in a fixed order into per-(chunk, group) partials, and the last CTA of each image
// (ticket counter) folds the partials -- again in a fixed order -- into final[n][group][2].
// The input may be the channel-concatenation of two tensors (UNet skip connections): octets below
// c1 come from x1, the rest from x2.
// scratch layout (floats): final[n*groups*2] | tickets[n] (as uint) | partial[n][chunks][groups][2]
__global__ void k_gn_stats(const __half* __restrict__ x1, const __half* __restrict__ x2, int c1,
                           float* __restrict__ scratch, int hw, int c, int groups, int rows_per_cta,
                           int rows_per_pass, const GnDist dist) {
    pdl_trigger();
    pdl_wait();
    extern __shared__ float sh[];  // [rows_per_pass][c][2] then [c][2]
    __shared__ bool is_last;
    const int octs = c >> 3;
    const int n = blockIdx.y, nimg = gridDim.y, chunks = gridDim.x;
    const int r0 = blockIdx.x * rows_per_cta;
    const int r1 = min(hw, r0 + rows_per_cta);
    const int oct = threadIdx.x % octs;
    const int rp = threadIdx.x / octs;
    float* fin = scratch;
    unsigned int* tickets = reinterpret_cast<unsigned int*>(scratch + size_t(nimg) * groups * 2);
    float* partial = scratch + size_t(nimg) * groups * 2 + nimg + 1;

    if (rp < rows_per_pass) {
        float s[8], ss[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) s[i] = ss[i] = 0.0f;
        const bool first = oct * 8 < c1;
        const int cs = first ? c1 : c - c1;
        const __half* base =
            (first ? x1 + oct * 8 : x2 + (oct * 8 - c1)) + (size_t(n) * hw) * cs;
        // four independent 16-byte loads in flight per thread (the loop is latency-, not bandwidth-bound)
        int r = r0 + rp;
        for (; r + 3 * rows_per_pass < r1; r += 4 * rows_per_pass) {
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                v[u] = *reinterpret_cast<const uint4*>(base + size_t(r + u * rows_per_pass) * cs);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = unpack_h2(w[j]);
                    s[2 * j] += f.x;
                    ss[2 * j] += f.x * f.x;
                    s[2 * j + 1] += f.y;
                    ss[2 * j + 1] += f.y * f.y;
                }
            }
        }
        for (; r < r1; r += rows_per_pass) {
            const uint4 v = *reinterpret_cast<const uint4*>(base + size_t(r) * cs);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_h2(w[j]);
                s[2 * j] += f.x;
                ss[2 * j] += f.x * f.x;
                s[2 * j + 1] += f.y;
                ss[2 * j + 1] += f.y * f.y;
            }
        }
        float* dst = sh + (size_t(rp) * c + oct * 8) * 2;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            dst[2 * i] = s[i];
            dst[2 * i + 1] = ss[i];
        }
    }
    __syncthreads();
    // per-channel totals over the row passes (fixed order)
    float* chs = sh + size_t(rows_per_pass) * c * 2;
    for (int ch = threadIdx.x; ch < c; ch += blockDim.x) {
        float a = 0.0f, b = 0.0f;
        for (int k = 0; k < rows_per_pass; ++k) {
            a += sh[(size_t(k) * c + ch) * 2];
            b += sh[(size_t(k) * c + ch) * 2 + 1];
        }
        chs[2 * ch] = a;
        chs[2 * ch + 1] = b;
    }
    __syncthreads();
    const int cg = c / groups;
    float* my_partial = partial + (size_t(n) * chunks + blockIdx.x) * groups * 2;
    for (int g = threadIdx.x; g < groups; g += blockDim.x) {
        float a = 0.0f, b = 0.0f;
        for (int k = 0; k < cg; ++k) {
            a += chs[2 * (g * cg + k)];
            b += chs[2 * (g * cg + k) + 1];
        }
        my_partial[2 * g] = a;
        my_partial[2 * g + 1] = b;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicAdd(&tickets[n], 1u);
        is_last = (t == unsigned(chunks) - 1u);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // last CTA of this image: fold the partials, one warp per group, lanes over chunks,
    // then a fixed shuffle tree.
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const float* pimg = partial + size_t(n) * chunks * groups * 2;
    for (int g = wid; g < groups; g += nw) {
        float a = 0.0f, b = 0.0f;
        for (int k = lane; k < chunks; k += 32) {
            a += __ldcg(pimg + (size_t(k) * groups + g) * 2);
            b += __ldcg(pimg + (size_t(k) * groups + g) * 2 + 1);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
        }
        if (lane == 0) {
            fin[(size_t(n) * groups + g) * 2] = a;
            fin[(size_t(n) * groups + g) * 2 + 1] = b;
            if (dist.world > 1) {
                // push this rank's sums into every peer's slot [parity][rank][n][g]
                const size_t slot = ((size_t(gn_epoch(dist) & 1) * dist.world + dist.rank) * nimg + n) * groups + g;
                for (int d = 0; d < dist.world; ++d) {
                    dist.peer_stats[d][slot * 2] = a;
                    dist.peer_stats[d][slot * 2 + 1] = b;
                }
            }
        }
    }
    if (threadIdx.x == 0) tickets[n] = 0u;  // leave the scratch ready for the next call
    if (dist.world > 1) {
        if (dist.sysfence)
            __threadfence_system();
        else
            __threadfence();  // device scope ahead of the ticket; the flag store below is the sys-scope release
        __syncthreads();
        if (threadIdx.x == 0) {
            // the image that finishes last raises this rank's flag on every peer
            const unsigned int t = atomicAdd(&tickets[nimg], 1u);
            if (t == unsigned(nimg) - 1u) {
                tickets[nimg] = 0u;
                const int epoch = gn_epoch(dist);
                for (int d = 0; d < dist.world; ++d) st_release_sys(dist.peer_flags[d] + dist.rank, epoch);
            }
        }
    }
}

// ---- GroupNorm apply: y = silu?((x-mean)*rstd*gamma+beta)
__global__ void k_gn_apply(const __half* __restrict__ x1, const __half* __restrict__ x2, int c1,
                           __half* __restrict__ y,
                           const __half* __restrict__ gamma, const __half* __restrict__ beta,
                           const float* __restrict__ stats, int hw, int c, int groups, float eps,
                           int silu, int rows_per_cta, int rows_per_pass, const GnDist dist) {
    pdl_trigger();
    pdl_wait();
    extern __shared__ float sh[];  // a[c], b[c]
    float* sa = sh;
    float* sb = sh + c;
    const int n = blockIdx.y;
    const int cg = c / groups;
    const float inv_cnt = 1.0f / (float(hw) * float(cg) * float(dist.world > 1 ? dist.world : 1));
    const int epoch = dist.world > 1 ? gn_epoch(dist) : 0;
    if (dist.world > 1) {
        if (threadIdx.x < dist.world) wait_flag_ge(dist.peer_flags[dist.rank] + threadIdx.x, epoch);
        __syncthreads();
    }
    for (int ch = threadIdx.x; ch < c; ch += blockDim.x) {
        const int g = ch / cg;
        float sum, sq;
        if (dist.world > 1) {
            sum = sq = 0.0f;
            const float* mine = dist.peer_stats[dist.rank];
            for (int src = 0; src < dist.world; ++src) {  // fixed rank order: same bits on every rank
                const size_t slot = ((size_t(epoch & 1) * dist.world + src) * gridDim.y + n) * groups + g;
                sum += __ldcg(mine + slot * 2);
                sq += __ldcg(mine + slot * 2 + 1);
            }
        } else {
            sum = stats[(size_t(n) * groups + g) * 2];
            sq = stats[(size_t(n) * groups + g) * 2 + 1];
        }
        const float mean = sum * inv_cnt;
        const float var = fmaxf(sq * inv_cnt - mean * mean, 0.0f);
        const float rstd = rsqrtf(var + eps);
        const float ga = gamma ? __half2float(gamma[ch]) : 1.0f;
        const float be = beta ? __half2float(beta[ch]) : 0.0f;
        sa[ch] = rstd * ga;
        sb[ch] = be - mean * rstd * ga;
    }
    __syncthreads();
    const int octs = c >> 3;
    const int oct = threadIdx.x % octs;
    const int rp = threadIdx.x / octs;
    if (rp >= rows_per_pass) return;
    const int r0 = blockIdx.x * rows_per_cta;
    const int r1 = min(hw, r0 + rows_per_cta);
    float a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        a[i] = sa[oct * 8 + i];
        b[i] = sb[oct * 8 + i];
    }
    const size_t base = (size_t(n) * hw) * c + oct * 8;
    const bool first = oct * 8 < c1;
    const int cs = first ? c1 : c - c1;
    const __half* src = (first ? x1 + oct * 8 : x2 + (oct * 8 - c1)) + (size_t(n) * hw) * cs;
    auto apply_row = [&](const uint4& v, int r) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_h2(w[j]);
            float u0 = fmaf(f.x, a[2 * j], b[2 * j]);
            float u1 = fmaf(f.y, a[2 * j + 1], b[2 * j + 1]);
            if (silu) {
                u0 = silu_f(u0);
                u1 = silu_f(u1);
            }
            o[j] = pack_h2(u0, u1);
        }
        *reinterpret_cast<uint4*>(y + base + size_t(r) * c) = make_uint4(o[0], o[1], o[2], o[3]);
    };
    int r = r0 + rp;
    for (; r + 3 * rows_per_pass < r1; r += 4 * rows_per_pass) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = *reinterpret_cast<const uint4*>(src + size_t(r + u * rows_per_pass) * cs);
#pragma unroll
        for (int u = 0; u < 4; ++u) apply_row(v[u], r + u * rows_per_pass);
    }
    for (; r < r1; r += rows_per_pass) apply_row(*reinterpret_cast<const uint4*>(src + size_t(r) * cs), r);
}

// ---- Fused GroupNorm: ONE launch, the tensor is read from global memory once.
// One thread-block cluster per image: each CTA parks its rows in shared memory while it accumulates the
// statistics, pushes its per-group partial sums into every CTA of the cluster through distributed shared
// memory, and after one hardware cluster barrier folds the partials in rank order (deterministic) and
// normalises its rows straight out of shared memory.  No global scratch, no atomics, no second read.
// The host takes this path whenever an image fits in the cluster's shared memory (clusters of up to
// 16 CTAs): every GroupNorm of the SD / SDXL UNets except the widest skip-concats at full resolution.
__global__ void __launch_bounds__(1024)
    k_gn_fused(const __half* __restrict__ x1, const __half* __restrict__ x2, int c1, __half* __restrict__ y,
               const __half* __restrict__ gamma, const __half* __restrict__ beta, int hw, int c, int groups, float eps,
               int silu, int rows_per_cta, int rows_per_pass, int c_total) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    if (threadIdx.x == 0) {
        TR(0);
        TRG(11);
    }
    pdl_trigger();
    pdl_wait();
    if (threadIdx.x == 0) TR(1);
    extern __shared__ __align__(16) uint8_t gsm[];
    __shared__ float s_all[16 * 64 * 2];  // [rank][group][2], written by every CTA of the cluster
    __shared__ float s_mean[64], s_rstd[64];
    const int octs = c >> 3;
    const int n = blockIdx.y;
    const int csize = int(cluster.num_blocks());
    const int crank = int(cluster.block_rank());
    const int r0 = crank * rows_per_cta;
    const int r1 = min(hw, r0 + rows_per_cta);
    const int oct = threadIdx.x % octs;
    const int rp = threadIdx.x / octs;
    uint4* slab = reinterpret_cast<uint4*>(gsm);
    float* sh = reinterpret_cast<float*>(slab + size_t(rows_per_cta) * octs);  // [rows_per_pass][c][2]
    float* chs = sh + size_t(rows_per_pass) * c * 2;                            // [c][2]

    // blockIdx.z selects a slice of whole groups (`c` channels, `groups` groups of the c_total-wide tensor):
    // groups are independent, so slices only multiply the number of SMs pulling the tensor in
    const int ch0 = int(blockIdx.z) * c + oct * 8;  // first of this thread's 8 channels, in the full tensor
    const bool active = rp < rows_per_pass;
    const bool first = ch0 < c1;
    const int cs = first ? c1 : c_total - c1;
    const __half* src = (first ? x1 + ch0 : x2 + (ch0 - c1)) + (size_t(n) * hw) * cs;
    // affine parameters of this thread's 8 channels: in flight while the rows stream in
    uint4 gv = make_uint4(0x3C003C00u, 0x3C003C00u, 0x3C003C00u, 0x3C003C00u), bv = make_uint4(0, 0, 0, 0);
    if (active) {
        if (gamma) gv = *reinterpret_cast<const uint4*>(gamma + ch0);
        if (beta) bv = *reinterpret_cast<const uint4*>(beta + ch0);
    }
    if (active) {
        float s[8], ss[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) s[i] = ss[i] = 0.0f;
        auto acc = [&](const uint4& v) {
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_h2(w[j]);
                s[2 * j] += f.x;
                ss[2 * j] += f.x * f.x;
                s[2 * j + 1] += f.y;
                ss[2 * j + 1] += f.y * f.y;
            }
        };
        int r = r0 + rp;
        for (; r + 7 * rows_per_pass < r1; r += 8 * rows_per_pass) {
            uint4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = *reinterpret_cast<const uint4*>(src + size_t(r + u * rows_per_pass) * cs);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                slab[size_t(r + u * rows_per_pass - r0) * octs + oct] = v[u];
                acc(v[u]);
            }
        }
        for (; r < r1; r += rows_per_pass) {
            const uint4 v = *reinterpret_cast<const uint4*>(src + size_t(r) * cs);
            slab[size_t(r - r0) * octs + oct] = v;
            acc(v);
        }
        float* dst = sh + (size_t(rp) * c + oct * 8) * 2;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            dst[2 * i] = s[i];
            dst[2 * i + 1] = ss[i];
        }
    }
    if (threadIdx.x == 0) TR(2);
    __syncthreads();
    if (threadIdx.x == 0) TR(3);
    // per-channel totals over the row slots, in a fixed order (bit-reproducible) but spread over the whole CTA:
    // stage 1 folds slots part, part + 8, ... into slot `part`; stage 2 folds those eight
    constexpr int kParts = 8;
    for (int i = threadIdx.x; i < c * kParts; i += blockDim.x) {
        const int ch = i % c, part = i / c;
        if (part < rows_per_pass) {
            float a = 0.0f, b = 0.0f;
            for (int k = part; k < rows_per_pass; k += kParts) {
                a += sh[(size_t(k) * c + ch) * 2];
                b += sh[(size_t(k) * c + ch) * 2 + 1];
            }
            // (slot `part` is only read by this thread in stage 1: safe to overwrite in place)
            sh[(size_t(part) * c + ch) * 2] = a;
            sh[(size_t(part) * c + ch) * 2 + 1] = b;
        }
    }
    __syncthreads();
    for (int ch = threadIdx.x; ch < c; ch += blockDim.x) {
        float a = 0.0f, b = 0.0f;
        const int np = rows_per_pass < kParts ? rows_per_pass : kParts;
        for (int k = 0; k < np; ++k) {
            a += sh[(size_t(k) * c + ch) * 2];
            b += sh[(size_t(k) * c + ch) * 2 + 1];
        }
        chs[2 * ch] = a;
        chs[2 * ch + 1] = b;
    }
    __syncthreads();
    const int cg_ = c / groups;
    // (group, destination rank) pairs: this CTA's partial of group g lands in slot [crank][g] of rank d
    for (int i = threadIdx.x; i < groups * csize; i += blockDim.x) {
        const int g = i % groups, d = i / groups;
        float a = 0.0f, b = 0.0f;
        for (int k = 0; k < cg_; ++k) {
            a += chs[2 * (g * cg_ + k)];
            b += chs[2 * (g * cg_ + k) + 1];
        }
        float* remote = cluster.map_shared_rank(s_all, d);
        remote[(crank * 64 + g) * 2] = a;
        remote[(crank * 64 + g) * 2 + 1] = b;
    }
    if (threadIdx.x == 0) TR(4);
    cluster.sync();  // release / acquire: every partial of the image is now in this CTA's s_all
    if (threadIdx.x == 0) TR(5);
    if (threadIdx.x < groups) {
        const int g = threadIdx.x;
        float a = 0.0f, b = 0.0f;
        for (int k = 0; k < csize; ++k) {  // rank order: the same bits in every CTA and on every launch
            a += s_all[(k * 64 + g) * 2];
            b += s_all[(k * 64 + g) * 2 + 1];
        }
        const float inv_cnt = 1.0f / (float(hw) * float(cg_));
        const float mean = a * inv_cnt;
        const float var = fmaxf(b * inv_cnt - mean * mean, 0.0f);
        s_mean[g] = mean;
        s_rstd[g] = rsqrtf(var + eps);
    }
    __syncthreads();
    if (!active) return;
    float a[8], b[8];
    {
        const uint32_t gw[4] = {gv.x, gv.y, gv.z, gv.w}, bw[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 gf = unpack_h2(gw[j]), bf = unpack_h2(bw[j]);
            const int g0 = (oct * 8 + 2 * j) / cg_, g1 = (oct * 8 + 2 * j + 1) / cg_;
            a[2 * j] = s_rstd[g0] * gf.x;
            b[2 * j] = bf.x - s_mean[g0] * a[2 * j];
            a[2 * j + 1] = s_rstd[g1] * gf.y;
            b[2 * j + 1] = bf.y - s_mean[g1] * a[2 * j + 1];
        }
    }
    __half* out = y + (size_t(n) * hw) * c_total + ch0;
    for (int r = r0 + rp; r < r1; r += rows_per_pass) {
        const uint4 v = slab[size_t(r - r0) * octs + oct];
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_h2(w[j]);
            float u0 = fmaf(f.x, a[2 * j], b[2 * j]);
            float u1 = fmaf(f.y, a[2 * j + 1], b[2 * j + 1]);
            if (silu) {
                u0 = silu_f(u0);
                u1 = silu_f(u1);
            }
            o[j] = pack_h2(u0, u1);
        }
        *reinterpret_cast<uint4*>(out + size_t(r) * c_total) = make_uint4(o[0], o[1], o[2], o[3]);
    }
#ifdef SW_TRACE
    if (threadIdx.x == 0) {
        TR(6);
        TRG(12);
        trace_value(13, trace_smid());
        trace_value(14, (unsigned long long)blockDim.x | ((unsigned long long)rows_per_cta << 16) |
                            ((unsigned long long)rows_per_pass << 32));
        trace_value(15, (unsigned long long)gridDim.x | ((unsigned long long)gridDim.y << 20) |
                            ((unsigned long long)gridDim.z << 40));
    }
#endif
}

// ---- LayerNorm: one warp per row, row cached in registers (c <= 2048).
template <int MAX_OCT>
__global__ void k_layernorm(const __half* __restrict__ x, __half* __restrict__ y,
                            const __half* __restrict__ gamma, const __half* __restrict__ beta,
                            int rows, int c, float eps) {
    pdl_trigger();
    pdl_wait();
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= rows) return;
    const int octs = c >> 3;
    const __half* xr = x + size_t(warp) * c;
    float v[MAX_OCT][8];
    float sum = 0.0f;
#pragma unroll
    for (int k = 0; k < MAX_OCT; ++k) {
        const int o = lane + k * 32;
        if (o < octs) {
            const uint4 u = *reinterpret_cast<const uint4*>(xr + o * 8);
            const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_h2(w[j]);
                v[k][2 * j] = f.x;
                v[k][2 * j + 1] = f.y;
                sum += f.x + f.y;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float mean = sum / float(c);
    float var = 0.0f;
#pragma unroll
    for (int k = 0; k < MAX_OCT; ++k) {
        if (lane + k * 32 < octs) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float d = v[k][j] - mean;
                var += d * d;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) var += __shfl_xor_sync(0xffffffffu, var, o);
    const float rstd = rsqrtf(var / float(c) + eps);
    __half* yr = y + size_t(warp) * c;
#pragma unroll
    for (int k = 0; k < MAX_OCT; ++k) {
        const int o = lane + k * 32;
        if (o < octs) {
            const uint4 g = *reinterpret_cast<const uint4*>(gamma + o * 8);
            const uint4 b = *reinterpret_cast<const uint4*>(beta + o * 8);
            const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
            const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
            uint32_t out[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 gf = unpack_h2(gw[j]);
                const float2 bf = unpack_h2(bw[j]);
                out[j] = pack_h2((v[k][2 * j] - mean) * rstd * gf.x + bf.x,
                                 (v[k][2 * j + 1] - mean) * rstd * gf.y + bf.y);
            }
            *reinterpret_cast<uint4*>(yr + o * 8) = make_uint4(out[0], out[1], out[2], out[3]);
        }
    }
}

// ---- row softmax (VAE mid-block attention: one head of dim 512, scores materialised by a GEMM)
__global__ void k_softmax_rows(const __half* __restrict__ x, __half* __restrict__ y, int cols,
                               float scale) {
    __shared__ float red[32];
    const __half* xr = x + size_t(blockIdx.x) * cols;
    __half* yr = y + size_t(blockIdx.x) * cols;
    const int octs = cols >> 3;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    float mx = -INFINITY;
    for (int o = threadIdx.x; o < octs; o += blockDim.x) {
        const uint4 u = *reinterpret_cast<const uint4*>(xr + o * 8);
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_h2(w[j]);
            mx = fmaxf(mx, fmaxf(f.x, f.y));
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) red[wid] = mx;
    __syncthreads();
    mx = red[0];
    for (int i = 1; i < nw; ++i) mx = fmaxf(mx, red[i]);
    __syncthreads();
    float sum = 0.0f;
    for (int o = threadIdx.x; o < octs; o += blockDim.x) {
        const uint4 u = *reinterpret_cast<const uint4*>(xr + o * 8);
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_h2(w[j]);
            sum += __expf((f.x - mx) * scale) + __expf((f.y - mx) * scale);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) red[wid] = sum;
    __syncthreads();
    sum = 0.0f;
    for (int i = 0; i < nw; ++i) sum += red[i];
    const float inv = 1.0f / sum;
    for (int o = threadIdx.x; o < octs; o += blockDim.x) {
        const uint4 u = *reinterpret_cast<const uint4*>(xr + o * 8);
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
        uint32_t out[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_h2(w[j]);
            out[j] = pack_h2(__expf((f.x - mx) * scale) * inv, __expf((f.y - mx) * scale) * inv);
        }
        *reinterpret_cast<uint4*>(yr + o * 8) = make_uint4(out[0], out[1], out[2], out[3]);
    }
}

}  // namespace

// Launch geometry shared by the statistics and apply kernels.
struct GnGeom {
    int rows_per_pass, threads, rows_per_cta, chunks;
};
static GnGeom gn_geom(int hw, int c) {
    GnGeom g;
    const int octs = c / 8;
    g.rows_per_pass = 256 / octs;
    if (g.rows_per_pass < 1) g.rows_per_pass = 1;
    g.threads = (octs * g.rows_per_pass + 31) / 32 * 32;
    // ~32 KB of activations per CTA, but never more than 256 CTAs per image
    int rpc = (32768 / (c * 2)) / g.rows_per_pass * g.rows_per_pass;
    if (rpc < g.rows_per_pass) rpc = g.rows_per_pass;
    const int min_rpc = ((hw + 255) / 256 + g.rows_per_pass - 1) / g.rows_per_pass * g.rows_per_pass;
    if (rpc < min_rpc) rpc = min_rpc;
    g.rows_per_cta = rpc;
    g.chunks = (hw + rpc - 1) / rpc;
    return g;
}

// Geometry of the fused single-launch path; cluster == 0 when the tensor does not qualify.
struct GnFusedGeom {
    int rows_per_pass, threads, rows_per_cta, cluster, chunks, slices;
    size_t smem;
};
static GnFusedGeom gn_fused_geom(int n, int hw, int c, int groups) {
    GnFusedGeom g;
    memset(&g, 0, sizeof(g));
    static int enabled = -1;  // SW_GN_FUSED=0 keeps the two-kernel path (A/B runs)
    if (enabled < 0) {
        const char* e = getenv("SW_GN_FUSED");
        enabled = e ? atoi(e) : 1;
    }
    if (!enabled || groups > 64 || c / 8 > 1024 || n > 65535) return g;
    const int sms = swh::sm_count();
    const int cgw = c / groups;
    // try the finest channel slicing first: slices of whole groups whose width is a multiple of 8
    for (int sl = 8; sl >= 1; sl >>= 1) {
        if (groups % sl || ((groups / sl) * cgw) % 8) continue;
        const int cs = c / sl, octs = cs / 8;
        int rpp_max = 512 / octs;
        if (rpp_max < 1) rpp_max = 1;
        if (rpp_max > hw) rpp_max = hw;
        // largest cluster (<= 16 CTAs) whose CTAs all own rows; the rows of a CTA are walked in equal passes
        // (rows per pass fitted to the CTA's share instead of a fixed 512 threads' worth: with the fixed pass the
        // 64x64x320 map only qualified for clusters of 4 = 64 CTAs, and the SiLU arithmetic of the apply phase --
        // ~8 elements per clock per SM -- ran on 64 SMs; scripts/trace_gn.py).  More CTAs per image = more SMs
        // pulling the tensor in and sharing the arithmetic.
        // clusters of 4 first, then larger, then smaller: measured (scripts/trace_gn.py sweep, profiles/r2_*) a
        // cluster of 4 wins whenever an image's slice fits four CTAs' shared memory (the all-to-all exchange of
        // partial sums and the co-scheduling of 16 CTAs in one GPC cost more than the extra SMs give); bigger
        // tensors need the bigger clusters anyway
        static int first_cl = -1;  // SW_GN_MAXCL: cluster size tried first (tuning hook)
        if (first_cl < 0) {
            const char* e = getenv("SW_GN_MAXCL");
            first_cl = e ? atoi(e) : 4;
            if (first_cl != 1 && first_cl != 2 && first_cl != 4 && first_cl != 8 && first_cl != 16) first_cl = 4;
        }
        int order[5], no = 0;
        for (int cl = first_cl; cl <= 16; cl <<= 1) order[no++] = cl;
        for (int cl = first_cl >> 1; cl >= 1; cl >>= 1) order[no++] = cl;
        for (int oi = 0; oi < no; ++oi) {
            const int cl = order[oi];
            const int rows = (hw + cl - 1) / cl;
            const int passes = (rows + rpp_max - 1) / rpp_max;
            const int rpp = (rows + passes - 1) / passes;
            const int rpc = rpp * passes;
            if (cl > 1 && rpc * (cl - 1) >= hw) continue;  // trailing CTAs would own no rows
            const size_t smem = size_t(rpc) * cs * 2 + (size_t(rpp) * cs * 2 + size_t(cs) * 2) * sizeof(float);
            if (smem > 216 * 1024) continue;
            if (sl > 1 && long(n) * cl * sl > 2L * sms) break;  // enough CTAs already: use coarser slices
            g.rows_per_pass = rpp;
            g.threads = (octs * rpp + 31) / 32 * 32;
            g.rows_per_cta = rpc;
            g.cluster = cl;
            g.chunks = cl;
            g.slices = sl;
            g.smem = smem;
            return g;
        }
    }
    return g;
}

extern "C" int sw_groupnorm_kernel_count(int n, int hw, int c, int groups) {
    if (c <= 0 || c % 8 || hw <= 0 || n <= 0 || groups <= 0) return -1;
    return gn_fused_geom(n, hw, c, groups).cluster > 0 ? 1 : 2;
}

extern "C" int64_t sw_groupnorm_scratch_floats(int n, int hw, int c, int groups) {
    if (c <= 0 || c % 8 || hw <= 0 || n <= 0 || groups <= 0) return -1;
    const GnGeom g = gn_geom(hw, c);
    const GnFusedGeom f = gn_fused_geom(n, hw, c, groups);
    (void)f;  // the single-launch path needs no scratch
    return int64_t(n) * groups * 2 + n + 1 + int64_t(n) * g.chunks * groups * 2;
}

int sw_groupnorm_launch(const void* x, const void* x2, void* y, const void* gamma, const void* beta,
                        int n, int hw, int c1, int c2, int groups, float eps, int silu,
                        float* scratch, cudaStream_t st, const sw_peers* peers);

// Shared by the public entry points; `scratch` must already be usable device memory of at least
// sw_groupnorm_scratch_floats() floats.
static int gn_launch(const void* x, const void* x2, void* y, const void* gamma, const void* beta,
                     int n, int hw, int c1, int c2, int groups, float eps, int silu,
                     float* scratch, cudaStream_t st, const sw_peers* peers) {
    const int c = c1 + c2;
    GnDist dist;
    memset(&dist, 0, sizeof(dist));
    dist.world = 1;
    if (peers && peers->world > 1) {
        if (peers->world > 8) return swh::fail(-1, "sw_groupnorm_nhwc_dist: world > 8");
        dist.world = peers->world;
        dist.rank = peers->rank;
        dist.epoch = peers->epoch;
        dist.epoch_base = peers->epoch_base;
        dist.epoch_mul = peers->epoch_mul;
        {
            static int env = -1;
            if (env < 0) {
                const char* e = getenv("SW_PEER_SYSFENCE");
                env = e ? atoi(e) : 0;
            }
            dist.sysfence = env;
        }
        for (int d = 0; d < peers->world; ++d) {
            dist.peer_stats[d] = static_cast<float*>(peers->peer[d]);
            dist.peer_flags[d] = peers->flags[d];
        }
    }
    if (c1 <= 0 || c2 < 0 || c1 % 8 || (c2 > 0 && !x2))
        return swh::fail(-1, "sw_groupnorm_nhwc: bad channel split c1=%d c2=%d", c1, c2);
    if (c % 8 || groups <= 0 || c % groups)
        return swh::fail(-1, "sw_groupnorm_nhwc: c=%d groups=%d unsupported", c, groups);
    if (c / 8 > 1024) return swh::fail(-1, "sw_groupnorm_nhwc: c=%d too wide", c);
    if (dist.world == 1) {
        const GnFusedGeom f = gn_fused_geom(n, hw, c, groups);
        if (f.cluster > 0) {
            static bool configured[64] = {false};
            int dev = 0;
            SW_CUDA_OK(cudaGetDevice(&dev));
            if (dev < 64 && !configured[dev]) {
                SW_CUDA_OK(cudaFuncSetAttribute(k_gn_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
                SW_CUDA_OK(cudaFuncSetAttribute(k_gn_fused, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
                configured[dev] = true;
            }
            SW_CUDA_OK(swh::launch_pdl(k_gn_fused, dim3(f.cluster, n, f.slices), dim3(f.threads), f.smem, st, f.cluster,
                                       static_cast<const __half*>(x), static_cast<const __half*>(x2), c1,
                                       static_cast<__half*>(y), static_cast<const __half*>(gamma),
                                       static_cast<const __half*>(beta), hw, c / f.slices, groups / f.slices, eps, silu,
                                       f.rows_per_cta, f.rows_per_pass, c));
            return 0;
        }
    }
    const GnGeom g = gn_geom(hw, c);
    // only the ticket counters have to start at zero
    SW_CUDA_OK(cudaMemsetAsync(scratch + size_t(n) * groups * 2, 0, sizeof(float) * size_t(n + 1), st));
    dim3 grid(g.chunks, n);
    const size_t smem_stats = (size_t(g.rows_per_pass) * c * 2 + size_t(c) * 2) * sizeof(float);
    SW_CUDA_OK(swh::launch_pdl(k_gn_stats, grid, dim3(g.threads), smem_stats, st, 1, static_cast<const __half*>(x),
                               static_cast<const __half*>(x2), c1, scratch, hw, c, groups, g.rows_per_cta,
                               g.rows_per_pass, dist));
    SW_CUDA_OK(swh::launch_pdl(k_gn_apply, grid, dim3(g.threads), size_t(2 * c * sizeof(float)), st, 1,
                               static_cast<const __half*>(x), static_cast<const __half*>(x2), c1,
                               static_cast<__half*>(y), static_cast<const __half*>(gamma),
                               static_cast<const __half*>(beta), static_cast<const float*>(scratch), hw, c, groups,
                               eps, silu, g.rows_per_cta, g.rows_per_pass, dist));
    return 0;
}

int sw_groupnorm_launch(const void* x, const void* x2, void* y, const void* gamma, const void* beta,
                        int n, int hw, int c1, int c2, int groups, float eps, int silu,
                        float* scratch, cudaStream_t st, const sw_peers* peers) {
    return gn_launch(x, x2, y, gamma, beta, n, hw, c1, c2, groups, eps, silu, scratch, st, peers);
}

extern "C" int sw_groupnorm_nhwc(const void* x, const void* x2, void* y, const void* gamma,
                                 const void* beta, int n, int hw, int c1, int c2, int groups,
                                 float eps, int silu, float* stats_scratch, void* stream) {
    if (!x || !y || !stats_scratch) return swh::fail(-1, "sw_groupnorm_nhwc: null argument");
    return gn_launch(x, x2, y, gamma, beta, n, hw, c1, c2, groups, eps, silu, stats_scratch,
                     static_cast<cudaStream_t>(stream), nullptr);
}

extern "C" int sw_groupnorm_nhwc_zeroed(const void* x, const void* x2, void* y, const void* gamma,
                                        const void* beta, int n, int hw, int c1, int c2, int groups,
                                        float eps, int silu, float* stats_scratch, void* stream) {
    if (!x || !y || !stats_scratch) return swh::fail(-1, "sw_groupnorm_nhwc_zeroed: null argument");
    return gn_launch(x, x2, y, gamma, beta, n, hw, c1, c2, groups, eps, silu, stats_scratch,
                     static_cast<cudaStream_t>(stream), nullptr);
}

extern "C" int sw_groupnorm_nhwc_dist(const void* x, const void* x2, void* y, const void* gamma,
                                      const void* beta, int n, int hw, int c1, int c2, int groups,
                                      float eps, int silu, float* stats_scratch, const sw_peers* peers,
                                      void* stream) {
    if (!x || !y || !stats_scratch || !peers) return swh::fail(-1, "sw_groupnorm_nhwc_dist: null argument");
    return sw_groupnorm_launch(x, x2, y, gamma, beta, n, hw, c1, c2, groups, eps, silu,
                               stats_scratch, static_cast<cudaStream_t>(stream), peers);
}

// ---- the two halves of a GroupNorm as separate kernels: what a framework has to run when the statistics
// all-reduce in between is a library call (NCCL).  They exist as the like-for-like baseline of
// sw_groupnorm_nhwc_dist (parallel/collectives.py::groupnorm_nccl_native), not as a product path.
namespace {
__global__ void __launch_bounds__(256) k_gn_sums(const __half* __restrict__ x, float* __restrict__ sums, int hw, int c,
                                                 int groups, int rows_per_cta) {
    // grid (row chunks, n): per-channel-pair partial sums over this CTA's rows, folded per group, one atomicAdd
    // per (CTA, group, moment)
    extern __shared__ float sh[];  // [2][groups]
    const int n = blockIdx.y, cpg = c / groups;
    for (int i = threadIdx.x; i < 2 * groups; i += blockDim.x) sh[i] = 0.0f;
    __syncthreads();
    const int r0 = blockIdx.x * rows_per_cta, r1 = min(hw, r0 + rows_per_cta);
    const int pairs = c >> 1;
    for (int pc = threadIdx.x; pc < pairs; pc += blockDim.x) {
        float su = 0.0f, sq = 0.0f;
        const __half2* col = reinterpret_cast<const __half2*>(x + (size_t(n) * hw + r0) * c) + pc;
        for (int r = r0; r < r1; ++r, col += pairs) {
            const float2 v = __half22float2(*col);
            su += v.x + v.y;
            sq += v.x * v.x + v.y * v.y;
        }
        const int g = (2 * pc) / cpg;
        atomicAdd(&sh[g], su);
        atomicAdd(&sh[groups + g], sq);
    }
    __syncthreads();
    for (int g = threadIdx.x; g < groups; g += blockDim.x) {
        atomicAdd(&sums[(size_t(n) * groups + g) * 2], sh[g]);
        atomicAdd(&sums[(size_t(n) * groups + g) * 2 + 1], sh[groups + g]);
    }
}

__global__ void __launch_bounds__(256) k_gn_apply_sums(const __half* __restrict__ x, const float* __restrict__ sums,
                                                       const __half* __restrict__ gamma, const __half* __restrict__ beta,
                                                       __half* __restrict__ y, int hw, int c, int groups, float inv_count,
                                                       float eps, int silu, long long total_pairs) {
    const int cpg = c / groups, pairs = c >> 1;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total_pairs;
         i += (long long)gridDim.x * blockDim.x) {
        const int pc = int(i % pairs);
        const long long row = i / pairs;
        const int n = int(row / hw), g = (2 * pc) / cpg;
        const float mean = sums[(size_t(n) * groups + g) * 2] * inv_count;
        const float var = fmaxf(sums[(size_t(n) * groups + g) * 2 + 1] * inv_count - mean * mean, 0.0f);
        const float rstd = rsqrtf(var + eps);
        const float2 v = __half22float2(reinterpret_cast<const __half2*>(x)[i]);
        const float2 ga = __half22float2(reinterpret_cast<const __half2*>(gamma)[pc]);
        const float2 be = __half22float2(reinterpret_cast<const __half2*>(beta)[pc]);
        float u0 = (v.x - mean) * rstd * ga.x + be.x, u1 = (v.y - mean) * rstd * ga.y + be.y;
        if (silu) {
            u0 = silu_f(u0);
            u1 = silu_f(u1);
        }
        reinterpret_cast<__half2*>(y)[i] = __floats2half2_rn(u0, u1);
    }
}
}  // namespace

extern "C" int sw_gn_partial_sums(const void* x, float* sums, int n, int hw, int c, int groups, void* stream) {
    if (!x || !sums || c % 2 || groups <= 0 || c % groups || (c / groups) % 2)
        return swh::fail(-1, "sw_gn_partial_sums: bad arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SW_CUDA_OK(cudaMemsetAsync(sums, 0, size_t(n) * groups * 2 * sizeof(float), st));
    const int rows_per_cta = 64;
    k_gn_sums<<<dim3((hw + rows_per_cta - 1) / rows_per_cta, n), 256, 2 * groups * sizeof(float), st>>>(
        static_cast<const __half*>(x), sums, hw, c, groups, rows_per_cta);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int sw_gn_apply_sums(const void* x, const float* sums, const void* gamma, const void* beta, void* y, int n,
                                int hw, int c, int groups, long long count, float eps, int silu, void* stream) {
    if (!x || !sums || !gamma || !beta || !y || c % 2 || c % groups || (c / groups) % 2 || count <= 0)
        return swh::fail(-1, "sw_gn_apply_sums: bad arguments");
    const long long total = (long long)n * hw * (c / 2);
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_gn_apply_sums<<<int(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __half*>(x), sums, static_cast<const __half*>(gamma), static_cast<const __half*>(beta),
        static_cast<__half*>(y), hw, c, groups, 1.0f / float(count), eps, silu, total);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int sw_layernorm(const void* x, void* y, const void* gamma, const void* beta, int rows,
                            int c, float eps, void* stream) {
    if (!x || !y || !gamma || !beta) return swh::fail(-1, "sw_layernorm: null argument");
    if (c % 8 || c > 2048) return swh::fail(-1, "sw_layernorm: c=%d unsupported", c);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int threads = 256;
    const int blocks = (rows * 32 + threads - 1) / threads;
    const int octs = c / 8;
    const __half* xx = static_cast<const __half*>(x);
    __half* yy = static_cast<__half*>(y);
    const __half* g = static_cast<const __half*>(gamma);
    const __half* b = static_cast<const __half*>(beta);
    if (octs <= 64)
        SW_CUDA_OK(swh::launch_pdl(k_layernorm<2>, dim3(blocks), dim3(threads), 0, st, 1, xx, yy, g, b, rows, c, eps));
    else if (octs <= 160)
        SW_CUDA_OK(swh::launch_pdl(k_layernorm<5>, dim3(blocks), dim3(threads), 0, st, 1, xx, yy, g, b, rows, c, eps));
    else
        SW_CUDA_OK(swh::launch_pdl(k_layernorm<8>, dim3(blocks), dim3(threads), 0, st, 1, xx, yy, g, b, rows, c, eps));
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int sw_softmax_rows(const void* x, void* y, int rows, int cols, float scale, void* stream) {
    if (!x || !y) return swh::fail(-1, "sw_softmax_rows: null argument");
    if (cols % 8) return swh::fail(-1, "sw_softmax_rows: cols=%d not a multiple of 8", cols);
    k_softmax_rows<<<rows, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __half*>(x), static_cast<__half*>(y), cols, scale);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}

// Probe builds (-DSW_TRACE) only: per-CTA phase stamps of the last GroupNorm launch.
SW_TRACE_DUMP_FN(sw_trace_dump_gn)
