// Measures cycles per tcgen05.mma (kind::f16, SS operands, SWIZZLE_128B K-major) for several N,
// for single-CTA (M=128) and CTA-pair (M=256) issue, with operands resident in shared memory.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../csrc/common.cuh"
using namespace sw;

__device__ __forceinline__ void alloc1(uint32_t* slot, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void alloc2(uint32_t* slot, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void mma2(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit2(uint64_t* bar) {
    const uint16_t mask = 3;
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void csync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <bool PAIR>
__global__ void __launch_bounds__(128) probe(int n, int iters, int kadv, long long* out, int commit_every) {
    extern __shared__ uint8_t raw[];
    const uint32_t ra = smem_u32(raw);
    uint8_t* smem = raw + (((ra + 1023u) & ~1023u) - ra);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 32768);
    uint64_t* bar2 = bar + 1;
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
    uint32_t rank = 0;
    if (PAIR) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_init(bar2, 1); fence_mbar_init(); }
    fence_proxy_async_smem();
    if (threadIdx.x < 32) { if (PAIR) alloc2(slot, 256); else alloc1(slot, 256); }
    tc_fence_before();
    if (PAIR) csync(); else __syncthreads();
    tc_fence_after();
    const uint32_t tm = *slot;
    long long t0 = 0, t1 = 0;
    if (threadIdx.x == 0 && rank == 0) {
        const uint32_t idesc = umma_idesc(PAIR ? 256 : 128, n, 0, 0, 0);
        const uint64_t ad = umma_desc_sw128(smem_u32(smem), 16, 1024);
        const uint64_t bd = umma_desc_sw128(smem_u32(smem + 16384), 16, 1024);
        t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            const uint64_t off = kadv ? uint64_t((i & 3) * 2) : 0;
            if (PAIR) mma2(tm, ad + off, bd + off, idesc, i > 0);
            else umma_f16(tm, ad + off, bd + off, idesc, i > 0);
            if (commit_every && (i % commit_every) == commit_every - 1) {
                if (PAIR) commit2(bar2); else umma_commit(bar2);
                if (commit_every < 0) {}
            }
        }
        if (PAIR) commit2(bar); else umma_commit(bar);
        mbar_wait(bar, 0);
        t1 = clock64();
        out[blockIdx.x] = t1 - t0;
    } else if (PAIR && threadIdx.x == 0) {
        mbar_wait(bar, 0);
    }
    tc_fence_before();
    if (PAIR) csync(); else __syncthreads();
    if (threadIdx.x < 32) {
        if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(256));
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(256));
    }
}

int main() {
    long long* out;
    cudaMalloc(&out, 1024 * sizeof(long long));
    const int smem = 16384 + 32768 + 1024 + 64;
    cudaFuncSetAttribute(probe<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(probe<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int iters = 512;
    for (int grid : {148}) {
        for (int ce : {2, 8, 16, 64}) {
        for (int kadv : {1}) {
            for (int n : {32, 64, 128, 256}) {
                for (int pair = 0; pair < 2; ++pair) {
                    if (pair && grid == 1) continue;
                    cudaMemset(out, 0, 1024 * sizeof(long long));
                    if (!pair) {
                        probe<false><<<grid, 128, smem>>>(n, iters, kadv, out, ce);
                    } else {
                        cudaLaunchConfig_t cfg = {};
                        cfg.gridDim = dim3(grid, 1, 1);
                        cfg.blockDim = dim3(128, 1, 1);
                        cfg.dynamicSmemBytes = smem;
                        cudaLaunchAttribute at[1];
                        at[0].id = cudaLaunchAttributeClusterDimension;
                        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                        cfg.attrs = at; cfg.numAttrs = 1;
                        cudaLaunchKernelEx(&cfg, probe<true>, n, iters, kadv, out, ce);
                    }
                    cudaError_t e = cudaDeviceSynchronize();
                    long long h[2];
                    cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
                    const double cyc = double(h[0]) / iters;
                    const double macs = (pair ? 256.0 : 128.0) * n * 16;
                    printf("commit_every=%d grid=%3d kadv=%d N=%3d %s: %7.1f cycles/MMA  -> %6.0f MAC/clk/SM  (%s)\n", ce, grid, kadv, n,
                           pair ? "M=256 2CTA" : "M=128 1CTA", cyc, macs / cyc / (pair ? 2 : 1), cudaGetErrorString(e));
                }
            }
        }
        }
    }
    return 0;
}
