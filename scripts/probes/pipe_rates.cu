// Per-SM instruction throughput of the pipes the softmax / SiLU / GELU epilogues lean on (B200, sm_100a):
// MUFU (ex2, rcp, tanh), FFMA (register and immediate forms), FMNMX, F2FP pack, HFMA2.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rates pipe_rates.cu && ./pipe_rates
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;
constexpr int UNROLL = 8;

template <int OP>
__global__ void k(float* out, float seed, long long* clk) {
    float v[UNROLL];
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) v[i] = seed + threadIdx.x * 1e-3f + i;
    float c1 = seed * 0.5f, c2 = seed * 0.25f;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < UNROLL; ++i) {
            if (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
            if (OP == 1) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
            if (OP == 2) asm volatile("tanh.approx.f32 %0, %0;" : "+f"(v[i]));
            if (OP == 3) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(v[i]) : "f"(c1), "f"(c2));
            if (OP == 4) asm volatile("fma.rn.f32 %0, %0, 0f3F800001, 0f3A83126F;" : "+f"(v[i]));
            if (OP == 5) asm volatile("max.f32 %0, %0, %1;" : "+f"(v[i]) : "f"(c1));
            if (OP == 6) {
                unsigned h;
                asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h) : "f"(v[i]), "f"(c1));
                v[i] = __uint_as_float(h);
            }
            if (OP == 7) {
                unsigned h = __float_as_uint(v[i]);
                asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(h) : "r"(__float_as_uint(c1)), "r"(__float_as_uint(c2)));
                v[i] = __uint_as_float(h);
            }
            if (OP == 8) asm volatile("add.f32 %0, %0, %1;" : "+f"(v[i]) : "f"(c1));
            if (OP == 9) asm volatile("mul.f32 %0, %0, %1;" : "+f"(v[i]) : "f"(c1));
            if (OP == 10) {
                unsigned h = __float_as_uint(v[i]);
                asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h));
                v[i] = __uint_as_float(h);
            }
            if (OP == 12) {  // per element of the attention softmax: scale-subtract, exp2, row sum (+ half a pack)
                float x;
                asm volatile("fma.rn.ftz.f32 %0, %1, %2, %3;" : "=f"(x) : "f"(v[i]), "f"(c1), "f"(c2));
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x));
                asm volatile("add.ftz.f32 %0, %0, %1;" : "+f"(v[i]) : "f"(x));
            }
            if (OP == 13) {  // 2^t without the MUFU: round-to-nearest split + cubic + exponent insertion
                float t = fminf(fmaxf(v[i], -126.0f), 0.0f);
                const float r = t + 12582912.0f;
                const float f = t - (r - 12582912.0f);
                float p = fmaf(f, 0.0550093f, 0.24221097f);
                p = fmaf(p, f, 0.69328285f);
                p = fmaf(p, f, 1.0f);
                v[i] = __int_as_float(__float_as_int(p) + (__float_as_int(r) << 23)) - 1.5f;
            }
            if (OP == 11) {
                unsigned h = __float_as_uint(v[i]);
                asm volatile("tanh.approx.f16x2 %0, %0;" : "+r"(h));
                v[i] = __uint_as_float(h);
            }
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, float* out, long long* clk, int threads = 1024) {
    const int blocks = 148;
    k<OP><<<blocks, threads>>>(out, 0.5f, clk);
    k<OP><<<blocks, threads>>>(out, 0.5f, clk);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < blocks; ++i) avg += h[i];
    avg /= blocks;
    const double ops = double(threads) * ITERS * UNROLL;
    printf("%-28s %4d thr/SM  %7.2f thread-ops / clk / SM   (%.0f clk)\n", name, threads, ops / avg, avg);
}

int main() {
    float* out;
    long long* clk;
    cudaMalloc(&out, 148 * 1024 * sizeof(float));
    cudaMalloc(&clk, 148 * sizeof(long long));
    run<0>("MUFU ex2.approx.ftz.f32", out, clk);
    run<0>("MUFU ex2 (1 warp / scheduler)", out, clk, 128);
    run<0>("MUFU ex2 (2 warps / scheduler)", out, clk, 256);
    run<0>("MUFU ex2 (4 warps / scheduler)", out, clk, 512);
    run<12>("softmax mix (1 warp / sched.)", out, clk, 128);
    run<12>("softmax mix (2 warps / sched.)", out, clk, 256);
    run<12>("softmax mix (4 warps / sched.)", out, clk, 512);
    run<13>("exp2 cubic on FMA pipe (1 w/s)", out, clk, 128);
    run<13>("exp2 cubic on FMA pipe (2 w/s)", out, clk, 256);
    run<13>("exp2 cubic on FMA pipe (4 w/s)", out, clk, 512);
    run<1>("MUFU rcp.approx.ftz.f32", out, clk);
    run<2>("MUFU tanh.approx.f32", out, clk);
    run<10>("ex2.approx.f16x2 (x2 elems)", out, clk);
    run<11>("tanh.approx.f16x2 (x2 elems)", out, clk);
    run<3>("FFMA reg,reg,reg", out, clk);
    run<4>("FFMA reg,imm,imm", out, clk);
    run<8>("FADD", out, clk);
    run<9>("FMUL", out, clk);
    run<5>("FMNMX", out, clk);
    run<6>("F2FP.F16.F32.PACK_AB", out, clk);
    run<7>("HFMA2 (x2 elems)", out, clk);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
