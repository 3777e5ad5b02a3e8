// Memory-bound glue kernels: layout converters, nearest upsample, im2col for strided convs,
// timestep embedding, fused CFG + scheduler update, uint8 image head.
#include "../include/swarm_b200.h"
#include "common.cuh"
#include "host_util.h"

namespace {

using namespace sw;

inline int blocks_for(long long work, int threads) {
    long long b = (work + threads - 1) / threads;
    if (b < 1) b = 1;
    if (b > 148LL * 32) b = 148LL * 32;  // grid-stride beyond 32 CTAs per SM
    return int(b);
}

__global__ void k_upsample2x(const uint4* __restrict__ x, uint4* __restrict__ y, int n, int h, int w,
                             int octs) {
    pdl_trigger();
    pdl_wait();
    const long long total = (long long)n * (2 * h) * (2 * w) * octs;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int o = int(i % octs);
        long long t = i / octs;
        const int ox = int(t % (2 * w));
        t /= (2 * w);
        const int oy = int(t % (2 * h));
        const int ni = int(t / (2 * h));
        y[i] = x[(((long long)ni * h + (oy >> 1)) * w + (ox >> 1)) * octs + o];
    }
}

__global__ void k_im2col(const uint4* __restrict__ x, uint4* __restrict__ out, int n, int h, int w,
                         int octs, int kh, int kw, int stride, int pad_top, int pad_left, int oh,
                         int ow) {
    pdl_trigger();
    pdl_wait();
    const long long total = (long long)n * oh * ow * kh * kw * octs;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int o = int(i % octs);
        long long t = i / octs;
        const int kx = int(t % kw);
        t /= kw;
        const int ky = int(t % kh);
        t /= kh;
        const int ox = int(t % ow);
        t /= ow;
        const int oy = int(t % oh);
        const int ni = int(t / oh);
        const int iy = oy * stride + ky - pad_top;
        const int ix = ox * stride + kx - pad_left;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (iy >= 0 && iy < h && ix >= 0 && ix < w)
            v = x[(((long long)ni * h + iy) * w + ix) * octs + o];
        out[i] = v;
    }
}

template <typename T>
__global__ void k_nchw_to_nhwc(const T* __restrict__ x, __half* __restrict__ y, int n, int c, int hw,
                               int c_pad) {
    const long long total = (long long)n * hw * c_pad;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int ch = int(i % c_pad);
        const long long t = i / c_pad;
        const int p = int(t % hw);
        const int ni = int(t / hw);
        float v = 0.0f;
        if (ch < c) v = float(x[((long long)ni * c + ch) * hw + p]);
        y[i] = __float2half_rn(v);
    }
}

template <typename T>
__global__ void k_nhwc_to_nchw(const __half* __restrict__ x, T* __restrict__ y, int n, int c, int hw,
                               int c_pad) {
    const long long total = (long long)n * c * hw;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int p = int(i % hw);
        const long long t = i / hw;
        const int ch = int(t % c);
        const int ni = int(t / c);
        y[i] = T(__half2float(x[((long long)ni * hw + p) * c_pad + ch]));
    }
}

__global__ void k_timestep_embedding(const float* __restrict__ t_table,
                                     const int* __restrict__ step_idx, __half* __restrict__ out,
                                     int n, int dim, float max_period, int flip, float freq_shift) {
    pdl_trigger();
    pdl_wait();
    const int half = dim / 2;
    const float t = t_table[step_idx ? *step_idx : 0];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n * half; i += gridDim.x * blockDim.x) {
        const int k = i % half;
        const int row = i / half;
        const float freq = __expf(-__logf(max_period) * float(k) / (float(half) - freq_shift));
        const float a = t * freq;
        const float s = sinf(a), c = cosf(a);
        __half* o = out + (size_t)row * dim;
        if (flip) {
            o[k] = __float2half_rn(c);
            o[half + k] = __float2half_rn(s);
        } else {
            o[k] = __float2half_rn(s);
            o[half + k] = __float2half_rn(c);
        }
    }
}

__global__ void k_advance(int* c) {
    pdl_trigger();
    pdl_wait(); *c += 1; }

__global__ void k_add(const uint4* __restrict__ a, const uint4* __restrict__ b, uint4* __restrict__ y,
                      long long n8) {
    pdl_trigger();
    pdl_wait();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8;
         i += (long long)gridDim.x * blockDim.x) {
        const uint4 u = a[i], v = b[i];
        const uint32_t uw[4] = {u.x, u.y, u.z, u.w}, vw[4] = {v.x, v.y, v.z, v.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 p = unpack_h2(uw[j]), q = unpack_h2(vw[j]);
            o[j] = pack_h2(p.x + q.x, p.y + q.y);
        }
        y[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

__global__ void k_silu(const uint4* __restrict__ a, uint4* __restrict__ y, long long n8) {
    pdl_trigger();
    pdl_wait();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8;
         i += (long long)gridDim.x * blockDim.x) {
        const uint4 u = a[i];
        const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 p = unpack_h2(uw[j]);
            o[j] = pack_h2(silu_f(p.x), silu_f(p.y));
        }
        y[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

__global__ void k_sched_step(const sw_sched_params p) {
    pdl_trigger();
    pdl_wait();
    const int step = p.step_idx ? *p.step_idx : 0;
    const float* cf = p.coef + (size_t)step * 8;
    const float e0 = cf[0], e1 = cf[1], c0 = cf[2], c1 = cf[3], c2 = cf[4], c3 = cf[5], s_in = cf[6];
    // guidance scales: a device slot when given (a captured graph then serves every guidance value), else scalars
    const float g_text = p.guidance ? p.guidance[0] : p.g_text;
    const float g_image = p.guidance ? p.guidance[1] : p.g_image;
    const long long total = (long long)p.rows * p.c;
    const __half* eps = static_cast<const __half*>(p.eps);
    __half* xin = static_cast<__half*>(p.x_model_in);
    const long long cond_stride = (long long)p.rows * p.eps_ld;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int ch = int(i % p.c);
        const long long row = i / p.c;
        const long long ei = row * p.eps_ld + ch;
        float e;
        if (p.n_cond == 1) {
            e = __half2float(eps[ei]);
        } else if (p.n_cond == 2) {
            const float eu = __half2float(eps[ei]);
            const float ec = __half2float(eps[cond_stride + ei]);
            e = eu + g_text * (ec - eu);
        } else {
            const float et = __half2float(eps[ei]);
            const float em = __half2float(eps[cond_stride + ei]);
            const float eu = __half2float(eps[2 * cond_stride + ei]);
            e = eu + g_text * (et - em) + g_image * (em - eu);
        }
        const float x = p.x[i];
        const float x0 = e0 * x + e1 * e;
        float xn = c0 * x + c1 * x0;
        if (c2 != 0.0f) xn += c2 * p.x0_prev[i];
        if (p.noise && c3 != 0.0f) xn += c3 * p.noise[(size_t)step * total + i];
        p.x0_prev[i] = x0;
        p.x[i] = xn;
        if (xin) {
            const __half hv = __float2half_rn(xn * s_in);
            for (int k = 0; k < p.n_cond; ++k)
                xin[((long long)k * p.rows + row) * p.xin_ld + ch] = hv;
        }
    }
}

__global__ void k_image_to_u8(const __half* __restrict__ x, uint8_t* __restrict__ out,
                              long long pixels, int c_pad) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < pixels * 3;
         i += (long long)gridDim.x * blockDim.x) {
        const long long px = i / 3;
        const int ch = int(i - px * 3);
        float v = __half2float(x[px * c_pad + ch]) * 0.5f + 0.5f;
        v = fminf(fmaxf(v, 0.0f), 1.0f);
        out[i] = (uint8_t)__float2int_rn(v * 255.0f);
    }
}

}  // namespace

extern "C" int sw_upsample2x_nhwc(const void* x, void* y, int n, int h, int w, int c, void* stream) {
    if (c % 8) return swh::fail(-1, "sw_upsample2x_nhwc: c=%d not a multiple of 8", c);
    const long long total = (long long)n * 4 * h * w * (c / 8);
    SW_CUDA_OK(swh::launch_pdl(k_upsample2x, dim3(blocks_for(total, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream),
                               1, static_cast<const uint4*>(x), static_cast<uint4*>(y), n, h, w, c / 8));
    return 0;
}

extern "C" int sw_im2col_nhwc(const void* x, void* out, int n, int h, int w, int c, int kh, int kw,
                              int stride, int pad_top, int pad_left, int oh, int ow, void* stream) {
    if (c % 8) return swh::fail(-1, "sw_im2col_nhwc: c=%d not a multiple of 8", c);
    const long long total = (long long)n * oh * ow * kh * kw * (c / 8);
    SW_CUDA_OK(swh::launch_pdl(k_im2col, dim3(blocks_for(total, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), 1,
                               static_cast<const uint4*>(x), static_cast<uint4*>(out), n, h, w, c / 8, kh, kw, stride,
                               pad_top, pad_left, oh, ow));
    return 0;
}

extern "C" int sw_nchw_to_nhwc(const void* x, void* y, int n, int c, int h, int w, int c_pad,
                               int src_fp32, void* stream) {
    const long long total = (long long)n * h * w * c_pad;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (src_fp32)
        k_nchw_to_nhwc<float><<<blocks_for(total, 256), 256, 0, st>>>(
            static_cast<const float*>(x), static_cast<__half*>(y), n, c, h * w, c_pad);
    else
        k_nchw_to_nhwc<__half><<<blocks_for(total, 256), 256, 0, st>>>(
            static_cast<const __half*>(x), static_cast<__half*>(y), n, c, h * w, c_pad);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int sw_nhwc_to_nchw(const void* x, void* y, int n, int c, int h, int w, int c_pad,
                               int dst_fp32, void* stream) {
    const long long total = (long long)n * h * w * c;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dst_fp32)
        k_nhwc_to_nchw<float><<<blocks_for(total, 256), 256, 0, st>>>(
            static_cast<const __half*>(x), static_cast<float*>(y), n, c, h * w, c_pad);
    else
        k_nhwc_to_nchw<__half><<<blocks_for(total, 256), 256, 0, st>>>(
            static_cast<const __half*>(x), static_cast<__half*>(y), n, c, h * w, c_pad);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int sw_timestep_embedding(const float* t_table, const int32_t* step_idx, void* out, int n,
                                     int dim, float max_period, int flip_sin_to_cos,
                                     float freq_shift, void* stream) {
    if (dim % 2) return swh::fail(-1, "sw_timestep_embedding: odd dim");
    SW_CUDA_OK(swh::launch_pdl(k_timestep_embedding, dim3(blocks_for((long long)n * dim / 2, 128)), dim3(128), 0,
                               static_cast<cudaStream_t>(stream), 1, t_table, static_cast<const int*>(step_idx),
                               static_cast<__half*>(out), n, dim, max_period, flip_sin_to_cos, freq_shift));
    return 0;
}

extern "C" int sw_advance_counter(int32_t* counter, void* stream) {
    SW_CUDA_OK(swh::launch_pdl(k_advance, dim3(1), dim3(1), 0, static_cast<cudaStream_t>(stream), 1,
                               static_cast<int*>(counter)));
    return 0;
}

extern "C" int sw_add(const void* a, const void* b, void* y, int64_t n, void* stream) {
    if (n % 8) return swh::fail(-1, "sw_add: n not a multiple of 8");
    SW_CUDA_OK(swh::launch_pdl(k_add, dim3(blocks_for(n / 8, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), 1,
                               static_cast<const uint4*>(a), static_cast<const uint4*>(b), static_cast<uint4*>(y),
                               (long long)(n / 8)));
    return 0;
}

extern "C" int sw_silu(const void* x, void* y, int64_t n, void* stream) {
    if (n % 8) return swh::fail(-1, "sw_silu: n not a multiple of 8");
    SW_CUDA_OK(swh::launch_pdl(k_silu, dim3(blocks_for(n / 8, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), 1,
                               static_cast<const uint4*>(x), static_cast<uint4*>(y), (long long)(n / 8)));
    return 0;
}

extern "C" int sw_sched_step(const sw_sched_params* p, void* stream) {
    if (!p || !p->eps || !p->x || !p->x0_prev || !p->coef)
        return swh::fail(-1, "sw_sched_step: null argument");
    if (p->n_cond < 1 || p->n_cond > 3) return swh::fail(-1, "sw_sched_step: n_cond=%d", p->n_cond);
    const long long total = (long long)p->rows * p->c;
    SW_CUDA_OK(swh::launch_pdl(k_sched_step, dim3(blocks_for(total, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream),
                               1, *p));
    return 0;
}

extern "C" int sw_image_to_u8(const void* x_nhwc, uint8_t* out, int64_t pixels, int c_pad,
                              void* stream) {
    k_image_to_u8<<<blocks_for(pixels * 3, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __half*>(x_nhwc), out, pixels, c_pad);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}
