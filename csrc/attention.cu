// Flash attention (non-causal) for sm_100a: both GEMMs run on tcgen05 with TMEM accumulators.
//
// One CTA owns 128 query rows of one (batch, head).  Per 128-key tile:
//   warp 1 (one lane) issues  S = Q K^T  (TMEM cols [0,128)), then  O += P V  (TMEM cols [128,..));
//   warps 2..5 (one thread per query row) read S with tcgen05.ld, keep the running max / sum in
//   registers, write P = exp2(S*scale - m) as fp16 into a SWIZZLE_128B smem tile that the second
//   MMA consumes as its A operand, and only touch O in TMEM when the running max moved by more
//   than 2^8 (lazy rescale);
//   warp 0 (one lane) streams K/V tiles with TMA.  Head dims that are not multiples of 64
//   (SD1.5: 40, 80, 160) rely on TMA zero-fill: the box is always 64 wide.
// With head_dim <= 64 two CTAs fit per SM, so one CTA's softmax overlaps the other's MMAs.
#include "../include/swarm_b200.h"
#include "common.cuh"
#include "host_util.h"

namespace {

using namespace sw;

constexpr int BQ = 128;
constexpr int BKV = 128;
constexpr int SLAB_BYTES = 128 * 128;  // 128 rows x 64 fp16
constexpr int kThreads = 192;
constexpr int O_COL0 = 128;

struct AttnArgs {
    int batch, heads, seq_q, seq_kv, head_dim;
    int nslab, dpad, kv_stages, tmem_cols;
    float scale_log2;
    __half* out;
    long long o_stride_b, o_stride_s;
    uint32_t idesc_qk, idesc_pv;
};

__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ void tmem_alloc_rt(uint32_t* slot, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(slot)),
                 "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_rt(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols)
                 : "memory");
}

__global__ void __launch_bounds__(kThreads)
    k_flash_attn(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const AttnArgs p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int nslab = p.nslab;
    const int stages = p.kv_stages;
    const int tile_bytes = nslab * SLAB_BYTES;

    uint8_t* sQ = smem;
    uint8_t* sK = sQ + tile_bytes;
    uint8_t* sV = sK + stages * tile_bytes;
    uint8_t* sP = sV + stages * tile_bytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * SLAB_BYTES);
    uint64_t* q_full = bars;
    uint64_t* s_full = bars + 1;
    uint64_t* p_full = bars + 2;
    uint64_t* pv_done = bars + 3;
    uint64_t* k_full = bars + 4;
    uint64_t* v_full = k_full + stages;
    uint64_t* kv_empty = v_full + stages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(kv_empty + stages);

    const int q0 = blockIdx.x * BQ;
    const int head = blockIdx.y;
    const int b = blockIdx.z;
    const int nkv = (p.seq_kv + BKV - 1) / BKV;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
    }
    if (warp == 1) {
        if (lane == 0) {
            mbar_init(q_full, 1);
            mbar_init(s_full, 1);
            mbar_init(p_full, 128);
            mbar_init(pv_done, 1);
            for (int s = 0; s < stages; ++s) {
                mbar_init(&k_full[s], 1);
                mbar_init(&v_full[s], 1);
                mbar_init(&kv_empty[s], 1);
            }
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc_rt(tmem_slot, uint32_t(p.tmem_cols));
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(q_full, uint32_t(tile_bytes));
            for (int sl = 0; sl < nslab; ++sl)
                tma_load_4d(sQ + sl * SLAB_BYTES, &tmQ, q_full, sl * 64, head, q0, b);
            int s = 0;
            uint32_t ph = 0;
            for (int j = 0; j < nkv; ++j) {
                mbar_wait(&kv_empty[s], ph ^ 1);
                mbar_expect_tx(&k_full[s], uint32_t(tile_bytes));
                for (int sl = 0; sl < nslab; ++sl)
                    tma_load_4d(sK + s * tile_bytes + sl * SLAB_BYTES, &tmK, &k_full[s], sl * 64,
                                head, j * BKV, b);
                mbar_expect_tx(&v_full[s], uint32_t(tile_bytes));
                for (int sl = 0; sl < nslab; ++sl)
                    tma_load_4d(sV + s * tile_bytes + sl * SLAB_BYTES, &tmV, &v_full[s], sl * 64,
                                head, j * BKV, b);
                if (++s == stages) {
                    s = 0;
                    ph ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const int kq = (p.head_dim + 15) / 16;
            const uint32_t tS = tmem_base;
            const uint32_t tO = tmem_base + O_COL0;
            auto issue_qk = [&](int s) {
                const uint32_t qa = smem_u32(sQ);
                const uint32_t ka = smem_u32(sK + s * tile_bytes);
                for (int kk = 0; kk < kq; ++kk) {
                    const uint32_t off = uint32_t(kk >> 2) * SLAB_BYTES;
                    const uint64_t ad = umma_desc_sw128(qa + off, 16, 1024) + uint64_t((kk & 3) * 2);
                    const uint64_t bd = umma_desc_sw128(ka + off, 16, 1024) + uint64_t((kk & 3) * 2);
                    umma_f16(tS, ad, bd, p.idesc_qk, kk > 0 ? 1u : 0u);
                }
            };
            mbar_wait(q_full, 0);
            mbar_wait(&k_full[0], 0);
            tc_fence_after();
            issue_qk(0);
            umma_commit(s_full);
            int s = 0;
            uint32_t ph = 0;
            for (int j = 0; j < nkv; ++j) {
                int s1 = s + 1;
                uint32_t ph1 = ph;
                if (s1 == stages) {
                    s1 = 0;
                    ph1 ^= 1;
                }
                mbar_wait(p_full, uint32_t(j & 1));  // softmax j has drained S and published P
                tc_fence_after();
                // With two K/V stages the next S tile is issued ahead of P.V so the softmax warps
                // can start on it while P.V runs; with one stage K[j+1] only lands after P.V[j]
                // released the buffer, so the order is reversed (head_dim 160: <= 2 KV tiles).
                if (stages > 1 && j + 1 < nkv) {
                    mbar_wait(&k_full[s1], ph1);
                    tc_fence_after();
                    issue_qk(s1);
                    umma_commit(s_full);
                }
                mbar_wait(&v_full[s], ph);
                tc_fence_after();
                const int kv_valid = min(BKV, p.seq_kv - j * BKV);
                const int kp = (kv_valid + 15) / 16;
                const uint32_t pa = smem_u32(sP);
                const uint32_t va = smem_u32(sV + s * tile_bytes);
                for (int kk = 0; kk < kp; ++kk) {
                    const uint64_t ad = umma_desc_sw128(pa + uint32_t(kk >> 2) * SLAB_BYTES, 16, 1024) +
                                        uint64_t((kk & 3) * 2);
                    // V is MN-major: 16 kv rows (K) per step = 2048 bytes; slabs of 64 dims are
                    // SLAB_BYTES apart (LBO), 8-row groups 1024 bytes apart (SBO).
                    const uint64_t bd = umma_desc_sw128(va + uint32_t(kk) * 2048, SLAB_BYTES, 1024);
                    umma_f16(tO, ad, bd, p.idesc_pv, (j > 0 || kk > 0) ? 1u : 0u);
                }
                umma_commit(&kv_empty[s]);
                umma_commit(pv_done);
                if (stages == 1 && j + 1 < nkv) {
                    mbar_wait(&k_full[s1], ph1);
                    tc_fence_after();
                    issue_qk(s1);
                    umma_commit(s_full);
                }
                s = s1;
                ph = ph1;
            }
        }
    } else {
        // ---------------- softmax + output: thread <-> query row
        const int q = warp & 3;
        const int r = q * 32 + lane;
        const uint32_t lane_off = uint32_t(q * 32) << 16;
        const uint32_t tS = tmem_base + lane_off;
        const uint32_t tO = tmem_base + lane_off + O_COL0;
        uint8_t* prow = sP + (r >> 3) * 1024 + (r & 7) * 128;
        const int rsw = r & 7;

        float m_ref = -INFINITY;
        float l = 0.0f;
        for (int j = 0; j < nkv; ++j) {
            const int kv_valid = min(BKV, p.seq_kv - j * BKV);
            mbar_wait(s_full, uint32_t(j & 1));
            tc_fence_after();
            // pass 1: row max of the raw scores
            float mx = -INFINITY;
#pragma unroll 1
            for (int c = 0; c < BKV; c += 32) {
                if (c >= kv_valid) break;  // warp-uniform
                uint32_t v[32];
                tmem_ld32(tS + c, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    if (c + i < kv_valid) mx = fmaxf(mx, __uint_as_float(v[i]));
            }
            const float m_new = fmaxf(m_ref, mx * p.scale_log2);
            if (j > 0) {
                mbar_wait(pv_done, uint32_t((j - 1) & 1));  // O is stable, P tile is free again
                tc_fence_after();
                const bool need = m_new > m_ref + 8.0f;
                if (__any_sync(0xffffffffu, need)) {
                    const float alpha = need ? ex2(m_ref - m_new) : 1.0f;
#pragma unroll 1
                    for (int c = 0; c < p.dpad; c += 16) {
                        uint32_t o[16];
                        tmem_ld16(tO + c, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        tmem_st16(tO + c, o);
                    }
                    tmem_st_wait();
                    l *= alpha;
                    if (need) m_ref = m_new;
                }
            } else {
                m_ref = m_new;
            }
            // pass 2: P = exp2(S*scale - m_ref) -> fp16 -> swizzled smem; row sum
            float lsum = 0.0f;
#pragma unroll 1
            for (int c = 0; c < BKV; c += 32) {
                uint32_t pk[16];
                if (c < kv_valid) {
                    uint32_t v[32];
                    tmem_ld32(tS + c, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        float e0 = (c + i < kv_valid)
                                       ? ex2(fmaf(__uint_as_float(v[i]), p.scale_log2, -m_ref))
                                       : 0.0f;
                        float e1 = (c + i + 1 < kv_valid)
                                       ? ex2(fmaf(__uint_as_float(v[i + 1]), p.scale_log2, -m_ref))
                                       : 0.0f;
                        lsum += e0 + e1;
                        pk[i >> 1] = pack_h2(e0, e1);
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) pk[i] = 0u;
                }
                uint8_t* slab = prow + (c >> 6) * SLAB_BYTES;
                const int cc0 = (c & 63) >> 3;  // first 16-byte chunk of this 32-column group
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    *reinterpret_cast<uint4*>(slab + (((cc0 + g) ^ rsw) << 4)) =
                        make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
                }
            }
            l += lsum;
            fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the MMA (async proxy)
            tc_fence_before();
            mbar_arrive(p_full);
        }
        mbar_wait(pv_done, uint32_t((nkv - 1) & 1));
        tc_fence_after();
        const float inv_l = 1.0f / l;
        const int qi = q0 + r;
        __half* orow = p.out + (long long)b * p.o_stride_b + (long long)qi * p.o_stride_s +
                       (long long)head * p.head_dim;
#pragma unroll 1
        for (int c = 0; c < p.dpad; c += 16) {
            uint32_t o[16];
            tmem_ld16(tO + c, o);
            tmem_ld_wait();
            if (qi < p.seq_q) {
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    if (c + g * 8 < p.head_dim) {
                        const int i = g * 8;
                        *reinterpret_cast<uint4*>(orow + c + i) = make_uint4(
                            pack_h2(__uint_as_float(o[i]) * inv_l, __uint_as_float(o[i + 1]) * inv_l),
                            pack_h2(__uint_as_float(o[i + 2]) * inv_l, __uint_as_float(o[i + 3]) * inv_l),
                            pack_h2(__uint_as_float(o[i + 4]) * inv_l, __uint_as_float(o[i + 5]) * inv_l),
                            pack_h2(__uint_as_float(o[i + 6]) * inv_l, __uint_as_float(o[i + 7]) * inv_l));
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_rt(tmem_base, uint32_t(p.tmem_cols));
    }
}

int make_qkv_map(CUtensorMap* m, const void* base, int d, int heads, int seq, int batch,
                 long long sb, long long ss, long long sh) {
    const uint64_t dims[4] = {uint64_t(d), uint64_t(heads), uint64_t(seq), uint64_t(batch)};
    const uint64_t str[3] = {uint64_t(sh) * 2, uint64_t(ss) * 2, uint64_t(sb) * 2};
    const uint32_t box[4] = {64, 1, 128, 1};
    return swh::encode_tmap_f16(m, base, 4, dims, str, box);
}

}  // namespace

extern "C" int sw_flash_attn(const sw_attn_params* p, void* stream) {
    if (!p || !p->q || !p->k || !p->v || !p->out) return swh::fail(-1, "sw_flash_attn: null argument");
    if (p->head_dim <= 0 || p->head_dim % 8 || p->head_dim > 160)
        return swh::fail(-1, "sw_flash_attn: head_dim=%d unsupported (multiple of 8, <= 160)", p->head_dim);
    if (p->batch <= 0 || p->heads <= 0 || p->seq_q <= 0 || p->seq_kv <= 0)
        return swh::fail(-1, "sw_flash_attn: bad shape");
    if ((p->o_stride_s % 8) || (p->o_stride_b % 8))
        return swh::fail(-1, "sw_flash_attn: output strides must be multiples of 8 elements");

    AttnArgs a;
    memset(&a, 0, sizeof(a));
    a.batch = p->batch;
    a.heads = p->heads;
    a.seq_q = p->seq_q;
    a.seq_kv = p->seq_kv;
    a.head_dim = p->head_dim;
    a.nslab = (p->head_dim + 63) / 64;
    a.dpad = (p->head_dim + 15) / 16 * 16;
    const int nkv = (p->seq_kv + BKV - 1) / BKV;
    a.kv_stages = (a.nslab <= 2 && nkv > 1) ? 2 : 1;
    a.tmem_cols = (O_COL0 + a.dpad <= 256) ? 256 : 512;
    a.scale_log2 = p->scale * 1.4426950408889634f;
    a.out = static_cast<__half*>(p->out);
    a.o_stride_b = p->o_stride_b;
    a.o_stride_s = p->o_stride_s;
    a.idesc_qk = sw::umma_idesc(128, 128, 0, 0, 0);
    a.idesc_pv = sw::umma_idesc(128, a.dpad, 0, 0, 1);

    CUtensorMap tq, tk, tv;
    int rc = make_qkv_map(&tq, p->q, p->head_dim, p->heads, p->seq_q, p->batch, p->q_stride_b,
                          p->q_stride_s, p->q_stride_h);
    if (rc) return rc;
    rc = make_qkv_map(&tk, p->k, p->head_dim, p->heads, p->seq_kv, p->batch, p->k_stride_b,
                      p->k_stride_s, p->k_stride_h);
    if (rc) return rc;
    rc = make_qkv_map(&tv, p->v, p->head_dim, p->heads, p->seq_kv, p->batch, p->v_stride_b,
                      p->v_stride_s, p->v_stride_h);
    if (rc) return rc;

    const int tile_bytes = a.nslab * SLAB_BYTES;
    const int smem = tile_bytes * (1 + 2 * a.kv_stages) + 2 * SLAB_BYTES + 1024 + 256;
    static bool configured[64] = {false};
    int dev = 0;
    SW_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 64 && !configured[dev]) {
        SW_CUDA_OK(cudaFuncSetAttribute(k_flash_attn, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        227 * 1024));
        configured[dev] = true;
    }
    if (smem > 227 * 1024) return swh::fail(-1, "sw_flash_attn: shared memory budget exceeded (%d)", smem);
    dim3 grid((p->seq_q + BQ - 1) / BQ, p->heads, p->batch);
    k_flash_attn<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(tq, tk, tv, a);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}
