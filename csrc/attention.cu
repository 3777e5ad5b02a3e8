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

#include <stdlib.h>

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
    int k_stages, v_stages;  // v2 kernel: independent K and V rings
    // cross-GPU K/V gather (sequence parallel): K/V rows [src*s_local, (src+1)*s_local) live on rank src
    int world, rank, epoch, s_local;
    const int* flags_local;  // this rank's int32[world] epoch flags
    const int* epoch_base;   // optional device counter: epoch += *epoch_base * epoch_mul (graph replays)
    int epoch_mul;
    float scale_log2;
    __half* out;
    long long o_stride_b, o_stride_s;
    uint32_t idesc_qk, idesc_pv;
    // v1, short sequences (temporal attention: seq = frames): `pack` batch entries share one 128-row
    // tile (TMA box {64, 1, seq, pack}); the softmax masks everything outside a row's own entry
    int pack;
    int causal;  // v1: row q0 + r sees columns <= its own index
    // v2: clocks the second query tile's softmax warpgroup holds back before its first tile, so that the two
    // warpgroups sit half a period apart (one in the exponentials while the other loads / reduces / waits)
    int stagger;
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

    // (through a shuffle so that ptxas knows the role branches and everything inside them are warp-uniform:
    // the TMA / MMA loops then run on the uniform datapath instead of R2UR.BROADCAST waterfalls)
    const int warp = __shfl_sync(0xffffffffu, int(threadIdx.x >> 5), 0);
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

    const int pack = p.pack;  // > 1: blockIdx.z walks groups of `pack` batch entries
    const int q0 = blockIdx.x * BQ;
    const int head = blockIdx.y;
    const int b = blockIdx.z * pack;
    const int nkv = (p.seq_kv + BKV - 1) / BKV;
    // bytes one TMA box delivers per 64-wide slab (a packed box has pack * seq rows, not 128)
    const uint32_t q_tx = uint32_t(nslab) * uint32_t(pack > 1 ? pack * p.seq_q : BQ) * 128u;
    const uint32_t kv_tx = uint32_t(nslab) * uint32_t(pack > 1 ? pack * p.seq_kv : BKV) * 128u;

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
    if (pack > 1) {
        // rows past pack * seq of the operand tiles are never written by TMA: clear them once so the
        // P.V product cannot pick up stale NaNs (P is zero there, but 0 * NaN is not)
        for (int i = threadIdx.x; i < (1 + 2 * stages) * tile_bytes / 16; i += blockDim.x)
            reinterpret_cast<uint4*>(sQ)[i] = make_uint4(0, 0, 0, 0);
        fence_proxy_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    pdl_trigger();
    pdl_wait();

    if (warp == 0) {
        {
            // whole warp walks the loop, one elected lane issues
            if (elect_one()) {
                mbar_expect_tx(q_full, q_tx);
                for (int sl = 0; sl < nslab; ++sl)
                    tma_load_4d(sQ + sl * SLAB_BYTES, &tmQ, q_full, sl * 64, head, q0, b);
            }
            int s = 0;
            uint32_t ph = 0;
            for (int j = 0; j < nkv; ++j) {
                mbar_wait(&kv_empty[s], ph ^ 1);
                if (elect_one()) {
                    mbar_expect_tx(&k_full[s], kv_tx);
                    for (int sl = 0; sl < nslab; ++sl)
                        tma_load_4d(sK + s * tile_bytes + sl * SLAB_BYTES, &tmK, &k_full[s], sl * 64,
                                    head, j * BKV, b);
                    mbar_expect_tx(&v_full[s], kv_tx);
                    for (int sl = 0; sl < nslab; ++sl)
                        tma_load_4d(sV + s * tile_bytes + sl * SLAB_BYTES, &tmV, &v_full[s], sl * 64,
                                    head, j * BKV, b);
                }
                if (++s == stages) {
                    s = 0;
                    ph ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        {
            // Lean issue loop (whole warp, one elected lane -- always the same one -- issues): it is on the
            // critical path of every KV tile, so all descriptor halves are precomputed and the K loops are
            // unrolled with predicates.
            const int kq = (p.head_dim + 15) / 16;
            const uint32_t tS = tmem_base;
            const uint32_t tO = tmem_base + O_COL0;
            const uint32_t q_lo = desc_lo(smem_u32(sQ), 16);
            const uint32_t k_lo0 = desc_lo(smem_u32(sK), 16);
            const uint32_t p_lo = desc_lo(smem_u32(sP), 16);
            // V is MN-major: 16 kv rows (K) per step = 2048 bytes; slabs of 64 dims are SLAB_BYTES
            // apart (LBO), 8-row groups 1024 bytes apart (SBO).
            const uint32_t v_lo0 = desc_lo(smem_u32(sV), SLAB_BYTES);
            const uint32_t stage_step = uint32_t(tile_bytes) >> 4;
            const uint32_t idesc_qk = p.idesc_qk, idesc_pv = p.idesc_pv;
            auto issue_qk = [&](uint32_t k_lo) {
#pragma unroll
                for (int kk = 0; kk < 10; ++kk) {
                    if (kk < kq) {
                        const uint32_t off = uint32_t(kk >> 2) * (SLAB_BYTES >> 4) + uint32_t(kk & 3) * 2;
                        umma_f16_lohi(tS, q_lo + off, k_lo + off, kDescHiSw128, idesc_qk, kk > 0 ? 1u : 0u);
                    }
                }
            };
            mbar_wait(q_full, 0);
            mbar_wait(&k_full[0], 0);
            tc_fence_after();
            if (elect_one()) {
                issue_qk(k_lo0);
                umma_commit(s_full);
            }
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
                    if (elect_one()) {
                        issue_qk(k_lo0 + uint32_t(s1) * stage_step);
                        umma_commit(s_full);
                    }
                }
                mbar_wait(&v_full[s], ph);
                tc_fence_after();
                const int kv_valid = pack > 1 ? pack * p.seq_kv : min(BKV, p.seq_kv - j * BKV);
                const int kp = (kv_valid + 15) >> 4;
                const uint32_t v_lo = v_lo0 + uint32_t(s) * stage_step;
                const uint32_t acc0 = j > 0 ? 1u : 0u;
                if (elect_one()) {
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk) {
                        if (kk < kp) {
                            const uint32_t poff = uint32_t(kk >> 2) * (SLAB_BYTES >> 4) + uint32_t(kk & 3) * 2;
                            umma_f16_lohi(tO, p_lo + poff, v_lo + uint32_t(kk) * (2048 >> 4), kDescHiSw128, idesc_pv,
                                          kk > 0 ? 1u : acc0);
                        }
                    }
                    umma_commit(&kv_empty[s]);
                    umma_commit(pv_done);
                }
                if (stages == 1 && j + 1 < nkv) {
                    mbar_wait(&k_full[s1], ph1);
                    tc_fence_after();
                    if (elect_one()) {
                        issue_qk(k_lo0 + uint32_t(s1) * stage_step);
                        umma_commit(s_full);
                    }
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

        // columns of S this row may attend to: the whole tile, or (packed) its own batch entry's block
        const int grp = pack > 1 ? r / p.seq_q : 0;
        const int col_lo = pack > 1 ? grp * p.seq_kv : 0;
        float m_ref = -INFINITY;
        float l = 0.0f;
        for (int j = 0; j < nkv; ++j) {
            const int kv_valid = pack > 1 ? pack * p.seq_kv : min(BKV, p.seq_kv - j * BKV);
            int col_hi = pack > 1 ? col_lo + p.seq_kv : kv_valid;
            if (p.causal) col_hi = min(col_hi, q0 + r - j * BKV + 1);  // (an empty range leaves m_ref and l untouched)
            if (lane == 0) mbar_wait(s_full, uint32_t(j & 1));
            __syncwarp();
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
                    if (c + i >= col_lo && c + i < col_hi) mx = fmaxf(mx, __uint_as_float(v[i]));
            }
            const float m_new = fmaxf(m_ref, mx * p.scale_log2);
            if (j > 0) {
                if (lane == 0) mbar_wait(pv_done, uint32_t((j - 1) & 1));  // O is stable, P tile is free again
                __syncwarp();
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
                        float e0 = (c + i >= col_lo && c + i < col_hi)
                                       ? ex2(fmaf(__uint_as_float(v[i]), p.scale_log2, -m_ref))
                                       : 0.0f;
                        float e1 = (c + i + 1 >= col_lo && c + i + 1 < col_hi)
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
        if (lane == 0) mbar_wait(pv_done, uint32_t((nkv - 1) & 1));
        __syncwarp();
        tc_fence_after();
        const float inv_l = 1.0f / l;
        int qi = q0 + r, bo = b;
        if (pack > 1) {  // row r = (batch entry grp, position r - grp * seq)
            qi = (grp < pack && b + grp < p.batch) ? r - grp * p.seq_q : p.seq_q;  // seq_q = "do not store"
            bo = b + grp;
        }
        __half* orow = p.out + (long long)bo * p.o_stride_b + (long long)qi * p.o_stride_s +
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


// =================================================================================================
// v2: two query tiles per CTA, ping-pong (head_dim <= 128).
//
// The v1 kernel above serialises, per KV tile, MMA -> softmax -> MMA inside one CTA and relies on a
// second resident CTA for overlap; measured 4.5x slower than it should be at S = 4096.  Here one
// CTA owns 256 query rows as two 128-row tiles with their own S/O accumulators in TMEM
// (S0 S1 O0 O1 = 4 x 128 columns) and their own softmax warpgroup, so while warpgroup 0 turns S0
// into P0 the tensor core is busy with tile 1's GEMMs and vice versa.  The softmax is single-pass:
// P = exp2(S*scale - m_ref) is computed against the running reference max of earlier tiles while
// the tile max is tracked on the side; only if some row exceeds the reference by more than 2^8
// (rare after the first tiles) is O rescaled and the tile recomputed.
// Warps: 0 TMA, 1 MMA issuer + TMEM owner, 2-5 softmax of tile 0, 6-9 softmax of tile 1.
constexpr int kThreads2 = 320;

struct KvMaps {
    CUtensorMap k[8];
    CUtensorMap v[8];
};

__device__ __forceinline__ void attn_wait_flag_ge(const int* p, int v) {
    int cur;
    for (unsigned i = 0;; ++i) {
        asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(cur) : "l"(p) : "memory");
        if (cur >= v) break;
        if (i > (1u << 27)) __trap();
        __nanosleep(64);
    }
}

// tmKs / tmVs: one tensor map (local attention) or one per rank (K/V gathered in-kernel: the tile
// loop starts at this rank's own rows and walks the peers' shards through their NVLink mappings).
__device__ __forceinline__ void attn2_body(const CUtensorMap& tmQ, const CUtensorMap* tmKs,
                                           const CUtensorMap* tmVs, const AttnArgs& p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

    // (through a shuffle so that ptxas knows the role branches and everything inside them are warp-uniform:
    // the TMA / MMA loops then run on the uniform datapath instead of R2UR.BROADCAST waterfalls)
    const int warp = __shfl_sync(0xffffffffu, int(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int nslab = p.nslab;
    const int kst = p.k_stages, vst = p.v_stages;
    const int tile_bytes = nslab * SLAB_BYTES;

    uint8_t* sQ = smem;                        // [2][tile_bytes]
    uint8_t* sK = sQ + 2 * tile_bytes;         // [kst][tile_bytes]
    uint8_t* sV = sK + kst * tile_bytes;       // [vst][tile_bytes]
    uint8_t* sP = sV + vst * tile_bytes;       // [2][2 * SLAB_BYTES]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 4 * SLAB_BYTES);
    uint64_t* q_full = bars;        // 1
    uint64_t* s_full = bars + 1;    // 2
    uint64_t* p_full = bars + 3;    // 2 (128 arrivals each)
    uint64_t* pv_done = bars + 5;   // 2
    uint64_t* s_free = bars + 7;    // 2 (128 arrivals each): S has been copied to registers
    uint64_t* k_full = bars + 9;    // kst
    uint64_t* k_empty = k_full + kst;
    uint64_t* v_full = k_empty + kst;
    uint64_t* v_empty = v_full + vst;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(v_empty + vst);

    const int q0 = blockIdx.x * (2 * BQ);
    const int head = blockIdx.y;
    const int b = blockIdx.z;
    const int nkv = (p.seq_kv + BKV - 1) / BKV;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(tmKs);
        tma_prefetch_desc(tmVs);
    }
    if (warp == 1) {
        if (lane == 0) {
            mbar_init(q_full, 1);
            for (int t = 0; t < 2; ++t) {
                mbar_init(&s_full[t], 1);
                mbar_init(&p_full[t], 128);
                mbar_init(&pv_done[t], 1);
                mbar_init(&s_free[t], 128);
            }
            for (int s = 0; s < kst; ++s) {
                mbar_init(&k_full[s], 1);
                mbar_init(&k_empty[s], 1);
            }
            for (int s = 0; s < vst; ++s) {
                mbar_init(&v_full[s], 1);
                mbar_init(&v_empty[s], 1);
            }
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc_rt(tmem_slot, 512u);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    pdl_trigger();
    pdl_wait();

    if (warp == 0) {
        {
            // ---------------- TMA producer (whole warp walks the loop, one elected lane issues)
            if (elect_one()) {
                mbar_expect_tx(q_full, uint32_t(2 * tile_bytes));
                for (int t = 0; t < 2; ++t)
                    for (int sl = 0; sl < nslab; ++sl)
                        tma_load_4d(sQ + t * tile_bytes + sl * SLAB_BYTES, &tmQ, q_full, sl * 64, head, q0 + t * BQ, b);
            }
            int sk = 0, sv = 0;
            uint32_t phk = 1, phv = 1;
            const int per_rank = p.world > 1 ? p.s_local / BKV : nkv;
            int jj = p.world > 1 ? p.rank * per_rank : 0;  // local shard first
            int last_src = p.world > 1 ? p.rank : 0;
            const int epoch = p.epoch + (p.epoch_base ? __shfl_sync(0xffffffffu, __ldcg(p.epoch_base), 0) * p.epoch_mul : 0);
            for (int j = 0; j < nkv; ++j) {
                int src = 0, row = jj * BKV;
                if (p.world > 1) {
                    src = jj / per_rank;
                    row = (jj - src * per_rank) * BKV;
                    if (src != last_src) {  // first touch of a peer's shard: wait until it is published
                        attn_wait_flag_ge(p.flags_local + src, epoch);
                        // the flag was read through the generic proxy, the shard is about to be read by TMA (async
                        // proxy): order the two
                        asm volatile("fence.proxy.async;" ::: "memory");
                        last_src = src;
                    }
                }
                const CUtensorMap* mk = tmKs + src;
                const CUtensorMap* mv = tmVs + src;
                mbar_wait(&k_empty[sk], phk);
                if (elect_one()) {
                    mbar_expect_tx(&k_full[sk], uint32_t(tile_bytes));
                    for (int sl = 0; sl < nslab; ++sl)
                        tma_load_4d(sK + sk * tile_bytes + sl * SLAB_BYTES, mk, &k_full[sk], sl * 64, head, row, b);
                }
                if (++sk == kst) {
                    sk = 0;
                    phk ^= 1;
                }
                mbar_wait(&v_empty[sv], phv);
                if (elect_one()) {
                    mbar_expect_tx(&v_full[sv], uint32_t(tile_bytes));
                    for (int sl = 0; sl < nslab; ++sl)
                        tma_load_4d(sV + sv * tile_bytes + sl * SLAB_BYTES, mv, &v_full[sv], sl * 64, head, row, b);
                }
                if (++sv == vst) {
                    sv = 0;
                    phv ^= 1;
                }
                if (++jj == nkv) jj = 0;
            }
        }
    } else if (warp == 1) {
        {
            // ---------------- MMA issuer (lean: precomputed descriptor halves, unrolled K loops; whole warp walks
            // the loop and one elected lane -- always the same one -- issues the MMAs and their commits)
            const int kq = (p.head_dim + 15) / 16;
            const uint32_t q_lo0 = desc_lo(smem_u32(sQ), 16);
            const uint32_t k_lo0 = desc_lo(smem_u32(sK), 16);
            const uint32_t p_lo0 = desc_lo(smem_u32(sP), 16);
            const uint32_t v_lo0 = desc_lo(smem_u32(sV), SLAB_BYTES);  // MN-major: LBO = slab stride
            const uint32_t tile_step = uint32_t(tile_bytes) >> 4;
            const uint32_t p_step = uint32_t(2 * SLAB_BYTES) >> 4;
            const uint32_t idesc_qk = p.idesc_qk, idesc_pv = p.idesc_pv;
            auto issue_qk = [&](int t, uint32_t k_lo) {
                const uint32_t q_lo = q_lo0 + uint32_t(t) * tile_step;
                const uint32_t tS = tmem_base + uint32_t(t) * 128u;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    if (kk < kq) {
                        const uint32_t off = uint32_t(kk >> 2) * (SLAB_BYTES >> 4) + uint32_t(kk & 3) * 2;
                        umma_f16_lohi(tS, q_lo + off, k_lo + off, kDescHiSw128, idesc_qk, kk > 0 ? 1u : 0u);
                    }
                }
            };
            auto issue_pv = [&](int t, uint32_t v_lo, int kp, uint32_t acc0) {
                const uint32_t p_lo = p_lo0 + uint32_t(t) * p_step;
                const uint32_t tO = tmem_base + 256u + uint32_t(t) * 128u;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    if (kk < kp) {
                        const uint32_t poff = uint32_t(kk >> 2) * (SLAB_BYTES >> 4) + uint32_t(kk & 3) * 2;
                        umma_f16_lohi(tO, p_lo + poff, v_lo + uint32_t(kk) * (2048 >> 4), kDescHiSw128, idesc_pv,
                                      kk > 0 ? 1u : acc0);
                    }
                }
            };
            // The two Q tiles run half a period apart: a tile's S = Q K^T is issued when the OTHER tile has
            // just copied its scores to registers, so one tile drains TMEM (tcgen05.ld, ~64 B/clk) while
            // the other sits in the exponentials (MUFU) instead of both fighting for the same unit.
            mbar_wait(q_full, 0);
            mbar_wait(&k_full[0], 0);
            tc_fence_after();
            if (elect_one()) {
                issue_qk(0, k_lo0);
                umma_commit(&s_full[0]);
            }
            mbar_wait(&s_free[0], 0);  // tile 0 holds S0(0) in registers
            tc_fence_after();
            if (elect_one()) {
                issue_qk(1, k_lo0);
                umma_commit(&s_full[1]);
                umma_commit(&k_empty[0]);
            }
            int sk = (kst > 1) ? 1 : 0;           // ring slot of K[j+1]
            uint32_t phk = (kst > 1) ? 0u : 1u;   // its full-barrier parity
            int sv = 0;
            uint32_t phv = 0;
            for (int j = 0; j < nkv; ++j) {
                const bool more = j + 1 < nkv;
                const int kv_valid = min(BKV, p.seq_kv - j * BKV);
                const int kp = (kv_valid + 15) >> 4;
                const uint32_t acc0 = j > 0 ? 1u : 0u;
                const uint32_t k_lo = k_lo0 + uint32_t(sk) * tile_step;
                const uint32_t v_lo = v_lo0 + uint32_t(sv) * tile_step;
                if (more) {  // tile 1 has copied S1(j): tile 0 may start on S0(j+1)
                    mbar_wait(&s_free[1], uint32_t(j & 1));
                    mbar_wait(&k_full[sk], phk);
                    tc_fence_after();
                    if (elect_one()) {
                        issue_qk(0, k_lo);
                        umma_commit(&s_full[0]);
                    }
                }
                mbar_wait(&p_full[0], uint32_t(j & 1));
                mbar_wait(&v_full[sv], phv);
                tc_fence_after();
                if (elect_one()) {
                    issue_pv(0, v_lo, kp, acc0);
                    umma_commit(&pv_done[0]);
                }
                if (more) {  // tile 0 has copied S0(j+1): tile 1 may start on S1(j+1)
                    mbar_wait(&s_free[0], uint32_t((j + 1) & 1));
                    tc_fence_after();
                    if (elect_one()) {
                        issue_qk(1, k_lo);
                        umma_commit(&s_full[1]);
                        umma_commit(&k_empty[sk]);
                    }
                    if (++sk == kst) {
                        sk = 0;
                        phk ^= 1;
                    }
                }
                mbar_wait(&p_full[1], uint32_t(j & 1));
                tc_fence_after();
                if (elect_one()) {
                    issue_pv(1, v_lo, kp, acc0);
                    umma_commit(&v_empty[sv]);
                    umma_commit(&pv_done[1]);
                }
                if (++sv == vst) {
                    sv = 0;
                    phv ^= 1;
                }
            }
        }
    } else {
        // ---------------- softmax warpgroups: thread <-> query row of tile t
        const int t = (warp - 2) >> 2;
        const int q = warp & 3;  // TMEM lane quadrant this warp may touch
        const int r = q * 32 + lane;
        const uint32_t lane_off = uint32_t(q * 32) << 16;
        const uint32_t tS = tmem_base + lane_off + uint32_t(t) * 128u;
        const uint32_t tO = tmem_base + lane_off + 256u + uint32_t(t) * 128u;
        uint8_t* prow = sP + t * (2 * SLAB_BYTES) + (r >> 3) * 1024 + (r & 7) * 128;
        const int rsw = r & 7;
        const float scale = p.scale_log2;

        float m_ref = -INFINITY;
        float l = 0.0f;
        // The exponentials (MUFU: 16 per clock per SM) are the scarce unit and the two warpgroups share it.  Left
        // alone they run in phase -- tile 1's first scores arrive right after tile 0 copied its own, and the lag
        // between two identical loops sharing one pipe is neutrally stable -- so both sit in the exponentials at
        // once and both wait at once (scripts/trace_attn.py: 1920 clocks of exponentials + 1280 of everything
        // else per KV tile, MUFU idle 40 % of the time).  Holding tile 1 back once by half a period keeps one
        // warpgroup in the exponentials while the other loads, reduces and waits.
        if (t == 1 && p.stagger > 0 && nkv > 2) {
            const long long t0 = clock64();
            while (clock64() - t0 < p.stagger) {
            }
        }
#ifdef SW_TRACE
        long long tr_acc[6] = {0, 0, 0, 0, 0, 0};  // wait S, load S, row max, wait PV, exponentials + P store, publish
        long long tr_t = clock64();
        const long long tr_begin = tr_t;
#define ATTN_TR(k)                       \
    {                                    \
        const long long now = clock64(); \
        tr_acc[k] += now - tr_t;         \
        tr_t = now;                      \
    }
#else
#define ATTN_TR(k)
#endif
        for (int j = 0; j < nkv; ++j) {
            const int kv_valid = min(BKV, p.seq_kv - j * BKV);
            if (lane == 0) mbar_wait(&s_full[t], uint32_t(j & 1));
            __syncwarp();
            tc_fence_after();
            ATTN_TR(0)
            // The whole score row moves to registers in one go and the TMEM tile is handed back at once:
            // the MMA warp computes the next S = Q K^T while this row is still in the exponentials.
            uint32_t s0[32], s1[32], s2[32], s3[32];
            tmem_ld32(tS, s0);
            tmem_ld32(tS + 32, s1);
            tmem_ld32(tS + 64, s2);
            tmem_ld32(tS + 96, s3);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&s_free[t]);
            ATTN_TR(1)
            // exact running maximum (online softmax); the reference only moves when the row outgrows
            // it by more than 2^8, so most tiles skip the O rescale
            // (four independent running maxima / sums per thread: with two softmax warps per scheduler a
            // 128-long dependent FMNMX or FADD chain is pure exposed latency)
            float mxa[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
            auto row_max = [&](const uint32_t (&v)[32], int c) {
                if (kv_valid == BKV) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) mxa[i & 3] = fmaxf(mxa[i & 3], __uint_as_float(v[i]));
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (c + i < kv_valid) mxa[i & 3] = fmaxf(mxa[i & 3], __uint_as_float(v[i]));
                }
            };
            row_max(s0, 0);
            row_max(s1, 32);
            row_max(s2, 64);
            row_max(s3, 96);
            const float mx = fmaxf(fmaxf(mxa[0], mxa[1]), fmaxf(mxa[2], mxa[3]));
            const float m_new = mx * scale;
            bool need = false;
            ATTN_TR(2)
            if (j == 0) {
                m_ref = m_new;
            } else {
                need = m_new > m_ref + 8.0f;
                if (lane == 0) mbar_wait(&pv_done[t], uint32_t((j - 1) & 1));  // O stable, P tile free
                __syncwarp();
                tc_fence_after();
                ATTN_TR(3)
                if (__any_sync(0xffffffffu, need)) {
                    const float alpha = need ? ex2(m_ref - m_new) : 1.0f;
                    if (need) m_ref = m_new;
#pragma unroll 1
                    for (int c = 0; c < p.dpad; c += 16) {
                        uint32_t o[16];
                        tmem_ld16(tO + c, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        tmem_st16(tO + c, o);
                    }
                    tmem_st_wait();
                    l *= alpha;
                }
            }
            // P = exp2(S*scale - m_ref) -> fp16 -> swizzled smem
            float ls[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            auto chunk = [&](const uint32_t (&v)[32], int c) {
                uint32_t pk[16];
                if (kv_valid == BKV) {
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        const float e0 = ex2(fmaf(__uint_as_float(v[i]), scale, -m_ref));
                        const float e1 = ex2(fmaf(__uint_as_float(v[i + 1]), scale, -m_ref));
                        ls[(i >> 1) & 3] += e0 + e1;
                        pk[i >> 1] = pack_h2(e0, e1);
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        const float e0 = c + i < kv_valid ? ex2(fmaf(__uint_as_float(v[i]), scale, -m_ref)) : 0.0f;
                        const float e1 = c + i + 1 < kv_valid ? ex2(fmaf(__uint_as_float(v[i + 1]), scale, -m_ref)) : 0.0f;
                        ls[(i >> 1) & 3] += e0 + e1;
                        pk[i >> 1] = pack_h2(e0, e1);
                    }
                }
                uint8_t* slab = prow + (c >> 6) * SLAB_BYTES;
                const int cc0 = (c & 63) >> 3;
#pragma unroll
                for (int g = 0; g < 4; ++g)
                    *reinterpret_cast<uint4*>(slab + (((cc0 + g) ^ rsw) << 4)) =
                        make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
            };
            chunk(s0, 0);
            chunk(s1, 32);
            chunk(s2, 64);
            chunk(s3, 96);
            l += (ls[0] + ls[1]) + (ls[2] + ls[3]);
            ATTN_TR(4)
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(&p_full[t]);
            ATTN_TR(5)
        }
#ifdef SW_TRACE
        if ((warp == 2 || warp == 6) && lane == 0) {
            for (int k = 0; k < 6; ++k) trace_value(t * 6 + k, (unsigned long long)tr_acc[k]);
            if (t == 0) trace_value(12, (unsigned long long)(clock64() - tr_begin));
            trace_value(13, trace_smid());
            trace_value(15, (unsigned long long)gridDim.x | ((unsigned long long)gridDim.y << 20) |
                                ((unsigned long long)gridDim.z << 40));
        }
#endif
        if (lane == 0) mbar_wait(&pv_done[t], uint32_t((nkv - 1) & 1));
        __syncwarp();
        tc_fence_after();
        const float inv_l = 1.0f / l;
        const int qi = q0 + t * BQ + r;
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
        tmem_dealloc_rt(tmem_base, 512u);
    }
}

__global__ void __launch_bounds__(kThreads2, 1)
    k_flash_attn2(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                  const __grid_constant__ CUtensorMap tmV, const AttnArgs p) {
    attn2_body(tmQ, &tmK, &tmV, p);
}

__global__ void __launch_bounds__(kThreads2, 1)
    k_flash_attn2_gather(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ KvMaps kv,
                         const AttnArgs p) {
    attn2_body(tmQ, kv.k, kv.v, p);
}

int make_qkv_map(CUtensorMap* m, const void* base, int d, int heads, int seq, int batch,
                 long long sb, long long ss, long long sh, int pack = 1) {
    const uint64_t dims[4] = {uint64_t(d), uint64_t(heads), uint64_t(seq), uint64_t(batch)};
    const uint64_t str[3] = {uint64_t(sh) * 2, uint64_t(ss) * 2, uint64_t(sb) * 2};
    const uint32_t box[4] = {64, 1, pack > 1 ? uint32_t(seq) : 128u, uint32_t(pack)};
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
    {
        static int env_stagger = -2;  // SW_ATTN_STAGGER: clocks (0 = off); default 1100
        if (env_stagger == -2) {
            const char* e = getenv("SW_ATTN_STAGGER");
            env_stagger = e ? atoi(e) : 1100;
        }
        a.stagger = env_stagger;
    }

    a.causal = p->causal ? 1 : 0;
    if (a.causal && (p->seq_q > BQ || p->seq_q != p->seq_kv))
        return swh::fail(-1, "sw_flash_attn: causal attention is defined for self-attention with seq <= %d", BQ);
    // short self-attention sequences (temporal attention over frames): several batch entries per tile
    a.pack = 1;
    // (queries and keys may differ in length: a frame-split clip attends from its own frames to all of them; row r of a
    // tile belongs to entry r / seq_q and sees columns [entry * seq_kv, (entry + 1) * seq_kv))
    const int seq_max = p->seq_q > p->seq_kv ? p->seq_q : p->seq_kv;
    if (seq_max <= 64 && p->batch > 1 && !a.causal) {
        a.pack = 128 / seq_max;
        if (a.pack > p->batch) a.pack = p->batch;
        if (a.pack > 128) a.pack = 128;
        static int no_pack = -1;  // SW_ATTN_PACK=0: one batch entry per tile (A/B runs)
        if (no_pack < 0) {
            const char* e = getenv("SW_ATTN_PACK");
            no_pack = (e && atoi(e) == 0) ? 1 : 0;
        }
        if (no_pack) a.pack = 1;
    }
    CUtensorMap tq, tk, tv;
    int rc = make_qkv_map(&tq, p->q, p->head_dim, p->heads, p->seq_q, p->batch, p->q_stride_b,
                          p->q_stride_s, p->q_stride_h, a.pack);
    if (rc) return rc;
    rc = make_qkv_map(&tk, p->k, p->head_dim, p->heads, p->seq_kv, p->batch, p->k_stride_b,
                      p->k_stride_s, p->k_stride_h, a.pack);
    if (rc) return rc;
    rc = make_qkv_map(&tv, p->v, p->head_dim, p->heads, p->seq_kv, p->batch, p->v_stride_b,
                      p->v_stride_s, p->v_stride_h, a.pack);
    if (rc) return rc;

    static bool configured[64] = {false};
    int dev = 0;
    SW_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 64 && !configured[dev]) {
        SW_CUDA_OK(cudaFuncSetAttribute(k_flash_attn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        SW_CUDA_OK(cudaFuncSetAttribute(k_flash_attn2, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        configured[dev] = true;
    }
    const int tile_bytes = a.nslab * SLAB_BYTES;
    static int force_v1 = -1;
    if (force_v1 < 0) force_v1 = getenv("SW_ATTN_V1") ? 1 : 0;
    if (!force_v1 && a.nslab <= 2 && p->seq_q > BQ) {
        // two query tiles per CTA, ping-pong softmax warpgroups
        a.k_stages = nkv > 1 ? 2 : 1;
        a.v_stages = (a.nslab == 1 && nkv > 1) ? 2 : 1;
        const int smem2 = tile_bytes * (2 + a.k_stages + a.v_stages) + 4 * SLAB_BYTES + 1024 + 256;
        if (smem2 <= 227 * 1024) {
            dim3 grid2((p->seq_q + 2 * BQ - 1) / (2 * BQ), p->heads, p->batch);
            SW_CUDA_OK(swh::launch_pdl(k_flash_attn2, grid2, dim3(kThreads2), size_t(smem2),
                                       static_cast<cudaStream_t>(stream), 1, tq, tk, tv, a));
            return 0;
        }
    }
    const int smem = tile_bytes * (1 + 2 * a.kv_stages) + 2 * SLAB_BYTES + 1024 + 256;
    if (smem > 227 * 1024) return swh::fail(-1, "sw_flash_attn: shared memory budget exceeded (%d)", smem);
    dim3 grid((p->seq_q + BQ - 1) / BQ, p->heads, (p->batch + a.pack - 1) / a.pack);
    SW_CUDA_OK(swh::launch_pdl(k_flash_attn, grid, dim3(kThreads), size_t(smem), static_cast<cudaStream_t>(stream), 1,
                               tq, tk, tv, a));
    return 0;
}

extern "C" int sw_flash_attn_gather(const sw_attn_params* p, const sw_peers* kpeers, const void* const* v_ptrs,
                                    int s_local, void* stream) {
    if (!p || !p->q || !p->out || !kpeers || !v_ptrs) return swh::fail(-1, "sw_flash_attn_gather: null argument");
    const int world = kpeers->world;
    if (world < 1 || world > 8) return swh::fail(-1, "sw_flash_attn_gather: world=%d", world);
    if (p->head_dim <= 0 || p->head_dim % 8 || p->head_dim > 128)
        return swh::fail(-1, "sw_flash_attn_gather: head_dim=%d unsupported (multiple of 8, <= 128)", p->head_dim);
    if (s_local <= 0 || s_local % BKV || p->seq_kv != s_local * world)
        return swh::fail(-1, "sw_flash_attn_gather: s_local must be a multiple of 128 and seq_kv == world*s_local");
    AttnArgs a;
    memset(&a, 0, sizeof(a));
    a.batch = p->batch;
    a.heads = p->heads;
    a.seq_q = p->seq_q;
    a.seq_kv = p->seq_kv;
    a.head_dim = p->head_dim;
    a.nslab = (p->head_dim + 63) / 64;
    a.dpad = (p->head_dim + 15) / 16 * 16;
    a.scale_log2 = p->scale * 1.4426950408889634f;
    a.out = static_cast<__half*>(p->out);
    a.o_stride_b = p->o_stride_b;
    a.o_stride_s = p->o_stride_s;
    a.idesc_qk = sw::umma_idesc(128, 128, 0, 0, 0);
    a.idesc_pv = sw::umma_idesc(128, a.dpad, 0, 0, 1);
    a.stagger = 1100;
    a.world = world;
    a.rank = kpeers->rank;
    a.epoch = kpeers->epoch;
    a.epoch_base = kpeers->epoch_base;
    a.epoch_mul = kpeers->epoch_mul;
    a.s_local = s_local;
    a.flags_local = kpeers->flags[kpeers->rank];
    const int nkv = p->seq_kv / BKV;
    a.k_stages = nkv > 1 ? 2 : 1;
    a.v_stages = (a.nslab == 1 && nkv > 1) ? 2 : 1;

    CUtensorMap tq;
    KvMaps kv;
    int rc = make_qkv_map(&tq, p->q, p->head_dim, p->heads, p->seq_q, p->batch, p->q_stride_b, p->q_stride_s,
                          p->q_stride_h);
    if (rc) return rc;
    for (int d = 0; d < 8; ++d) {
        const int s = d < world ? d : 0;
        rc = make_qkv_map(&kv.k[d], kpeers->peer[s], p->head_dim, p->heads, s_local, p->batch, p->k_stride_b,
                          p->k_stride_s, p->k_stride_h);
        if (rc) return rc;
        rc = make_qkv_map(&kv.v[d], v_ptrs[s], p->head_dim, p->heads, s_local, p->batch, p->v_stride_b,
                          p->v_stride_s, p->v_stride_h);
        if (rc) return rc;
    }
    static bool configured[64] = {false};
    int dev = 0;
    SW_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 64 && !configured[dev]) {
        SW_CUDA_OK(cudaFuncSetAttribute(k_flash_attn2_gather, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        227 * 1024));
        configured[dev] = true;
    }
    const int tile_bytes = a.nslab * SLAB_BYTES;
    const int smem2 = tile_bytes * (2 + a.k_stages + a.v_stages) + 4 * SLAB_BYTES + 1024 + 256;
    if (smem2 > 227 * 1024) return swh::fail(-1, "sw_flash_attn_gather: shared memory budget exceeded");
    dim3 grid2((p->seq_q + 2 * BQ - 1) / (2 * BQ), p->heads, p->batch);
    k_flash_attn2_gather<<<grid2, kThreads2, smem2, static_cast<cudaStream_t>(stream)>>>(tq, kv, a);
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}

// Probe builds (-DSW_TRACE) only: per-CTA softmax phase totals of the last flash-attention launch.
SW_TRACE_DUMP_FN(sw_trace_dump_attn)
