// Implicit-GEMM convolution / linear layer for sm_100a.
//
// One CTA computes a 128 x BN output tile: 128 output pixels (a BW x BH x BI box of the
// channels-last activation) by BN output channels.  The K loop walks (filter tap, 64-channel
// slab): for every step one TMA box load fetches the *shifted* activation window straight from
// the NHWC tensor (out-of-bounds rows/cols are zero-filled by the TMA unit, which is the conv
// padding) and one fetches the matching 64-wide weight slab.  tcgen05.mma accumulates in TMEM;
// four epilogue warps drain TMEM with tcgen05.ld and apply bias / time-embedding row bias /
// activation / GEGLU / residual before a vectorised fp16 store.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2..5 = epilogue (TMEM lane quadrant = warp_id % 4).
#include "../include/swarm_b200.h"
#include "common.cuh"
#include "host_util.h"

#include <stdlib.h>

namespace {

using namespace sw;

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kThreads = 192;
constexpr int kThreadsLean = 320;  // lean kernel: TMA warp, MMA warp, eight epilogue warps
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KB

struct IGemmArgs {
    int c1, c2;
    int n, h, w;
    int kh, kw, pad, pad_w, pad_y1;
    int sx;  // convolution stride (1 or 2): tile origin in input pixels = output origin * sx
    int phases;  // lean kernel: blockIdx.z = output phase (a, b) of a folded nearest-2x upsample + conv3x3
    int cout;
    int bw, bh, bi;  // tile box: bw*bh*bi <= 128 pixels
    int tiles_x, tiles_y;
    int stages;
    uint32_t idesc;
    const __half* bias;
    const __half* row_bias;
    int row_bias_ld;
    const __half* residual;
    int ldr;
    __half* out;
    int ldc;
    int act;
    float out_scale;
    int bf16;
    int bn;             // N tile (multiple of 32, <= 256)
    int tmem_cols;      // power of two >= bn
    // LayerNorm folded into the GEMM (weights pre-multiplied by gamma, bias' = W.beta + b on the host side):
    //   out = rstd_r * (acc - mean_r * col_sums[c]) + bias'[c]
    // with (mean, rstd) of each activation row rebuilt from per-column-tile partial sums that the GEMM
    // producing the activation wrote in its own epilogue (row_stats_out), in tile order = deterministic.
    float* row_stats_out;        // producer: [column tiles][rows][2] (sum, sum of squares) of the stored output
    const float* row_stats_in;   // consumer: the producer's buffer
    const float* col_sums;       // consumer: fp32 [cout], sum over K of the folded weight row
    int stats_tiles, ln_width;   // consumer: partial slots per row, elements per normalised row
    float ln_eps;
    int ptiles, total_tiles;  // persistent kernel: pixel tiles, pixel tiles x column tiles
    int tma_epi;        // epilogue stages 64-column slabs in shared memory and writes them with TMA stores
    int epi_alias;      // lean kernel: the epilogue boxes overlay the drained operand ring (else dedicated)
    int w_kmajor;       // weights packed k-block-major [k-blocks][cout][64]: tmB is 3-D {64, cout, k-blocks}
    int cluster_k;      // splits == 2 and the two splits of a tile form a cluster: partials meet in DSMEM
    int splits;         // split-K factor (grid.z)
    int kb_per_split;   // k-blocks per split
    unsigned int* tickets;  // [tiles] arrival counters (split-K)
    unsigned int* done;     // [tiles] departure counters: the last split to leave zeroes both
    float* partials;        // [tile][split][bn/32][128][32] fp32
    // ---- fused collectives (plain GEMMs only; see sw_igemm_ag / sw_igemm_rs)
    int mode;               // 0 local, 1 all-gather fused into the A loads, 2 reduce-scatter push epilogue
    int world, rank, epoch;
    int m_local;            // rows of the activation shard each rank owns
    __half* rs_out[8];      // mode 2: rank d's staging buffer [world][m_local][ldc]
    int* flags[8];          // rank d's int32[world] epoch flags
    unsigned int* rs_ticket;
    // ---- lean fused collectives (igemm_lean_body<.., MODE>): copy / reduce roles inside the GEMM grid, per-tile flags
    __half* gathered[8];    // MODE 1: rank d's gathered activation [world * m_local][a_ld] (symmetric)
    int a_ld;               // MODE 1: row pitch of the gathered buffer (elements)
    int shard_tiles;        // 128-row tiles per rank shard (m_local / 128)
    const int* epoch_base;  // optional device counter (CUDA-graph replays): epoch += *epoch_base * epoch_mul
    int epoch_mul;
    const __half* rs_stage; // MODE 2: this rank's staging [world][m_local][ldc] (what the peers push into)
    __half* rs_final;       // MODE 2: this rank's reduced rows [m_local][ld_final]
    int ld_final;
    int dbg;                // SW_FUSED_DBG bit mask (timing bisection only): 1 no completion wait / flag, 2 no reduce role
};

struct AgMaps {
    CUtensorMap m[8];       // mode 1: tensor map of every rank's activation shard (peer mappings)
};

__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void wait_flag_ge(const int* p, int v) {
    int cur;
    for (unsigned i = 0;; ++i) {
        asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(cur) : "l"(p) : "memory");
        if (cur >= v) break;
        if (i > (1u << 27)) __trap();
        __nanosleep(64);
    }
}
// timing probe variants of the wait (SW_FUSED_DBG bits 16 / 32 / 64)
__device__ __forceinline__ void wait_flag_ge_dbg(const int* p, int v, int dbg) {
    if (!(dbg & (16 | 32 | 64))) {
        wait_flag_ge(p, v);
        return;
    }
    int cur;
    for (unsigned i = 0;; ++i) {
        if (dbg & 64)
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(cur) : "l"(p) : "memory");
        else
            asm volatile("ld.relaxed.sys.global.s32 %0, [%1];" : "=r"(cur) : "l"(p) : "memory");
        if (cur >= v) break;
        if (i > (1u << 27)) __trap();
        __nanosleep(64);
    }
    if (dbg & 32) asm volatile("fence.acq_rel.sys;" ::: "memory");
    if (dbg & 16) asm volatile("fence.acq_rel.gpu;" ::: "memory");
}

// ---------------------------------------------------------------- raw-address PTX helpers
// The producer and MMA roles are single threads: every instruction they execute sits on the
// critical path of a K block (a lone thread retires a dependent instruction every ~5-30 cycles;
// a runtime division alone costs more than a whole MMA).  So their loops carry incremental
// counters and precomputed 32-bit shared addresses / descriptor halves -- no div/mod, no
// generic->shared conversions, no descriptor assembly inside the loop.
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // bit 24 of a shared::cluster address: CTA in the pair

__device__ __forceinline__ void bar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
    for (uint32_t i = 0; !ok; ++i) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
        if (i > SW_SPIN_LIMIT) __trap();
    }
}
__device__ __forceinline__ void bar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
template <bool PAIR>
__device__ __forceinline__ void tma4(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2, int c3) {
    if (PAIR)
        asm volatile(
            "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
            "[%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
            "l"(reinterpret_cast<uint64_t>(m)), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
            : "memory");
    else
        asm volatile(
            "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
            "[%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
            "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
            : "memory");
}
template <bool PAIR>
__device__ __forceinline__ void tma2(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
    if (PAIR)
        asm volatile(
            "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
            "[%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
            "l"(reinterpret_cast<uint64_t>(m)), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1)
            : "memory");
    else
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
            "[%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
            "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1)
            : "memory");
}
template <bool PAIR>
__device__ __forceinline__ void tma3(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2) {
    if (PAIR)
        asm volatile(
            "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
            "[%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
            "l"(reinterpret_cast<uint64_t>(m)), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2)
            : "memory");
    else
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
            "[%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
            "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
            : "memory");
}
// weight tile of k-block `kb`: plain [cout][K] layout (columns kcol..kcol+63 of rows bcol0..) or the k-block-major
// packing [k-blocks][cout][64] in which a tile is one contiguous run of memory (one DRAM page burst per 1 KB
// instead of a 128-byte segment per page)
template <bool PAIR>
__device__ __forceinline__ void tma_b(uint32_t dst, const CUtensorMap* m, uint32_t bar, int kmajor, int kcol, int kb,
                                      int bcol0) {
    if (kmajor)
        tma3<PAIR>(dst, m, bar, 0, bcol0, kb);
    else
        tma2<PAIR>(dst, m, bar, kcol, bcol0);
}
template <bool PAIR>
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi,
                                        uint32_t idesc, uint32_t accumulate) {
    if (PAIR)
        asm volatile(
            "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
            "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
            "setp.ne.b32 p, %5, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, p;\n\t}" ::"r"(tmem_d),
            "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate)
            : "memory");
    else
        asm volatile(
            "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
            "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
            "setp.ne.b32 p, %5, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}" ::"r"(tmem_d),
            "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate)
            : "memory");
}
// Arrive on `bar` (same offset in both CTAs for a pair) once all MMAs issued so far are done.
template <bool PAIR>
__device__ __forceinline__ void mma_commit(uint32_t bar) {
    if (PAIR) {
        const uint16_t mask = 0x3;
        asm volatile(
            "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::
                "r"(bar),
            "h"(mask)
            : "memory");
    } else {
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
                     : "memory");
    }
}
template <bool PAIR>
__device__ __forceinline__ void tmem_alloc_any(uint32_t* slot, uint32_t cols) {
    if (PAIR) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)),
                     "r"(cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)),
                     "r"(cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
}
template <bool PAIR>
__device__ __forceinline__ void tmem_dealloc_any(uint32_t taddr, uint32_t cols) {
    if (PAIR)
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
    else
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
template <bool PAIR>
__device__ __forceinline__ void cta_or_cluster_sync() {
    if (PAIR) {
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    } else {
        __syncthreads();
    }
}

// PAIR == false: one CTA computes a 128 x BN tile (tcgen05 cta_group::1).
// PAIR == true : a cluster of two CTAs computes a 256 x BN tile (cta_group::2): each CTA stages its
// own 128 pixels of A and half of the weight tile, the leader issues the M = 256 MMAs which read
// both halves of B, so every SM pulls only half of the weight operand; accumulators land in both
// CTAs' TMEM and each CTA runs its own epilogue.
template <bool PAIR>
__device__ __forceinline__ void igemm_body(const CUtensorMap& tmA1, const CUtensorMap& tmA2,
                                           const CUtensorMap& tmB, const CUtensorMap& tmOut,
                                           const CUtensorMap& tmRes, const CUtensorMap& tmOut32, const IGemmArgs& p) {
    const int BN = p.bn;
    const int B_STAGE_BYTES = (PAIR ? BN / 2 : BN) * BK * 2;
    const int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
    __shared__ __align__(16) float s_bias[256];
    __shared__ __align__(16) float s_cs[256];

    extern __shared__ uint8_t smem_raw[];
    // SWIZZLE_128B operands need 1024-byte aligned stage bases.
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

    // (warp index through a shuffle: ptxas then knows it is warp-uniform, see the note on the role loops)
    const int warp = __shfl_sync(0xffffffffu, int(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int stages = p.stages;
    if (threadIdx.x == 0) {
        TR(0);
        TRG(11);
    }
    uint32_t cta_rank = 0;
    if (PAIR) {
        asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
        cta_rank = __shfl_sync(0xffffffffu, cta_rank, 0);
    }
    const bool leader = cta_rank == 0;

    // (the TMA-store epilogue stages 4 x 16 KB boxes over the drained ring: keep the barriers clear of them)
    const int ring_bytes = (p.tma_epi && stages * STAGE_BYTES < 65536) ? 65536 : stages * STAGE_BYTES;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + ring_bytes);
    uint64_t* empty = full + stages;
    uint64_t* acc_full = empty + stages;
    uint64_t* res_full = acc_full + 1;  // [2] residual slabs (TMA-store epilogue)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_full + 2);

    // spatial tile -> origin pixel
    int t = blockIdx.x;
    if (p.mode != 0) {
        // fused collectives: rotate the tile order by rank so that (all-gather) local rows are
        // consumed first and (reduce-scatter) the ranks push to different owners at any moment
        const int per_rank = p.m_local / BM;
        t = (t + (p.rank + (p.mode == 2 ? 1 : 0)) * per_rank) % int(gridDim.x);
    }
    const int tx = t % p.tiles_x;
    t /= p.tiles_x;
    const int ty = t % p.tiles_y;
    const int tn = t / p.tiles_y;
    const int x0 = tx * p.bw, y0 = ty * p.bh, n0 = tn * p.bi;
    const int xs0 = x0 * p.sx, ys0 = y0 * p.sx;  // tile origin in input pixels
    const int col0 = blockIdx.y * BN;
    const int bcol0 = col0 + (PAIR ? int(cta_rank) * (BN / 2) : 0);
    // mode 1 (all-gather fused): this tile's rows live on rank `ag_src`; they are fetched straight
    // from that rank's memory by TMA over NVLink, local rows first (tile order is rotated by rank).
    const CUtensorMap* pA1 = &tmA1;
    int ax0 = x0;
    int ag_src = -1;
    if (p.mode == 1) {
        ag_src = x0 / p.m_local;
        ax0 = x0 - ag_src * p.m_local;
        pA1 = &tmA1 + ag_src;  // tmA1 is element 0 of an AgMaps array
    }

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA1);
        tma_prefetch_desc(&tmA2);
        tma_prefetch_desc(&tmB);
        if (p.tma_epi) {
            tma_prefetch_desc(&tmOut);
            tma_prefetch_desc(&tmRes);
            tma_prefetch_desc(&tmOut32);
        }
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < stages; ++s) {
                mbar_init(&full[s], 1);
                mbar_init(&empty[s], 1);
            }
            mbar_init(acc_full, 1);
            mbar_init(&res_full[0], 1);
            mbar_init(&res_full[1], 1);
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc_any<PAIR>(tmem_slot, uint32_t(p.tmem_cols));
    }
    tc_fence_before();
    cta_or_cluster_sync<PAIR>();  // (pair: peer barriers must exist before any remote arrive)
    tc_fence_after();
    const uint32_t tmem_acc = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    if (threadIdx.x == 0) TR(1);
    pdl_trigger();  // the next kernel may start its prologue
    pdl_wait();     // our inputs (and the buffers we overwrite) belong to the previous kernel until here
    if (threadIdx.x == 0) TR(2);

    const int nkb1 = (p.c1 + BK - 1) / BK;
    const int nkb2 = (p.c2 + BK - 1) / BK;
    const int per_tap = nkb1 + nkb2;
    const int ctot = p.c1 + p.c2;
    const int total_all = p.kh * p.kw * per_tap;
    const int it_begin = blockIdx.z * p.kb_per_split;
    const int it_end = min(total_all, it_begin + p.kb_per_split);
    const uint32_t a_bytes = uint32_t(p.bw * p.bh * p.bi) * (BK * 2);
    const uint32_t smem0 = smem_u32(smem);
    const uint32_t full0 = smem_u32(full), empty0 = smem_u32(empty);

    if (warp == 0) {
        {
            // ---------------- TMA producer (lean loop, see note above).  The whole warp walks the loop with
            // warp-uniform values and ONE elected lane issues: with the loop under `lane == 0` ptxas treated every
            // operand as divergent and wrapped each UTMALDG in an ELECT / R2UR.BROADCAST / BRA.U.ANY waterfall
            // (a k-block cost ~565 clocks of issue whatever the tile size: scripts/trace_igemm.py).
            int tap = it_begin / per_tap;
            int j = it_begin - tap * per_tap;
            int ky = tap / p.kw, kx = tap - ky * p.kw;
            int tapbase = tap * ctot;
            const uint32_t tx_bytes = (PAIR ? 2u : 1u) * (a_bytes + uint32_t(B_STAGE_BYTES));
            uint32_t sa = smem0, fb = full0, eb = empty0;
            int s = 0;
            uint32_t ph = 1;
            if (p.mode == 1 && ag_src != p.rank) {
                wait_flag_ge(p.flags[p.rank] + ag_src, p.epoch);
                asm volatile("fence.proxy.async;" ::: "memory");  // generic-proxy flag read -> async-proxy (TMA) loads
            }
            if (lane == 0) TR(3);
            for (int it = it_begin; it < it_end; ++it) {
                bar_wait(eb, ph);
                if (elect_one()) {
                    // both CTAs' loads complete on the leader's barrier; only the leader arms it
                    if (!PAIR || leader) bar_expect_tx(fb, tx_bytes);
                    if (j < nkb1) {
                        tma4<PAIR>(sa, pA1, fb, j * BK, ax0 * p.sx + kx - p.pad_w, y0 * p.sx + ky - p.pad_y1, n0);
                        tma_b<PAIR>(sa + A_STAGE_BYTES, &tmB, fb, p.w_kmajor, tapbase + j * BK, it, bcol0);
                    } else {
                        tma4<PAIR>(sa, &tmA2, fb, (j - nkb1) * BK, xs0 + kx - p.pad_w, ys0 + ky - p.pad, n0);
                        tma_b<PAIR>(sa + A_STAGE_BYTES, &tmB, fb, p.w_kmajor, tapbase + p.c1 + (j - nkb1) * BK, it, bcol0);
                    }
                }
                if (++j == per_tap) {
                    j = 0;
                    tapbase += ctot;
                    if (++kx == p.kw) {
                        kx = 0;
                        ++ky;
                    }
                }
                if (++s == stages) {
                    s = 0;
                    ph ^= 1;
                    sa = smem0;
                    fb = full0;
                    eb = empty0;
                } else {
                    sa += STAGE_BYTES;
                    fb += 8;
                    eb += 8;
                }
            }
            if (lane == 0) TR(4);
        }
        if (!PAIR && p.cluster_k) cluster_sync_all();  // (whole cluster: see the epilogue)
    } else if (warp == 1) {
        if (!PAIR || leader) {
            // ---------------- MMA issuer (lean loop; whole warp walks it, one elected lane issues -- the same lane
            // every time, so the commits track the MMAs)
            // descriptor = {lo: addr>>4 | LBO(16 B)>>4 << 16, hi: SBO(1024 B)>>4 | version 1 | SWIZZLE_128B}
            const uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
            const uint32_t a_lo0 = ((smem0 & 0x3FFFFu) >> 4) | (1u << 16);
            const uint32_t b_off = uint32_t(A_STAGE_BYTES) >> 4;
            const uint32_t stage_step = uint32_t(STAGE_BYTES) >> 4;
            // K steps (of 16) in the last slab of each input; all other slabs have 4
            const int ks1 = (p.c1 - (nkb1 - 1) * BK + 15) / 16;
            const int ks2 = nkb2 ? (p.c2 - (nkb2 - 1) * BK + 15) / 16 : 4;
            const int last1 = nkb1 - 1, last2 = per_tap - 1;
            int j = it_begin % per_tap;
            uint32_t a_lo = a_lo0, fb = full0, eb = empty0;
            int s = 0;
            uint32_t ph = 0, acc = 0;
            const uint32_t idesc = p.idesc;
            for (int it = it_begin; it < it_end; ++it) {
                bar_wait(fb, ph);
                tc_fence_after();
#ifdef SW_TRACE
                if (it == it_begin && lane == 0) TR(5);
#endif
                int ksteps = 4;
                if (j == last1) ksteps = ks1 < 4 ? ks1 : 4;
                if (nkb2 && j == last2) ksteps = ks2 < 4 ? ks2 : 4;
                if (elect_one()) {
                    // +32 bytes (2 descriptor units) per 16-element K step inside the 128-byte swizzle row
                    mma_f16<PAIR>(tmem_acc, a_lo, a_lo + b_off, desc_hi, idesc, acc);
                    if (ksteps > 1) mma_f16<PAIR>(tmem_acc, a_lo + 2, a_lo + b_off + 2, desc_hi, idesc, 1u);
                    if (ksteps > 2) mma_f16<PAIR>(tmem_acc, a_lo + 4, a_lo + b_off + 4, desc_hi, idesc, 1u);
                    if (ksteps > 3) mma_f16<PAIR>(tmem_acc, a_lo + 6, a_lo + b_off + 6, desc_hi, idesc, 1u);
                    mma_commit<PAIR>(eb);
                }
                acc = 1u;
                if (++j == per_tap) j = 0;
                if (++s == stages) {
                    s = 0;
                    ph ^= 1;
                    a_lo = a_lo0;
                    fb = full0;
                    eb = empty0;
                } else {
                    a_lo += stage_step;
                    fb += 8;
                    eb += 8;
                }
            }
            if (elect_one()) mma_commit<PAIR>(smem_u32(acc_full));
            if (lane == 0) TR(6);
        }
        if (!PAIR && p.cluster_k) cluster_sync_all();
    } else {
        // ---------------- epilogue: one thread per accumulator row
        const int q = warp & 3;
        const int r = q * 32 + lane;
        const int xl = r % p.bw;
        const int r2 = r / p.bw;
        const int yl = r2 % p.bh;
        const int nl = r2 / p.bh;
        const int x = x0 + xl, y = y0 + yl, ni = n0 + nl;
        const bool row_ok = (nl < p.bi) && x < p.w && y < p.h && ni < p.n;
        const size_t m = (size_t(ni) * p.h + y) * p.w + x;

        // per-column addends of this tile (bias, plus the time-embedding row bias when the whole tile
        // lies in one image) go to shared memory while the main loop runs: the per-chunk epilogue then
        // has no global-load latency on its critical path
        const bool fold_rb = p.row_bias != nullptr && p.bi == 1;
        for (int i = threadIdx.x - 64; i < BN; i += 128) {
            const int col = col0 + i;
            float b = 0.0f;
            if (col < p.cout) {
                if (p.bias) b = __half2float(p.bias[col]);
                if (fold_rb && n0 < p.n) b += __half2float(p.row_bias[size_t(n0) * p.row_bias_ld + col]);
            }
            s_bias[i] = b;
            s_cs[i] = (p.col_sums && col < p.cout) ? p.col_sums[col] : 0.0f;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        float ln_mean = 0.0f, ln_rstd = 1.0f;
        if (p.row_stats_in && row_ok) {
            const size_t rows_total = size_t(p.n) * p.h * p.w;
            float su = 0.0f, sq = 0.0f;
            for (int ti = 0; ti < p.stats_tiles; ++ti) {
                const float2 pr = *reinterpret_cast<const float2*>(p.row_stats_in + (size_t(ti) * rows_total + m) * 2);
                su += pr.x;
                sq += pr.y;
            }
            const float inv_w = 1.0f / float(p.ln_width);
            ln_mean = su * inv_w;
            ln_rstd = rsqrtf(fmaxf(sq * inv_w - ln_mean * ln_mean, 0.0f) + p.ln_eps);
        }
        float st_sum = 0.0f, st_sq = 0.0f;  // producer side: this row's sums over the tile's columns

        if (threadIdx.x == 64) TR(7);
        if (lane == 0) mbar_wait(acc_full, 0);  // one poller per warp: 128 spinning threads steal issue slots
        __syncwarp();
        tc_fence_after();
        if (threadIdx.x == 64) TR(8);
        const uint32_t trow = tmem_acc + (uint32_t(q * 32) << 16);

        // ---- split-K: every split parks its fp32 tile in scratch, then ALL splits of the tile (they are
        // co-resident: the host sizes the grid to one wave) wait for each other and each reduces its own
        // band of rows -- sums in split order, so the result is bit-reproducible -- and runs the epilogue
        // for that band.  The reduction is spread over every participating SM instead of serialising
        // on the CTA that happened to arrive last (that version lost to no split at all beyond 4 splits).
        if (p.splits > 1 && !p.cluster_k) {
            const int tile_id = blockIdx.x * gridDim.y + blockIdx.y;
            const int nchunk = BN / 32;
            float* tile_base = p.partials + size_t(tile_id) * p.splits * nchunk * 4096;
            float* mine = tile_base + size_t(blockIdx.z) * nchunk * 4096 + r * 32;
            uint32_t vn[32];
            tmem_ld32(trow, vn);
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t v[32];
                tmem_ld_wait_dep(vn);
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = vn[j];
                if (c + 32 < BN) tmem_ld32(trow + c + 32, vn);
                float4* dst = reinterpret_cast<float4*>(mine + (c / 32) * 4096);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    dst[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                         __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
            }
            __threadfence();
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (threadIdx.x == 64) {
                atomicAdd(&p.tickets[tile_id], 1u);
                const volatile unsigned int* t = p.tickets + tile_id;
                for (unsigned int i = 0; *t < unsigned(p.splits); ++i) {
                    if (i > (1u << 26)) __trap();  // a split that never became resident must not hang the box
                    __nanosleep(20);
                }
                __threadfence();
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            // my band of rows; items = (row, 8-column group), consecutive threads -> consecutive columns
            const int band = (128 + p.splits - 1) / p.splits;
            const int row_begin = int(blockIdx.z) * band;
            const int row_end = min(128, row_begin + band);
            const int g8s = BN >> 3;
            const int items = max(0, row_end - row_begin) * g8s;
            for (int i = threadIdx.x - 64; i < items; i += 128) {
                const int rr = row_begin + i / g8s;
                const int c8 = (i - (i / g8s) * g8s) << 3;
                const int col = col0 + c8;
                const int xl2 = rr % p.bw, t2 = rr / p.bw;
                const int yl2 = t2 % p.bh, nl2 = t2 / p.bh;
                const int xx = x0 + xl2, yy = y0 + yl2, nn = n0 + nl2;
                if (!(nl2 < p.bi && xx < p.w && yy < p.h && nn < p.n) || col >= p.cout) continue;
                const size_t mm = (size_t(nn) * p.h + yy) * p.w + xx;
                float f[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = 0.0f;
                const float* src = tile_base + size_t(c8 >> 5) * 4096 + rr * 32 + (c8 & 31);
                const size_t zstride = size_t(nchunk) * 4096;
                // four splits' partials in flight per thread (the loop was one dependent L2 round trip per split:
                // ~700 clocks x splits, most of a split-K layer's time -- scripts/trace_igemm.py conv rows); the sums
                // are still taken in split order, so the result is bit-identical
                int z = 0;
                for (; z + 4 <= p.splits; z += 4) {
                    float4 a0[4], a1[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        a0[u] = __ldcg(reinterpret_cast<const float4*>(src + size_t(z + u) * zstride));
                        a1[u] = __ldcg(reinterpret_cast<const float4*>(src + size_t(z + u) * zstride) + 1);
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        f[0] += a0[u].x, f[1] += a0[u].y, f[2] += a0[u].z, f[3] += a0[u].w;
                        f[4] += a1[u].x, f[5] += a1[u].y, f[6] += a1[u].z, f[7] += a1[u].w;
                    }
                }
                for (; z < p.splits; ++z) {
                    const float4 a0 = __ldcg(reinterpret_cast<const float4*>(src + size_t(z) * zstride));
                    const float4 a1 = __ldcg(reinterpret_cast<const float4*>(src + size_t(z) * zstride) + 1);
                    f[0] += a0.x, f[1] += a0.y, f[2] += a0.z, f[3] += a0.w;
                    f[4] += a1.x, f[5] += a1.y, f[6] += a1.z, f[7] += a1.w;
                }
                auto add_h8 = [&](const __half* ptr) {
                    const uint4 b = *reinterpret_cast<const uint4*>(ptr);
                    const float2 b0 = unpack_h2(b.x), b1 = unpack_h2(b.y), b2 = unpack_h2(b.z), b3 = unpack_h2(b.w);
                    f[0] += b0.x, f[1] += b0.y, f[2] += b1.x, f[3] += b1.y;
                    f[4] += b2.x, f[5] += b2.y, f[6] += b3.x, f[7] += b3.y;
                };
                if (p.bias) add_h8(p.bias + col);
                if (p.row_bias) add_h8(p.row_bias + size_t(nn) * p.row_bias_ld + col);
                if (p.act == SW_ACT_SILU) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] = silu_f(f[j]);
                } else if (p.act == SW_ACT_GELU) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] = gelu_f(f[j]);
                }
                if (p.out_scale != 1.0f) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] *= p.out_scale;
                }
                if (p.residual) add_h8(p.residual + mm * p.ldr + col);
                *reinterpret_cast<uint4*>(p.out + mm * p.ldc + col) =
                    make_uint4(pack_h2(f[0], f[1]), pack_h2(f[2], f[3]), pack_h2(f[4], f[5]), pack_h2(f[6], f[7]));
            }
            // departure: the last split to finish its band re-arms both counters for the next launch
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (threadIdx.x == 64) {
                if (atomicAdd(&p.done[tile_id], 1u) == unsigned(p.splits) - 1u) {
                    p.tickets[tile_id] = 0u;
                    p.done[tile_id] = 0u;
                }
            }
        }
        // ---- split-K over a cluster of two CTAs (small GEMMs: every CTA otherwise streams the whole
        // 128 x K activation slab at ~40 B/clk): each CTA accumulates half of K, ships the OTHER CTA's half
        // of the columns through distributed shared memory, and after one cluster barrier finishes its own
        // half.  No global round trip, no spinning; a + b is the same in both orders, so still deterministic.
        int kc_begin = 0, kc_end = BN;
        const float* kstage = nullptr;
        if (!PAIR && p.cluster_k) {
            const uint32_t zr = blockIdx.z;  // == rank in the (1,1,2) cluster
            const int half = BN >> 1;
            float* stage = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(full) + 256);
            uint32_t remote;
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(stage)), "r"(zr ^ 1u));
            const int pc0 = int(zr ^ 1u) * half;
#pragma unroll 1
            for (int k = 0; k < half; k += 32) {
                uint32_t v[32];
                tmem_ld32(trow + pc0 + k, v);
                tmem_ld_wait();
                const uint32_t dst = remote + uint32_t(((k >> 5) * 128 + r) * 128);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + 16 * j), "r"(v[4 * j]),
                                 "r"(v[4 * j + 1]), "r"(v[4 * j + 2]), "r"(v[4 * j + 3])
                                 : "memory");
            }
            cluster_sync_all();
            kc_begin = int(zr) * half;
            kc_end = kc_begin + half;
            kstage = stage;
        }
        const bool do_epilogue = p.splits == 1 || p.cluster_k;
        if (!do_epilogue) {
            // another split owns the epilogue of this tile
        } else if (p.tma_epi) {
            // ---- TMA-store epilogue.  Row-per-thread 16-byte global stores touch 32 different lines per
            // instruction and serialise in the LSU (measured ~0.4 us per 32 columns); instead each thread
            // writes its row of a 64-column slab into a 128B-swizzled shared-memory box (conflict-free),
            // and one thread hands the box to the TMA unit, which also clips rows / columns that fall
            // outside the tensor.  The residual arrives the same way (TMA load, two slabs ahead).  The
            // staging boxes alias the operand ring: every MMA has retired once acc_full fired.
            const bool geglu = p.act == SW_ACT_GEGLU;
            const int ncols_tile = geglu ? BN / 2 : BN;
            const int ocol0 = geglu ? int(blockIdx.y) * (BN / 2) : col0;
            int nvalid = (geglu ? p.cout / 2 : p.cout) - ocol0;
            if (nvalid > ncols_tile) nvalid = ncols_tile;
            const int cend = (nvalid + 31) & ~31;
            const int nslab = (nvalid + 63) >> 6;
            // a last slab of only 32 columns (BN = 160, 224) must not spill a 64-wide box into the next
            // tile's columns: it is staged unswizzled and stored through a 32-column tensor map
            const bool tail32 = (cend & 63) == 32;
            const bool has_res = p.residual != nullptr;
            const bool lead = threadIdx.x == 64;
            const uint32_t rfull0 = smem_u32(res_full);
            const uint32_t res_stage = smem0 + 32768u;
            if (has_res && lead) {
                for (int sl = 0; sl < 2 && sl < nslab; ++sl) {
                    bar_expect_tx(rfull0 + 8 * sl, a_bytes);
                    tma4<false>(res_stage + sl * 16384, &tmRes, rfull0 + 8 * sl, ocol0 + sl * 64, x0, y0, n0);
                }
            }
            const uint32_t row_off = uint32_t(r) * 128u, swz = uint32_t(r) & 7u;
            const bool plain = (p.act == SW_ACT_NONE || geglu) && p.out_scale == 1.0f && (!p.row_bias || fold_rb);

            // begin / end of a 64-column slab (all 128 epilogue threads take the same path)
            auto slab_begin = [&](int sl) {
                if (sl >= 2) {  // the store issued two slabs ago must have finished reading this box
                    if (lead) bulk_wait_read<1>();
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                }
                if (has_res) bar_wait(rfull0 + 8 * (sl & 1), uint32_t(sl >> 1) & 1u);
            };
            auto slab_end = [&](int sl) {
                fence_proxy_async_smem();
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (lead) {
                    tma_store_4d((tail32 && sl == nslab - 1) ? &tmOut32 : &tmOut, smem0 + uint32_t(sl & 1) * 16384u,
                                 ocol0 + sl * 64, x0, y0, n0);
                    bulk_commit();
                    if (has_res && sl + 2 < nslab) {
                        bar_expect_tx(rfull0 + 8 * (sl & 1), a_bytes);
                        tma4<false>(res_stage + (sl & 1) * 16384, &tmRes, rfull0 + 8 * (sl & 1), ocol0 + (sl + 2) * 64,
                                    x0, y0, n0);
                    }
                }
            };
            // 8 output columns f[0..8) at column c8 (multiple of 8) of the tile -> staging box
            auto put8 = [&](int c8, float (&f)[8]) {
                const uint32_t off =
                    (uint32_t(c8 >> 6) & 1u) * 16384u + row_off + (((uint32_t(c8 >> 3) & 7u) ^ swz) << 4);
                if (!plain) {
                    if (p.row_bias && !fold_rb && row_ok) {
                        const uint4 b =
                            *reinterpret_cast<const uint4*>(p.row_bias + size_t(ni) * p.row_bias_ld + col0 + c8);
                        const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float2 t2 = unpack_h2(bw[j]);
                            f[2 * j] += t2.x;
                            f[2 * j + 1] += t2.y;
                        }
                    }
                    if (p.act == SW_ACT_SILU) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) f[j] = silu_f(f[j]);
                    } else if (p.act == SW_ACT_GELU) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) f[j] = gelu_f(f[j]);
                    }
                    if (p.out_scale != 1.0f) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) f[j] *= p.out_scale;
                    }
                }
                if (has_res) {
                    const uint4 rv = lds128(res_stage + off);
                    const float2 r0 = unpack_h2(rv.x), r1 = unpack_h2(rv.y), r2 = unpack_h2(rv.z),
                                 r3 = unpack_h2(rv.w);
                    f[0] += r0.x, f[1] += r0.y, f[2] += r1.x, f[3] += r1.y;
                    f[4] += r2.x, f[5] += r2.y, f[6] += r3.x, f[7] += r3.y;
                }
                if (p.row_stats_out) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        st_sum += f[j];
                        st_sq = fmaf(f[j], f[j], st_sq);
                    }
                }
                uint32_t dst = smem0 + off;
                if (tail32 && (c8 >> 6) == nslab - 1)
                    dst = smem0 + (uint32_t(c8 >> 6) & 1u) * 16384u + uint32_t(r) * 64u + ((uint32_t(c8 >> 3) & 3u) << 4);
                sts128(dst, make_uint4(pack_h2(f[0], f[1]), pack_h2(f[2], f[3]), pack_h2(f[4], f[5]),
                                       pack_h2(f[6], f[7])));
            };
            if (geglu) {
                // columns [0,BN/2) of the accumulator are values, [BN/2,BN) their gates
                const int HALF = BN / 2;
                uint32_t vn[16], gn[16];
                tmem_ld16(trow, vn);
                tmem_ld16(trow + HALF, gn);
#pragma unroll 1
                for (int c = 0; c < cend; c += 16) {
                    if ((c & 63) == 0) slab_begin(c >> 6);
                    uint32_t v[16], g[16];
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] = vn[j], g[j] = gn[j];
                    if (c + 16 < cend) {
                        tmem_ld16(trow + c + 16, vn);
                        tmem_ld16(trow + HALF + c + 16, gn);
                    }
#pragma unroll
                    for (int h8 = 0; h8 < 2; ++h8) {
                        float f[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float a = __uint_as_float(v[h8 * 8 + j]) + s_bias[c + h8 * 8 + j];
                            const float gt = __uint_as_float(g[h8 * 8 + j]) + s_bias[HALF + c + h8 * 8 + j];
                            f[j] = a * gelu_f(gt);
                        }
                        put8(c + h8 * 8, f);
                    }
                    if ((c & 63) == 48 || c + 16 >= cend) slab_end(c >> 6);
                }
            } else {
                uint32_t vn[32];
                tmem_ld32(trow, vn);
#pragma unroll 1
                for (int c = 0; c < cend; c += 32) {
                    if ((c & 63) == 0) slab_begin(c >> 6);
                    uint32_t v[32];
                    tmem_ld_wait_dep(vn);
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = vn[j];
                    if (c + 32 < cend) tmem_ld32(trow + c + 32, vn);
#pragma unroll
                    for (int g8 = 0; g8 < 4; ++g8) {
                        const float4 b0 = *reinterpret_cast<const float4*>(&s_bias[c + g8 * 8]);
                        const float4 b1 = *reinterpret_cast<const float4*>(&s_bias[c + g8 * 8 + 4]);
                        const float4 k0 = *reinterpret_cast<const float4*>(&s_cs[c + g8 * 8]);
                        const float4 k1 = *reinterpret_cast<const float4*>(&s_cs[c + g8 * 8 + 4]);
                        // (mean = 0, rstd = 1, col sums = 0 when no LayerNorm is folded in: plain acc + bias)
                        float f[8] = {fmaf(ln_rstd, fmaf(-ln_mean, k0.x, __uint_as_float(v[g8 * 8 + 0])), b0.x),
                                      fmaf(ln_rstd, fmaf(-ln_mean, k0.y, __uint_as_float(v[g8 * 8 + 1])), b0.y),
                                      fmaf(ln_rstd, fmaf(-ln_mean, k0.z, __uint_as_float(v[g8 * 8 + 2])), b0.z),
                                      fmaf(ln_rstd, fmaf(-ln_mean, k0.w, __uint_as_float(v[g8 * 8 + 3])), b0.w),
                                      fmaf(ln_rstd, fmaf(-ln_mean, k1.x, __uint_as_float(v[g8 * 8 + 4])), b1.x),
                                      fmaf(ln_rstd, fmaf(-ln_mean, k1.y, __uint_as_float(v[g8 * 8 + 5])), b1.y),
                                      fmaf(ln_rstd, fmaf(-ln_mean, k1.z, __uint_as_float(v[g8 * 8 + 6])), b1.z),
                                      fmaf(ln_rstd, fmaf(-ln_mean, k1.w, __uint_as_float(v[g8 * 8 + 7])), b1.w)};
                        put8(c + g8 * 8, f);
                    }
                    if ((c & 63) == 32 || c + 32 >= cend) slab_end(c >> 6);
                }
            }
            if (p.row_stats_out && row_ok) {
                const size_t rows_total = size_t(p.n) * p.h * p.w;
                *reinterpret_cast<float2*>(p.row_stats_out + (size_t(blockIdx.y) * rows_total + m) * 2) =
                    make_float2(st_sum, st_sq);
            }
            if (lead) bulk_wait_read<0>();
        } else if (p.act == SW_ACT_GEGLU) {
            // columns [0,BN/2) of the tile are values, [BN/2,BN) their gates; the TMEM loads of chunk
            // c+16 are in flight while chunk c is converted and stored
            const int HALF = BN / 2;
            const int ocol0 = blockIdx.y * HALF;
            const int ncols_out = p.cout / 2;
            auto emit = [&](int c, const uint32_t (&v)[16], const uint32_t (&g)[16]) {
                if (!(row_ok && ocol0 + c < ncols_out)) return;
                uint32_t o[8];
#pragma unroll
                for (int j = 0; j < 16; j += 2) {
                    const float a0 = fmaf(ln_rstd, fmaf(-ln_mean, s_cs[c + j], __uint_as_float(v[j])), s_bias[c + j]);
                    const float a1 = fmaf(ln_rstd, fmaf(-ln_mean, s_cs[c + j + 1], __uint_as_float(v[j + 1])), s_bias[c + j + 1]);
                    const float g0 = fmaf(ln_rstd, fmaf(-ln_mean, s_cs[HALF + c + j], __uint_as_float(g[j])), s_bias[HALF + c + j]);
                    const float g1 = fmaf(ln_rstd, fmaf(-ln_mean, s_cs[HALF + c + j + 1], __uint_as_float(g[j + 1])),
                                          s_bias[HALF + c + j + 1]);
                    o[j / 2] = pack_h2(a0 * gelu_f(g0), a1 * gelu_f(g1));
                }
                uint4* dst = reinterpret_cast<uint4*>(p.out + m * p.ldc + ocol0 + c);
                dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
                dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
            };
            uint32_t v0[16], g0[16], v1[16], g1[16];
            tmem_ld16(trow, v0);
            tmem_ld16(trow + HALF, g0);
#pragma unroll 1
            for (int c = 0; c < HALF; c += 32) {
                tmem_ld_wait();
                const bool more1 = c + 16 < HALF;
                if (more1) {
                    tmem_ld16(trow + c + 16, v1);
                    tmem_ld16(trow + HALF + c + 16, g1);
                }
                emit(c, v0, g0);
                if (more1) {
                    tmem_ld_wait();
                    if (c + 32 < HALF) {
                        tmem_ld16(trow + c + 32, v0);
                        tmem_ld16(trow + HALF + c + 32, g0);
                    }
                    emit(c + 16, v1, g1);
                }
            }
        } else {
            // one 32-column chunk: addends from shared memory, residual already in registers
            // the common case (bias / folded row bias / residual only) gets a short straight-line body:
            // these kernels live for microseconds, so the instructions actually fetched matter
            const bool plain = p.act == SW_ACT_NONE && p.out_scale == 1.0f && p.mode != 2 && (!p.row_bias || fold_rb);
            const bool has_res = p.residual != nullptr;
            __half* const out_row = p.out + m * p.ldc + col0;
            auto emit = [&](int c, const uint32_t (&v)[32], const uint4 (&rs)[4]) {
                if (!row_ok) return;
                if (plain) {
#pragma unroll
                    for (int g8 = 0; g8 < 4; ++g8) {
                        if (col0 + c + g8 * 8 < p.cout) {
                            const float4 b0 = *reinterpret_cast<const float4*>(&s_bias[c + g8 * 8]);
                            const float4 b1 = *reinterpret_cast<const float4*>(&s_bias[c + g8 * 8 + 4]);
                            float f0 = __uint_as_float(v[g8 * 8 + 0]) + b0.x, f1 = __uint_as_float(v[g8 * 8 + 1]) + b0.y;
                            float f2 = __uint_as_float(v[g8 * 8 + 2]) + b0.z, f3 = __uint_as_float(v[g8 * 8 + 3]) + b0.w;
                            float f4 = __uint_as_float(v[g8 * 8 + 4]) + b1.x, f5 = __uint_as_float(v[g8 * 8 + 5]) + b1.y;
                            float f6 = __uint_as_float(v[g8 * 8 + 6]) + b1.z, f7 = __uint_as_float(v[g8 * 8 + 7]) + b1.w;
                            if (has_res) {
                                const float2 r0 = unpack_h2(rs[g8].x), r1 = unpack_h2(rs[g8].y);
                                const float2 r2 = unpack_h2(rs[g8].z), r3 = unpack_h2(rs[g8].w);
                                f0 += r0.x, f1 += r0.y, f2 += r1.x, f3 += r1.y;
                                f4 += r2.x, f5 += r2.y, f6 += r3.x, f7 += r3.y;
                            }
                            *reinterpret_cast<uint4*>(out_row + c + g8 * 8) =
                                make_uint4(pack_h2(f0, f1), pack_h2(f2, f3), pack_h2(f4, f5), pack_h2(f6, f7));
                        }
                    }
                    return;
                }
#pragma unroll
                for (int g8 = 0; g8 < 4; ++g8) {
                    const int col = col0 + c + g8 * 8;
                    if (col >= p.cout) break;
                    float f[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[g8 * 8 + j]) + s_bias[c + g8 * 8 + j];
                    if (p.row_bias && !fold_rb) {  // tile spans several images (tiny feature maps)
                        const uint4 b = *reinterpret_cast<const uint4*>(
                            p.row_bias + size_t(ni) * p.row_bias_ld + col);
                        const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float2 t2 = unpack_h2(bw[j]);
                            f[2 * j] += t2.x;
                            f[2 * j + 1] += t2.y;
                        }
                    }
                    if (p.act == SW_ACT_SILU) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) f[j] = silu_f(f[j]);
                    } else if (p.act == SW_ACT_GELU) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) f[j] = gelu_f(f[j]);
                    }
                    if (p.out_scale != 1.0f) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) f[j] *= p.out_scale;
                    }
                    if (p.residual) {
                        const uint32_t bw[4] = {rs[g8].x, rs[g8].y, rs[g8].z, rs[g8].w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float2 t2 = unpack_h2(bw[j]);
                            f[2 * j] += t2.x;
                            f[2 * j + 1] += t2.y;
                        }
                    }
                    __half* dst = p.out + m * p.ldc + col;
                    if (p.mode == 2) {
                        // reduce-scatter push: the partial row goes to its owner's staging slot [rank]
                        const int d = int(m) / p.m_local;
                        const size_t lr = size_t(p.rank) * p.m_local + (m - size_t(d) * p.m_local);
                        dst = p.rs_out[d] + lr * p.ldc + col;
                    }
                    *reinterpret_cast<uint4*>(dst) =
                        make_uint4(pack_h2(f[0], f[1]), pack_h2(f[2], f[3]), pack_h2(f[4], f[5]),
                                   pack_h2(f[6], f[7]));
                }
            };
            // residual of chunk c (read before this thread writes those columns: in-place adds are fine)
            auto fetch = [&](int c, uint4 (&rs)[4]) {
                if (!p.residual || !row_ok) return;
#pragma unroll
                for (int g8 = 0; g8 < 4; ++g8) {
                    const int col = col0 + c + g8 * 8;
                    if (col < p.cout) rs[g8] = *reinterpret_cast<const uint4*>(p.residual + m * p.ldr + col);
                }
            };
            {
                // TMEM chunk c+32 and its residual are in flight while chunk c is converted and stored
                uint32_t vn[32];
                uint4 rn[4];
                tmem_ld32(trow + kc_begin, vn);
                fetch(kc_begin, rn);
#pragma unroll 1
                for (int c = kc_begin; c < kc_end; c += 32) {
                    uint32_t v[32];
                    uint4 rs[4];
                    tmem_ld_wait_dep(vn);
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = vn[j];
#pragma unroll
                    for (int j = 0; j < 4; ++j) rs[j] = rn[j];
                    if (c + 32 < kc_end) {
                        tmem_ld32(trow + c + 32, vn);
                        fetch(c + 32, rn);
                    }
                    if (kstage != nullptr) {  // the other half of K, delivered by the cluster peer
                        const float4* ps = reinterpret_cast<const float4*>(kstage + (((c - kc_begin) >> 5) * 128 + r) * 32);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 a = ps[j];
                            v[4 * j] = __float_as_uint(__uint_as_float(v[4 * j]) + a.x);
                            v[4 * j + 1] = __float_as_uint(__uint_as_float(v[4 * j + 1]) + a.y);
                            v[4 * j + 2] = __float_as_uint(__uint_as_float(v[4 * j + 2]) + a.z);
                            v[4 * j + 3] = __float_as_uint(__uint_as_float(v[4 * j + 3]) + a.w);
                        }
                    }
                    emit(c, v, rs);
                }
            }
        }
    }

    if (threadIdx.x == 64) TR(9);
    if (p.mode == 2) {
        // every CTA publishes its pushes system-wide; the CTA that finishes last raises this
        // rank's epoch flag on every peer (their reduce kernels poll it)
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned int total = gridDim.x * gridDim.y;
            if (atomicAdd(p.rs_ticket, 1u) == total - 1u) {
                *p.rs_ticket = 0u;
                __threadfence();
                for (int d = 0; d < p.world; ++d) st_release_sys(p.flags[d] + p.rank, p.epoch);
            }
        }
    }
    tc_fence_before();
    cta_or_cluster_sync<PAIR>();  // (pair: both epilogues have drained TMEM)
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_any<PAIR>(tmem_acc, uint32_t(p.tmem_cols));
    }
#ifdef SW_TRACE
    if (threadIdx.x == 0) {
        TR(10);
        TRG(12);
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        const unsigned cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        if (cta < unsigned(kTraceCtas)) {
            g_sw_trace[cta * kTraceSlots + 13] = smid;
            g_sw_trace[cta * kTraceSlots + 14] = (unsigned long long)p.bn | ((unsigned long long)p.stages << 16) |
                                                 ((unsigned long long)p.splits << 24) | ((unsigned long long)(PAIR ? 1 : 0) << 32) |
                                                 ((unsigned long long)p.cluster_k << 33) | ((unsigned long long)p.tma_epi << 34);
            g_sw_trace[cta * kTraceSlots + 15] = (unsigned long long)gridDim.x | ((unsigned long long)gridDim.y << 20) |
                                                 ((unsigned long long)gridDim.z << 40);
        }
    }
#endif
}

__global__ void __launch_bounds__(kThreads)
    k_igemm(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmA2,
            const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
            const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmOut32,
    const IGemmArgs p) {
    igemm_body<false>(tmA1, tmA2, tmB, tmOut, tmRes, tmOut32, p);
}

__global__ void __launch_bounds__(kThreads)
    k_igemm_ag(const __grid_constant__ AgMaps maps, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmOut32,
    const IGemmArgs p) {
    igemm_body<false>(maps.m[0], maps.m[0], tmB, tmOut, tmRes, tmOut32, p);
}

__global__ void __launch_bounds__(kThreads)
    k_igemm2_ag(const __grid_constant__ AgMaps maps, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmOut32,
    const IGemmArgs p) {
    igemm_body<true>(maps.m[0], maps.m[0], tmB, tmOut, tmRes, tmOut32, p);
}

__global__ void __launch_bounds__(kThreads)
    k_igemm2(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmA2,
             const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
             const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmOut32,
    const IGemmArgs p) {
    igemm_body<true>(tmA1, tmA2, tmB, tmOut, tmRes, tmOut32, p);
}

// =================================================================================================
// Lean variant: the same tile / ring / TMEM design for the one epilogue that almost every layer of a UNet step
// uses -- bias (+ the per-image time-embedding bias when the tile lies in one image) (+ residual), fp16 out
// through swizzled staging boxes and TMA stores -- and nothing else.  scripts/trace_igemm.py (per-CTA clock
// stamps) showed where a small GEMM spends its life in the general kernel above: ~1500 clocks of set-up, ~600
// from the dependency wait to the first TMA (integer divisions), ~1000 of first-load latency, a main loop paced
// by the issuing threads, and an epilogue of ~5600-6900 clocks for a 128x128 tile -- half of the kernel --
// because one warp per scheduler walks ~1400 SASS instructions per 32 columns of a 127 KB kernel image whose
// every line is an instruction-cache miss in a kernel that lives for microseconds.  Here all run-time epilogue
// choices are template parameters or gone, everything that does not touch global memory is computed before the
// dependency wait, and (when the grid is a single wave anyway) the residual tile gets its own staging boxes so
// its TMA loads are issued before the main loop instead of after it.
//
// MODE 1 / 2 are the fused collectives of a tensor-parallel block (plain GEMMs, single-CTA tiles).  Both keep the
// math path above untouched and add a second ROLE to the same grid, tied to it by per-tile epoch flags:
//   MODE 1, all-gather -> GEMM: blocks with blockIdx.y == 0 (the first the hardware dispatches) copy this rank's
//     activation shard, tile by tile, into every peer's gathered buffer with 16-byte stores over NVLink and raise
//     flag [rank][tile] on the peer; a compute block whose 128 rows came from rank s waits for flag [s][tile] before
//     its first A load, and then reads LOCAL memory only -- a remote tile crosses NVLink exactly once, and the tile
//     order is rotated so every rank starts on rows it already holds.
//   MODE 2, GEMM -> reduce-scatter: a compute block hands its fp16 partial tile to the TMA store unit with the tensor
//     map of the rows' OWNER (its staging slot for this rank, over NVLink) and raises flag [rank][tile_m][tile_n]
//     there once the stores have completed; blocks with blockIdx.y == gridDim.y - 1 (the last dispatched) each own
//     some tiles of this rank's rows: they wait for the `world` flags of a tile and add the slots in rank order in
//     fp32 (+ bias, + residual) -- deterministic, and a tile is reduced as soon as its last partial lands instead
//     of after the whole GEMM and a kernel boundary.
// Neither role ever waits for a block with a higher dispatch index of its own grid, so the waits cannot cycle.
struct PeerMaps {
    CUtensorMap m[8];
};

__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ int lean_epoch(const IGemmArgs& p) {
    return p.epoch + (p.epoch_base ? __ldcg(p.epoch_base) * p.epoch_mul : 0);
}

__device__ __forceinline__ void lean_push_role(const IGemmArgs& p) {
    // MODE 1 copy role: tile blockIdx.x of this rank's shard -> the same rows of every peer's gathered buffer
    const int tile = blockIdx.x;
    pdl_trigger();
    pdl_wait();  // the shard is the previous kernel's output
    if (tile >= p.shard_tiles) return;
    const int cpr = p.c1 >> 3;                       // 16-byte chunks per row
    const int ld16 = p.a_ld >> 3;
    const size_t row0 = size_t(p.rank) * p.m_local + size_t(tile) * BM;
    const uint4* __restrict__ src = reinterpret_cast<const uint4*>(p.gathered[p.rank]) + row0 * ld16;
    const int total = BM * cpr;
    for (int i = threadIdx.x; i < total; i += 4 * int(blockDim.x)) {
        uint4 v[4];
        int off[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int ii = i + u * int(blockDim.x);
            const int r = ii / cpr;
            off[u] = ii < total ? r * ld16 + (ii - r * cpr) : -1;
            if (off[u] >= 0) v[u] = src[off[u]];
        }
        for (int d = 0; d < p.world; ++d) {
            if (d == p.rank) continue;
            uint4* __restrict__ dst = reinterpret_cast<uint4*>(p.gathered[d]) + row0 * ld16;
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (off[u] >= 0) dst[off[u]] = v[u];
        }
    }
    __syncthreads();  // (every thread's stores happen before the release stores below; no system fence needed)
    if (int(threadIdx.x) < p.world && int(threadIdx.x) != p.rank)
        st_release_sys(p.flags[threadIdx.x] + p.rank * p.shard_tiles + tile, lean_epoch(p));
}

__device__ __forceinline__ void lean_reduce_role(const IGemmArgs& p, int b) {
    // MODE 2 reduce role: block b of shard_tiles * ntile owns ONE tile of this rank's rows, in the order the
    // partials arrive; out = sum over src (rank order) of staging[src] (+ bias) (+ residual)
    const int ntile = (p.cout + p.bn - 1) / p.bn;
    const int tm = b / ntile;
    const int r0 = tm * BM;
    pdl_trigger();
    pdl_wait();  // (residual / out belong to earlier kernels until here)
    for (int tn = b % ntile; tn < ntile; tn += ntile) {
        if (int(threadIdx.x) < p.world && !(p.dbg & 4))
            wait_flag_ge_dbg(p.flags[p.rank] + (int(threadIdx.x) * p.shard_tiles + tm) * ntile + tn, lean_epoch(p), p.dbg);
        __syncthreads();
        const int c0 = tn * p.bn;
        const int ncol = min(p.bn, p.cout - c0);
        const int octs = ncol >> 3;
        for (int i = threadIdx.x; i < BM * octs; i += blockDim.x) {
            const int r = r0 + i / octs, c = c0 + ((i % octs) << 3);
            float acc[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = 0.0f;
            auto add8 = [&](const __half* ptr) {
                const uint4 v = __ldcg(reinterpret_cast<const uint4*>(ptr));
                const float2 a = unpack_h2(v.x), b = unpack_h2(v.y), cc = unpack_h2(v.z), d = unpack_h2(v.w);
                acc[0] += a.x, acc[1] += a.y, acc[2] += b.x, acc[3] += b.y;
                acc[4] += cc.x, acc[5] += cc.y, acc[6] += d.x, acc[7] += d.y;
            };
            for (int src = 0; src < p.world; ++src) add8(p.rs_stage + (size_t(src) * p.m_local + r) * p.ldc + c);
            if (p.bias) add8(p.bias + c);
            if (p.residual) add8(p.residual + size_t(r) * p.ldr + c);
            *reinterpret_cast<uint4*>(p.rs_final + size_t(r) * p.ld_final + c) =
                make_uint4(pack_h2(acc[0], acc[1]), pack_h2(acc[2], acc[3]), pack_h2(acc[4], acc[5]),
                           pack_h2(acc[6], acc[7]));
        }
        __syncthreads();
    }
}

template <bool PAIR, bool HAS_RES, int MODE = 0>
__device__ __forceinline__ void igemm_lean_body(const CUtensorMap& tmA1, const CUtensorMap& tmA2,
                                                const CUtensorMap& tmB, const CUtensorMap& tmOut,
                                                const CUtensorMap& tmRes, const CUtensorMap& tmOut32,
                                                const IGemmArgs& p, const PeerMaps* pm = nullptr) {
    // fused modes run on a 1-D grid so that the dispatch order is exactly: [copy blocks] -> compute blocks, row tile
    // major in rank-rotated order (every column tile of the rows this rank already holds before any row that has to
    // arrive; rows this rank owns last) -> [reduce blocks]
    const int ntile_f = MODE != 0 ? (p.cout + p.bn - 1) / p.bn : 1;
    int cblock = int(blockIdx.x);        // compute block index
    if (MODE == 1) {
        if (cblock < p.shard_tiles) {
            lean_push_role(p);
            return;
        }
        cblock -= p.shard_tiles;
    }
    if (MODE == 2 && cblock >= p.tiles_x * ntile_f) {
        if (!(p.dbg & 2)) lean_reduce_role(p, cblock - p.tiles_x * ntile_f);
        return;
    }
    const int ctile_f = MODE != 0 ? cblock % ntile_f : int(blockIdx.y);
    const int BN = p.bn;
    const int B_STAGE_BYTES = (PAIR ? BN / 2 : BN) * BK * 2;
    const int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
    __shared__ __align__(16) float s_bias[256];

    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

    const int warp = __shfl_sync(0xffffffffu, int(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int stages = p.stages;
    if (threadIdx.x == 0) {
        TR(0);
        TRG(11);
    }
    uint32_t cta_rank = 0;
    if (PAIR) {
        asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
        cta_rank = __shfl_sync(0xffffffffu, cta_rank, 0);
    }
    const bool leader = cta_rank == 0;

    // shared memory: operand ring | epilogue boxes (out[2], res[2]; they alias the drained ring when
    // p.epi_alias) | barriers
    const int ring_only = stages * STAGE_BYTES;
    const int ring_bytes = p.epi_alias ? (ring_only < 65536 ? 65536 : ring_only) : ring_only;
    const int box_off = p.epi_alias ? 0 : ring_bytes;
    const int bars_off = p.epi_alias ? ring_bytes : ring_bytes + (HAS_RES ? 65536 : 32768);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + bars_off);
    uint64_t* empty = full + stages;
    uint64_t* acc_full = empty + stages;
    uint64_t* res_full = acc_full + 1;  // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_full + 2);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA1);
        tma_prefetch_desc(&tmA2);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmOut);
        if (HAS_RES) tma_prefetch_desc(&tmRes);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < stages; ++s) {
                mbar_init(&full[s], 1);
                mbar_init(&empty[s], 1);
            }
            mbar_init(acc_full, 1);
            mbar_init(&res_full[0], 1);
            mbar_init(&res_full[1], 1);
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc_any<PAIR>(tmem_slot, uint32_t(p.tmem_cols));
    }

    // ---- everything that needs no global memory comes before the dependency wait
    int t = blockIdx.x;
    if (MODE != 0)  // rank-rotated tile order: (all-gather) own rows first, (reduce-scatter) own rows last
        t = (cblock / ntile_f + (p.rank + (MODE == 2 ? 1 : 0)) * p.shard_tiles) % p.tiles_x;
    const int tx = t % p.tiles_x;
    t /= p.tiles_x;
    const int ty = t % p.tiles_y;
    const int tn = t / p.tiles_y;
    const int x0 = tx * p.bw, y0 = ty * p.bh, n0 = tn * p.bi;
    const int xs0 = x0 * p.sx, ys0 = y0 * p.sx;  // tile origin in input pixels
    const int col0 = ctile_f * BN;
    const int bcol0 = col0 + (PAIR ? int(cta_rank) * (BN / 2) : 0);
    // folded upsample (p.phases): the four output phases (a, b) of out[2y+a, 2x+b] are the z slices of ONE grid; a
    // phase has its own taps (rows a-1.., columns b-1..), its own block of weight rows and its own output map
    // (the kernel arguments tmOut, tmRes, tmOut32, tmA2 carry the maps of phases 0..3: none of them has its usual
    // role in a bias-only, single-input, cout % 64 == 0 convolution)
    const int phase = (MODE == 0 && p.phases) ? int(blockIdx.z) : 0;
    const int pad_yp = p.pad_y1 - (phase >> 1), pad_xp = p.pad_w - (phase & 1);
    const int brow0 = bcol0 + phase * p.cout;
    const CUtensorMap* const out_map = phase == 0 ? &tmOut : phase == 1 ? &tmRes : phase == 2 ? &tmOut32 : &tmA2;
    const int peer_of_tile = MODE != 0 ? x0 / p.m_local : 0;                 // who produced (1) / owns (2) these rows
    const int tile_in_shard = MODE != 0 ? (x0 - peer_of_tile * p.m_local) / BM : 0;
    const int nkb1 = (p.c1 + BK - 1) / BK;
    const int nkb2 = (p.c2 + BK - 1) / BK;
    const int per_tap = nkb1 + nkb2;
    const int ctot = p.c1 + p.c2;
    const int total_kb = p.kh * p.kw * per_tap;
    const uint32_t a_bytes = uint32_t(p.bw * p.bh * p.bi) * (BK * 2);
    const uint32_t smem0 = smem_u32(smem);
    const uint32_t full0 = smem_u32(full), empty0 = smem_u32(empty);
    const uint32_t out_stage = smem0 + uint32_t(box_off);
    const uint32_t res_stage = out_stage + 32768u;
    const uint32_t rfull0 = smem_u32(res_full);
    int nvalid = p.cout - col0;
    if (nvalid > BN) nvalid = BN;
    const int cend = (nvalid + 31) & ~31;
    const int nslab = (nvalid + 63) >> 6;

    tc_fence_before();
    cta_or_cluster_sync<PAIR>();
    tc_fence_after();
    const uint32_t tmem_acc = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    if (threadIdx.x == 0) TR(1);
    pdl_trigger();
    pdl_wait();
    if (threadIdx.x == 0) TR(2);

    if (warp == 0) {
        // ---------------- TMA producer: the whole warp walks the loop, one elected lane issues
        int j = 0, ky = 0, kx = 0, tapbase = 0;
        const uint32_t tx_bytes = (PAIR ? 2u : 1u) * (a_bytes + uint32_t(B_STAGE_BYTES));
        uint32_t sa = smem0, fb = full0, eb = empty0;
        int s = 0;
        uint32_t ph = 1;
        if (MODE == 1 && peer_of_tile != p.rank) {
            // these rows are another rank's: its copy role raises the flag once they have landed in OUR buffer.
            // The data arrived through generic-proxy stores and is about to be read by the async proxy (TMA).
            wait_flag_ge(p.flags[p.rank] + peer_of_tile * p.shard_tiles + tile_in_shard, lean_epoch(p));
            fence_proxy_async_all();
        }
        if (lane == 0) TR(3);
        for (int it = 0; it < total_kb; ++it) {
            bar_wait(eb, ph);
            if (elect_one()) {
                if (!PAIR || leader) bar_expect_tx(fb, tx_bytes);
                if (j < nkb1) {
                    tma4<PAIR>(sa, &tmA1, fb, j * BK, xs0 + kx - pad_xp, ys0 + ky - pad_yp, n0);
                    tma_b<PAIR>(sa + A_STAGE_BYTES, &tmB, fb, p.w_kmajor, tapbase + j * BK, it, brow0);
                } else {
                    tma4<PAIR>(sa, &tmA2, fb, (j - nkb1) * BK, xs0 + kx - p.pad_w, ys0 + ky - p.pad, n0);
                    tma_b<PAIR>(sa + A_STAGE_BYTES, &tmB, fb, p.w_kmajor, tapbase + p.c1 + (j - nkb1) * BK, it, bcol0);
                }
            }
            if (++j == per_tap) {
                j = 0;
                tapbase += ctot;
                if (++kx == p.kw) {
                    kx = 0;
                    ++ky;
                }
            }
            if (++s == stages) {
                s = 0;
                ph ^= 1;
                sa = smem0;
                fb = full0;
                eb = empty0;
            } else {
                sa += STAGE_BYTES;
                fb += 8;
                eb += 8;
            }
        }
        if (lane == 0) TR(4);
    } else if (warp == 1) {
        if (!PAIR || leader) {
            // ---------------- MMA issuer (whole warp walks the loop; the same elected lane issues MMAs and commits)
            const uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
            const uint32_t a_lo0 = ((smem0 & 0x3FFFFu) >> 4) | (1u << 16);
            const uint32_t b_off = uint32_t(A_STAGE_BYTES) >> 4;
            const uint32_t stage_step = uint32_t(STAGE_BYTES) >> 4;
            const int ks1 = (p.c1 - (nkb1 - 1) * BK + 15) / 16;
            const int ks2 = nkb2 ? (p.c2 - (nkb2 - 1) * BK + 15) / 16 : 4;
            const int last1 = nkb1 - 1, last2 = per_tap - 1;
            int j = 0;
            uint32_t a_lo = a_lo0, fb = full0, eb = empty0;
            int s = 0;
            uint32_t ph = 0, acc = 0;
            const uint32_t idesc = p.idesc;
            for (int it = 0; it < total_kb; ++it) {
                bar_wait(fb, ph);
                tc_fence_after();
#ifdef SW_TRACE
                if (it == 0 && lane == 0) TR(5);
#endif
                int ksteps = 4;
                if (j == last1) ksteps = ks1 < 4 ? ks1 : 4;
                if (nkb2 && j == last2) ksteps = ks2 < 4 ? ks2 : 4;
                if (elect_one()) {
                    mma_f16<PAIR>(tmem_acc, a_lo, a_lo + b_off, desc_hi, idesc, acc);
                    if (ksteps > 1) mma_f16<PAIR>(tmem_acc, a_lo + 2, a_lo + b_off + 2, desc_hi, idesc, 1u);
                    if (ksteps > 2) mma_f16<PAIR>(tmem_acc, a_lo + 4, a_lo + b_off + 4, desc_hi, idesc, 1u);
                    if (ksteps > 3) mma_f16<PAIR>(tmem_acc, a_lo + 6, a_lo + b_off + 6, desc_hi, idesc, 1u);
                    mma_commit<PAIR>(eb);
                }
                acc = 1u;
                if (++j == per_tap) j = 0;
                if (++s == stages) {
                    s = 0;
                    ph ^= 1;
                    a_lo = a_lo0;
                    fb = full0;
                    eb = empty0;
                } else {
                    a_lo += stage_step;
                    fb += 8;
                    eb += 8;
                }
            }
            if (elect_one()) mma_commit<PAIR>(smem_u32(acc_full));
            if (lane == 0) TR(6);
        }
    } else {
        // ---------------- epilogue: EIGHT warps, two per TMEM lane quadrant.  Warps 2-5 drain the even 64-column
        // slabs of the accumulator, warps 6-9 the odd ones; each warpgroup has its own staging boxes, named
        // barrier and TMA store stream, so a 128-column tile leaves in half the time of the four-warp version
        // (the epilogue is latency-bound: one warp per scheduler walking a serial instruction stream).
        const int half = (warp - 2) >> 2;
        const int q = warp & 3;
        const int r = q * 32 + lane;
        const bool lead = lane == 0 && ((warp - 2) & 3) == 0;  // threads 64 and 192
        const uint32_t bar_id = 1u + uint32_t(half);
        const bool tail32 = (cend & 63) == 32;  // last slab is 32 columns wide: staged unswizzled, own tensor map
        const uint32_t box = uint32_t(half) * 16384u;
        const uint32_t rfull = rfull0 + 8u * uint32_t(half);
        // the residual boxes do not alias the ring: fetch this warpgroup's first slab while the main loop runs
        if (HAS_RES && !p.epi_alias && lead && half < nslab) {
            bar_expect_tx(rfull, a_bytes);
            tma4<false>(res_stage + box, &tmRes, rfull, col0 + half * 64, x0, y0, n0);
        }
        // per-column addends (bias, plus the time-embedding row bias: the host only picks this kernel when a
        // tile lies within one image) go to shared memory while the main loop runs
        for (int i = threadIdx.x - 64; i < BN; i += 256) {
            const int col = col0 + i;
            float b = 0.0f;
            if (col < p.cout && MODE != 2) {  // (reduce-scatter: bias and residual are added once, by the reduce role)
                if (p.bias) b = __half2float(p.bias[col]);
                if (p.row_bias && n0 < p.n) b += __half2float(p.row_bias[size_t(n0) * p.row_bias_ld + col]);
            }
            s_bias[i] = b;
        }
        asm volatile("bar.sync 3, 256;" ::: "memory");
        if (threadIdx.x == 64) TR(7);
        if (lane == 0) mbar_wait(acc_full, 0);
        __syncwarp();
        tc_fence_after();
        if (threadIdx.x == 64) TR(8);
        if (HAS_RES && p.epi_alias && lead && half < nslab) {  // the boxes overlay the ring: every MMA has retired
            bar_expect_tx(rfull, a_bytes);
            tma4<false>(res_stage + box, &tmRes, rfull, col0 + half * 64, x0, y0, n0);
        }
        const uint32_t trow = tmem_acc + (uint32_t(q * 32) << 16);
        const uint32_t row_off = uint32_t(r) * 128u, swz = uint32_t(r) & 7u;
        int k = 0;  // slabs this warpgroup has handled
#pragma unroll 1
        for (int sl = half; sl < nslab; sl += 2, ++k) {
            const int c0 = sl << 6;
            const bool two = c0 + 32 < cend;
            uint32_t v0[32], v1[32];
            tmem_ld32(trow + c0, v0);
            if (two) tmem_ld32(trow + c0 + 32, v1);
            if (k > 0) {  // this warpgroup's previous store must have finished reading the box
                if (lead) bulk_wait_read<0>();
                asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
            }
            if (HAS_RES) bar_wait(rfull, uint32_t(k) & 1u);
            const bool tail_slab = tail32 && sl == nslab - 1;
            tmem_ld_wait_dep(v0);
            auto chunk = [&](const uint32_t (&v)[32], int c) {
#pragma unroll
                for (int g8 = 0; g8 < 4; ++g8) {
                    const int c8 = c + g8 * 8;
                    const float4 b0 = *reinterpret_cast<const float4*>(&s_bias[c8]);
                    const float4 b1 = *reinterpret_cast<const float4*>(&s_bias[c8 + 4]);
                    float f0 = __uint_as_float(v[g8 * 8 + 0]) + b0.x, f1 = __uint_as_float(v[g8 * 8 + 1]) + b0.y;
                    float f2 = __uint_as_float(v[g8 * 8 + 2]) + b0.z, f3 = __uint_as_float(v[g8 * 8 + 3]) + b0.w;
                    float f4 = __uint_as_float(v[g8 * 8 + 4]) + b1.x, f5 = __uint_as_float(v[g8 * 8 + 5]) + b1.y;
                    float f6 = __uint_as_float(v[g8 * 8 + 6]) + b1.z, f7 = __uint_as_float(v[g8 * 8 + 7]) + b1.w;
                    const uint32_t off = box + row_off + (((uint32_t(c8 >> 3) & 7u) ^ swz) << 4);
                    if (HAS_RES) {
                        const uint4 rv = lds128(res_stage + off);
                        const float2 r0 = unpack_h2(rv.x), r1 = unpack_h2(rv.y), r2 = unpack_h2(rv.z), r3 = unpack_h2(rv.w);
                        f0 += r0.x, f1 += r0.y, f2 += r1.x, f3 += r1.y;
                        f4 += r2.x, f5 += r2.y, f6 += r3.x, f7 += r3.y;
                    }
                    const uint32_t dst = tail_slab ? out_stage + box + uint32_t(r) * 64u + ((uint32_t(c8 >> 3) & 3u) << 4)
                                                   : out_stage + off;
                    sts128(dst, make_uint4(pack_h2(f0, f1), pack_h2(f2, f3), pack_h2(f4, f5), pack_h2(f6, f7)));
                }
            };
            chunk(v0, c0);
            if (two) {
                tmem_ld_wait_dep(v1);  // (both loads were issued together: this is a no-op wait that pins the order)
                chunk(v1, c0 + 32);
            }
            fence_proxy_async_smem();
            asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
            if (lead) {
                if (MODE == 2)  // the owner's staging slot for this rank (peer memory), rows relative to its shard
                    tma_store_4d(&pm->m[peer_of_tile], out_stage + box, col0 + c0, x0 - peer_of_tile * p.m_local, 0, 0);
                else if (MODE == 0 && p.phases)
                    tma_store_4d(out_map, out_stage + box, col0 + c0, x0, y0, n0);
                else
                    tma_store_4d(tail_slab ? &tmOut32 : &tmOut, out_stage + box, col0 + c0, x0, y0, n0);
                bulk_commit();
                if (HAS_RES && sl + 2 < nslab) {
                    bar_expect_tx(rfull, a_bytes);
                    tma4<false>(res_stage + box, &tmRes, rfull, col0 + c0 + 128, x0, y0, n0);
                }
            }
        }
        if (MODE == 2 && (p.dbg & 1)) {
            if (lead) bulk_wait_read<0>();
        } else if (MODE == 2) {
            // both warpgroups' stores must have COMPLETED (not merely been read out of shared memory) before the
            // owner is told; async-proxy writes are ordered before the generic-proxy flag store by the fence
            if (lead) {
                bulk_wait_all();
                fence_proxy_async_all();
            }
            asm volatile("bar.sync 3, 256;" ::: "memory");
            if (threadIdx.x == 64 && !(p.dbg & 8)) {
                const int ntile = (p.cout + BN - 1) / BN;
                st_release_sys(p.flags[peer_of_tile] + (p.rank * p.shard_tiles + tile_in_shard) * ntile + ctile_f,
                               lean_epoch(p));
            }
        } else if (lead) {
            bulk_wait_read<0>();
        }
    }
    if (threadIdx.x == 64) TR(9);
    tc_fence_before();
    cta_or_cluster_sync<PAIR>();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_any<PAIR>(tmem_acc, uint32_t(p.tmem_cols));
    }
#ifdef SW_TRACE
    if (threadIdx.x == 0) {
        TR(10);
        TRG(12);
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        const unsigned cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        if (cta < unsigned(kTraceCtas)) {
            g_sw_trace[cta * kTraceSlots + 13] = smid;
            g_sw_trace[cta * kTraceSlots + 14] = (unsigned long long)p.bn | ((unsigned long long)p.stages << 16) | (1ull << 24) |
                                                 ((unsigned long long)(PAIR ? 1 : 0) << 32) | (1ull << 34) | (1ull << 36) |
                                                 ((unsigned long long)p.epi_alias << 37);
            g_sw_trace[cta * kTraceSlots + 15] = (unsigned long long)gridDim.x | ((unsigned long long)gridDim.y << 20) |
                                                 ((unsigned long long)gridDim.z << 40);
        }
    }
#endif
}

#define SW_LEAN_KERNEL(NAME, PAIR, RES)                                                                            \
    __global__ void __launch_bounds__(kThreadsLean, 2)                                                                \
        NAME(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmA2,                   \
             const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,                   \
             const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmOut32,               \
             const IGemmArgs p) {                                                                                  \
        igemm_lean_body<PAIR, RES>(tmA1, tmA2, tmB, tmOut, tmRes, tmOut32, p);                                     \
    }
SW_LEAN_KERNEL(k_igemm_lean, false, false)
SW_LEAN_KERNEL(k_igemm_lean_r, false, true)
SW_LEAN_KERNEL(k_igemm2_lean, true, false)
SW_LEAN_KERNEL(k_igemm2_lean_r, true, true)
#undef SW_LEAN_KERNEL

__global__ void __launch_bounds__(kThreadsLean, 2)
    k_igemm_lean_ag(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmOut32,
                    const IGemmArgs p) {
    igemm_lean_body<false, false, 1>(tmA1, tmA1, tmB, tmOut, tmOut, tmOut32, p);
}
__global__ void __launch_bounds__(kThreadsLean, 2)
    k_igemm_lean_rs(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ PeerMaps pm, const IGemmArgs p) {
    igemm_lean_body<false, false, 2>(tmA1, tmA1, tmB, pm.m[0], pm.m[0], pm.m[0], p, &pm);
}

// =================================================================================================
// Persistent variant (mode 0, single-CTA tiles, no split-K): each CTA walks tiles
// blockIdx.x, blockIdx.x + gridDim.x, ... with TWO accumulators in TMEM.  The kernel above pays its
// prologue (barrier init, TMEM allocation, first TMA round trip) and its epilogue once per tile with
// nothing to hide them behind except a second resident CTA; graph-timed sweeps put it 2-3x behind
// cuBLAS on the short-K transformer GEMMs for exactly that reason.  Here the producer warp runs ahead
// into the next tile's operands while the epilogue warps drain the previous accumulator, and the MMA
// warp switches accumulators per tile:
//   warp 0   TMA producer      ring of `stages` operand slots, continuous across tiles
//   warp 1   MMA issuer        waits acc_empty[b], accumulates tile i into TMEM buffer b = i & 1,
//                              commits acc_full[b]
//   warps 2-5 epilogue         wait acc_full[b], drain, arrive acc_empty[b] as soon as the last
//                              tcgen05.ld has landed, then finish the math / stores
// Output goes through 64-column swizzled staging boxes + TMA stores (dedicated shared memory: the ring
// is busy with the next tile), or straight from registers for GEGLU.
__global__ void __launch_bounds__(kThreads)
    k_igemm_p(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmA2,
              const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
              const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmOut32,
              const IGemmArgs p) {
    const int BN = p.bn;
    const int B_STAGE_BYTES = BN * BK * 2;
    const int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
    __shared__ __align__(16) float s_bias[2][256];
    __shared__ __align__(16) float s_csp[2][256];

    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

    const int warp = __shfl_sync(0xffffffffu, int(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int stages = p.stages;
    if (threadIdx.x == 0) {
        TR(0);
        TRG(11);
    }
    const int ring_bytes = stages * STAGE_BYTES;
    const int stage_area = p.tma_epi ? 65536 : 0;  // out[2] + res[2] boxes of 16 KB
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + ring_bytes + stage_area);
    uint64_t* empty = full + stages;
    uint64_t* acc_full = empty + stages;   // [2]
    uint64_t* acc_empty = acc_full + 2;    // [2], 128 arrivals
    uint64_t* res_full = acc_empty + 2;    // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_full + 2);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA1);
        tma_prefetch_desc(&tmA2);
        tma_prefetch_desc(&tmB);
        if (p.tma_epi) {
            tma_prefetch_desc(&tmOut);
            tma_prefetch_desc(&tmRes);
            tma_prefetch_desc(&tmOut32);
        }
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < stages; ++s) {
                mbar_init(&full[s], 1);
                mbar_init(&empty[s], 1);
            }
            for (int b = 0; b < 2; ++b) {
                mbar_init(&acc_full[b], 1);
                mbar_init(&acc_empty[b], 128);
                mbar_init(&res_full[b], 1);
            }
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc_any<false>(tmem_slot, uint32_t(2 * p.tmem_cols));
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    if (threadIdx.x == 0) TR(1);
    pdl_trigger();
    pdl_wait();
    if (threadIdx.x == 0) TR(2);

    const int nkb1 = (p.c1 + BK - 1) / BK;
    const int nkb2 = (p.c2 + BK - 1) / BK;
    const int per_tap = nkb1 + nkb2;
    const int ctot = p.c1 + p.c2;
    const int total_kb = p.kh * p.kw * per_tap;
    const uint32_t a_bytes = uint32_t(p.bw * p.bh * p.bi) * (BK * 2);
    const uint32_t smem0 = smem_u32(smem);
    const uint32_t full0 = smem_u32(full), empty0 = smem_u32(empty);
    const int ptiles = p.ptiles;          // pixel tiles (fastest-varying: concurrent CTAs share a weight tile)
    const int total_tiles = p.total_tiles;

    if (warp == 0) {
        {
            // ---------------- TMA producer (whole warp walks the loop, one elected lane issues)
            const uint32_t tx_bytes = a_bytes + uint32_t(B_STAGE_BYTES);
            uint32_t sa = smem0, fb = full0, eb = empty0;
            int s = 0;
            uint32_t ph = 1;
            if (lane == 0) TR(3);
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const int nt = tile / ptiles;
                int t = tile - nt * ptiles;
                const int tx = t % p.tiles_x;
                t /= p.tiles_x;
                const int ty = t % p.tiles_y;
                const int tn = t / p.tiles_y;
                const int x0 = tx * p.bw, y0 = ty * p.bh, n0 = tn * p.bi;
                const int bcol0 = nt * BN;
                int j = 0, ky = 0, kx = 0, tapbase = 0;
                for (int it = 0; it < total_kb; ++it) {
                    bar_wait(eb, ph);
                    if (elect_one()) {
                        bar_expect_tx(fb, tx_bytes);
                        if (j < nkb1) {
                            tma4<false>(sa, &tmA1, fb, j * BK, x0 * p.sx + kx - p.pad_w, y0 * p.sx + ky - p.pad_y1, n0);
                            tma_b<false>(sa + A_STAGE_BYTES, &tmB, fb, p.w_kmajor, tapbase + j * BK, it, bcol0);
                        } else {
                            tma4<false>(sa, &tmA2, fb, (j - nkb1) * BK, x0 * p.sx + kx - p.pad_w, y0 * p.sx + ky - p.pad, n0);
                            tma_b<false>(sa + A_STAGE_BYTES, &tmB, fb, p.w_kmajor, tapbase + p.c1 + (j - nkb1) * BK, it, bcol0);
                        }
                    }
                    if (++j == per_tap) {
                        j = 0;
                        tapbase += ctot;
                        if (++kx == p.kw) {
                            kx = 0;
                            ++ky;
                        }
                    }
                    if (++s == stages) {
                        s = 0;
                        ph ^= 1;
                        sa = smem0;
                        fb = full0;
                        eb = empty0;
                    } else {
                        sa += STAGE_BYTES;
                        fb += 8;
                        eb += 8;
                    }
                }
            }
            if (lane == 0) TR(4);
        }
    } else if (warp == 1) {
        {
            // ---------------- MMA issuer (whole warp, one elected lane issues)
            const uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
            const uint32_t a_lo0 = ((smem0 & 0x3FFFFu) >> 4) | (1u << 16);
            const uint32_t b_off = uint32_t(A_STAGE_BYTES) >> 4;
            const uint32_t stage_step = uint32_t(STAGE_BYTES) >> 4;
            const int ks1 = (p.c1 - (nkb1 - 1) * BK + 15) / 16;
            const int ks2 = nkb2 ? (p.c2 - (nkb2 - 1) * BK + 15) / 16 : 4;
            const int last1 = nkb1 - 1, last2 = per_tap - 1;
            uint32_t a_lo = a_lo0, fb = full0, eb = empty0;
            int s = 0;
            uint32_t ph = 0;
            const uint32_t idesc = p.idesc;
            const uint32_t accf0 = smem_u32(acc_full), acce0 = smem_u32(acc_empty);
            int i = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++i) {
                const uint32_t ab = uint32_t(i) & 1u;
                const uint32_t tacc = tmem_base + ab * uint32_t(p.tmem_cols);
                bar_wait(acce0 + 8 * ab, ((uint32_t(i) >> 1) & 1u) ^ 1u);  // epilogue has drained this buffer
                tc_fence_after();
                int j = 0;
                uint32_t acc = 0;
                for (int it = 0; it < total_kb; ++it) {
                    bar_wait(fb, ph);
                    tc_fence_after();
#ifdef SW_TRACE
                    if (it == 0 && i == 0 && lane == 0) TR(5);
#endif
                    int ksteps = 4;
                    if (j == last1) ksteps = ks1 < 4 ? ks1 : 4;
                    if (nkb2 && j == last2) ksteps = ks2 < 4 ? ks2 : 4;
                    if (elect_one()) {
                        mma_f16<false>(tacc, a_lo, a_lo + b_off, desc_hi, idesc, acc);
                        if (ksteps > 1) mma_f16<false>(tacc, a_lo + 2, a_lo + b_off + 2, desc_hi, idesc, 1u);
                        if (ksteps > 2) mma_f16<false>(tacc, a_lo + 4, a_lo + b_off + 4, desc_hi, idesc, 1u);
                        if (ksteps > 3) mma_f16<false>(tacc, a_lo + 6, a_lo + b_off + 6, desc_hi, idesc, 1u);
                        mma_commit<false>(eb);
                    }
                    acc = 1u;
                    if (++j == per_tap) j = 0;
                    if (++s == stages) {
                        s = 0;
                        ph ^= 1;
                        a_lo = a_lo0;
                        fb = full0;
                        eb = empty0;
                    } else {
                        a_lo += stage_step;
                        fb += 8;
                        eb += 8;
                    }
                }
                if (elect_one()) mma_commit<false>(accf0 + 8 * ab);
            }
            if (lane == 0) TR(6);
        }
    } else {
        // ---------------- epilogue: one thread per accumulator row
        const int q = warp & 3;
        const int r = q * 32 + lane;
        const int xl = r % p.bw;
        const int r2 = r / p.bw;
        const int yl = r2 % p.bh;
        const int nl = r2 / p.bh;
        const uint32_t lane_off = uint32_t(q * 32) << 16;
        const bool lead = threadIdx.x == 64;
        const bool geglu = p.act == SW_ACT_GEGLU;
        const bool has_res = p.residual != nullptr;
        const uint32_t out_stage = smem0 + uint32_t(ring_bytes);
        const uint32_t res_stage = out_stage + 32768u;
        const uint32_t rfull0 = smem_u32(res_full);
        const uint32_t accf0 = smem_u32(acc_full);
        const uint32_t row_off = uint32_t(r) * 128u, swz = uint32_t(r) & 7u;
        uint32_t rph[2] = {0u, 0u};  // parity of the next completion of each residual box
        int osl = 0;                 // output slabs staged so far (box = osl & 1), across tiles
        int i = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++i) {
            const int nt = tile / ptiles;
            int t = tile - nt * ptiles;
            const int tx = t % p.tiles_x;
            t /= p.tiles_x;
            const int ty = t % p.tiles_y;
            const int tn = t / p.tiles_y;
            const int x0 = tx * p.bw, y0 = ty * p.bh, n0 = tn * p.bi;
            const int col0 = nt * BN;
            const int x = x0 + xl, y = y0 + yl, ni = n0 + nl;
            const bool row_ok = (nl < p.bi) && x < p.w && y < p.h && ni < p.n;
            const size_t m = (size_t(ni) * p.h + y) * p.w + x;
            const uint32_t ab = uint32_t(i) & 1u;
            float* sb = s_bias[ab];
            float* scs = s_csp[ab];
            const bool fold_rb = p.row_bias != nullptr && p.bi == 1;
            for (int k = threadIdx.x - 64; k < BN; k += 128) {
                const int col = col0 + k;
                float b = 0.0f;
                if (col < p.cout) {
                    if (p.bias) b = __half2float(p.bias[col]);
                    if (fold_rb && n0 < p.n) b += __half2float(p.row_bias[size_t(n0) * p.row_bias_ld + col]);
                }
                sb[k] = b;
                scs[k] = (p.col_sums && col < p.cout) ? p.col_sums[col] : 0.0f;
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            float ln_mean = 0.0f, ln_rstd = 1.0f;
            if (p.row_stats_in && row_ok) {
                const size_t rows_total = size_t(p.n) * p.h * p.w;
                float su = 0.0f, sq = 0.0f;
                for (int ti = 0; ti < p.stats_tiles; ++ti) {
                    const float2 pr = *reinterpret_cast<const float2*>(p.row_stats_in + (size_t(ti) * rows_total + m) * 2);
                    su += pr.x;
                    sq += pr.y;
                }
                const float inv_w = 1.0f / float(p.ln_width);
                ln_mean = su * inv_w;
                ln_rstd = rsqrtf(fmaxf(sq * inv_w - ln_mean * ln_mean, 0.0f) + p.ln_eps);
            }
            const uint32_t trow = tmem_base + ab * uint32_t(p.tmem_cols) + lane_off;
            const uint32_t acce = smem_u32(&acc_empty[ab]);

            if (geglu) {
                // values in accumulator columns [0, BN/2), gates in [BN/2, BN); direct 16-byte stores
                const int HALF = BN / 2;
                const int ocol0 = nt * HALF;
                const int ncols_out = p.cout / 2;
                if (lane == 0) bar_wait(accf0 + 8 * ab, (uint32_t(i) >> 1) & 1u);
                __syncwarp();
                tc_fence_after();
                uint32_t vn[16], gn[16];
                tmem_ld16(trow, vn);
                tmem_ld16(trow + HALF, gn);
#pragma unroll 1
                for (int c = 0; c < HALF; c += 16) {
                    uint32_t v[16], g[16];
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] = vn[j], g[j] = gn[j];
                    if (c + 16 < HALF) {
                        tmem_ld16(trow + c + 16, vn);
                        tmem_ld16(trow + HALF + c + 16, gn);
                    } else {  // the accumulator is in registers: hand it back to the MMA warp
                        tc_fence_before();
                        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(acce) : "memory");
                    }
                    if (row_ok && ocol0 + c < ncols_out) {
                        uint32_t o[8];
#pragma unroll
                        for (int j = 0; j < 16; j += 2) {
                            const float a0 = fmaf(ln_rstd, fmaf(-ln_mean, scs[c + j], __uint_as_float(v[j])), sb[c + j]);
                            const float a1 = fmaf(ln_rstd, fmaf(-ln_mean, scs[c + j + 1], __uint_as_float(v[j + 1])), sb[c + j + 1]);
                            const float g0 = fmaf(ln_rstd, fmaf(-ln_mean, scs[HALF + c + j], __uint_as_float(g[j])), sb[HALF + c + j]);
                            const float g1 = fmaf(ln_rstd, fmaf(-ln_mean, scs[HALF + c + j + 1], __uint_as_float(g[j + 1])),
                                                  sb[HALF + c + j + 1]);
                            o[j / 2] = pack_h2(a0 * gelu_f(g0), a1 * gelu_f(g1));
                        }
                        uint4* dst = reinterpret_cast<uint4*>(p.out + m * p.ldc + ocol0 + c);
                        dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
                        dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
                    }
                }
                continue;
            }

            if (!p.tma_epi) {
                // ---- register -> global epilogue (two persistent CTAs per SM, no staging boxes): TMEM
                // chunk c+32 and its residual are in flight while chunk c is converted and stored
                int nv = p.cout - col0;
                if (nv > BN) nv = BN;
                const int ce = (nv + 31) & ~31;
                const bool pl = p.act == SW_ACT_NONE && p.out_scale == 1.0f && (!p.row_bias || fold_rb);
                __half* const out_row = p.out + m * p.ldc + col0;
                const __half* const res_row = has_res ? p.residual + m * p.ldr + col0 : nullptr;
                auto fetch = [&](int c, uint4 (&rs)[4]) {
                    if (!has_res || !row_ok) return;
#pragma unroll
                    for (int g8 = 0; g8 < 4; ++g8)
                        if (col0 + c + g8 * 8 < p.cout) rs[g8] = *reinterpret_cast<const uint4*>(res_row + c + g8 * 8);
                };
                if (lane == 0) bar_wait(accf0 + 8 * ab, (uint32_t(i) >> 1) & 1u);
                __syncwarp();
                tc_fence_after();
                uint32_t vn[32];
                uint4 rn[4];
                tmem_ld32(trow, vn);
                fetch(0, rn);
#pragma unroll 1
                for (int c = 0; c < ce; c += 32) {
                    uint32_t v[32];
                    uint4 rs[4];
                    tmem_ld_wait_dep(vn);
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = vn[j];
#pragma unroll
                    for (int j = 0; j < 4; ++j) rs[j] = rn[j];
                    if (c + 32 < ce) {
                        tmem_ld32(trow + c + 32, vn);
                        fetch(c + 32, rn);
                    } else {
                        tc_fence_before();
                        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(acce) : "memory");
                    }
                    if (!row_ok) continue;
#pragma unroll
                    for (int g8 = 0; g8 < 4; ++g8) {
                        const int c8 = c + g8 * 8;
                        if (col0 + c8 >= p.cout) break;
                        const float4 b0 = *reinterpret_cast<const float4*>(&sb[c8]);
                        const float4 b1 = *reinterpret_cast<const float4*>(&sb[c8 + 4]);
                        float f[8] = {__uint_as_float(v[g8 * 8 + 0]) + b0.x, __uint_as_float(v[g8 * 8 + 1]) + b0.y,
                                      __uint_as_float(v[g8 * 8 + 2]) + b0.z, __uint_as_float(v[g8 * 8 + 3]) + b0.w,
                                      __uint_as_float(v[g8 * 8 + 4]) + b1.x, __uint_as_float(v[g8 * 8 + 5]) + b1.y,
                                      __uint_as_float(v[g8 * 8 + 6]) + b1.z, __uint_as_float(v[g8 * 8 + 7]) + b1.w};
                        if (!pl) {
                            if (p.row_bias && !fold_rb) {
                                const uint4 b = *reinterpret_cast<const uint4*>(p.row_bias + size_t(ni) * p.row_bias_ld +
                                                                                col0 + c8);
                                const float2 t0 = unpack_h2(b.x), t1 = unpack_h2(b.y), t2 = unpack_h2(b.z),
                                             t3 = unpack_h2(b.w);
                                f[0] += t0.x, f[1] += t0.y, f[2] += t1.x, f[3] += t1.y;
                                f[4] += t2.x, f[5] += t2.y, f[6] += t3.x, f[7] += t3.y;
                            }
                            if (p.act == SW_ACT_SILU) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) f[j] = silu_f(f[j]);
                            } else if (p.act == SW_ACT_GELU) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) f[j] = gelu_f(f[j]);
                            }
                            if (p.out_scale != 1.0f) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) f[j] *= p.out_scale;
                            }
                        }
                        if (has_res) {
                            const float2 r0 = unpack_h2(rs[g8].x), r1 = unpack_h2(rs[g8].y), r2 = unpack_h2(rs[g8].z),
                                         r3 = unpack_h2(rs[g8].w);
                            f[0] += r0.x, f[1] += r0.y, f[2] += r1.x, f[3] += r1.y;
                            f[4] += r2.x, f[5] += r2.y, f[6] += r3.x, f[7] += r3.y;
                        }
                        *reinterpret_cast<uint4*>(out_row + c8) =
                            make_uint4(pack_h2(f[0], f[1]), pack_h2(f[2], f[3]), pack_h2(f[4], f[5]), pack_h2(f[6], f[7]));
                    }
                }
                continue;
            }

            // ---- 64-column slabs staged in swizzled shared memory, written by TMA stores
            int nvalid = p.cout - col0;
            if (nvalid > BN) nvalid = BN;
            const int cend = (nvalid + 31) & ~31;
            const int nslab = (nvalid + 63) >> 6;
            const bool tail32 = (cend & 63) == 32;
            const bool plain = p.act == SW_ACT_NONE && p.out_scale == 1.0f && (!p.row_bias || fold_rb);
            if (has_res && lead) {
                for (int sl = 0; sl < 2 && sl < nslab; ++sl) {
                    bar_expect_tx(rfull0 + 8 * sl, a_bytes);
                    tma4<false>(res_stage + sl * 16384, &tmRes, rfull0 + 8 * sl, col0 + sl * 64, x0, y0, n0);
                }
            }
            if (lane == 0) bar_wait(accf0 + 8 * ab, (uint32_t(i) >> 1) & 1u);
            __syncwarp();
            tc_fence_after();
            uint32_t vn[32];
            tmem_ld32(trow, vn);
#pragma unroll 1
            for (int c = 0; c < cend; c += 32) {
                const int sl = c >> 6;
                const uint32_t obox = out_stage + uint32_t(osl & 1) * 16384u;
                const uint32_t rbox = res_stage + uint32_t(sl & 1) * 16384u;
                if ((c & 63) == 0) {
                    if (osl >= 2) {  // the store issued two slabs ago must have finished reading this box
                        if (lead) bulk_wait_read<1>();
                        asm volatile("bar.sync 1, 128;" ::: "memory");
                    }
                    if (has_res) {
                        bar_wait(rfull0 + 8 * (sl & 1), rph[sl & 1]);
                        rph[sl & 1] ^= 1u;
                    }
                }
                uint32_t v[32];
                tmem_ld_wait_dep(vn);
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = vn[j];
                if (c + 32 < cend) {
                    tmem_ld32(trow + c + 32, vn);
                } else {  // whole accumulator read: the MMA warp may start the tile after next in it
                    tc_fence_before();
                    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(acce) : "memory");
                }
                const bool tail_slab = tail32 && sl == nslab - 1;
#pragma unroll
                for (int g8 = 0; g8 < 4; ++g8) {
                    const int c8 = c + g8 * 8;
                    const float4 b0 = *reinterpret_cast<const float4*>(&sb[c8]);
                    const float4 b1 = *reinterpret_cast<const float4*>(&sb[c8 + 4]);
                    float f[8] = {__uint_as_float(v[g8 * 8 + 0]) + b0.x, __uint_as_float(v[g8 * 8 + 1]) + b0.y,
                                  __uint_as_float(v[g8 * 8 + 2]) + b0.z, __uint_as_float(v[g8 * 8 + 3]) + b0.w,
                                  __uint_as_float(v[g8 * 8 + 4]) + b1.x, __uint_as_float(v[g8 * 8 + 5]) + b1.y,
                                  __uint_as_float(v[g8 * 8 + 6]) + b1.z, __uint_as_float(v[g8 * 8 + 7]) + b1.w};
                    if (!plain) {
                        if (p.row_bias && !fold_rb && row_ok) {
                            const uint4 b =
                                *reinterpret_cast<const uint4*>(p.row_bias + size_t(ni) * p.row_bias_ld + col0 + c8);
                            const float2 t0 = unpack_h2(b.x), t1 = unpack_h2(b.y), t2 = unpack_h2(b.z), t3 = unpack_h2(b.w);
                            f[0] += t0.x, f[1] += t0.y, f[2] += t1.x, f[3] += t1.y;
                            f[4] += t2.x, f[5] += t2.y, f[6] += t3.x, f[7] += t3.y;
                        }
                        if (p.act == SW_ACT_SILU) {
#pragma unroll
                            for (int j = 0; j < 8; ++j) f[j] = silu_f(f[j]);
                        } else if (p.act == SW_ACT_GELU) {
#pragma unroll
                            for (int j = 0; j < 8; ++j) f[j] = gelu_f(f[j]);
                        }
                        if (p.out_scale != 1.0f) {
#pragma unroll
                            for (int j = 0; j < 8; ++j) f[j] *= p.out_scale;
                        }
                    }
                    const uint32_t off = row_off + (((uint32_t(c8 >> 3) & 7u) ^ swz) << 4);
                    if (has_res) {
                        const uint4 rv = lds128(rbox + off);
                        const float2 r0 = unpack_h2(rv.x), r1 = unpack_h2(rv.y), r2 = unpack_h2(rv.z),
                                     r3 = unpack_h2(rv.w);
                        f[0] += r0.x, f[1] += r0.y, f[2] += r1.x, f[3] += r1.y;
                        f[4] += r2.x, f[5] += r2.y, f[6] += r3.x, f[7] += r3.y;
                    }
                    const uint32_t dst = tail_slab ? obox + uint32_t(r) * 64u + ((uint32_t(c8 >> 3) & 3u) << 4) : obox + off;
                    sts128(dst, make_uint4(pack_h2(f[0], f[1]), pack_h2(f[2], f[3]), pack_h2(f[4], f[5]),
                                           pack_h2(f[6], f[7])));
                }
                if ((c & 63) == 32 || c + 32 >= cend) {
                    fence_proxy_async_smem();
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    if (lead) {
                        tma_store_4d(tail_slab ? &tmOut32 : &tmOut, obox, col0 + sl * 64, x0, y0, n0);
                        bulk_commit();
                        if (has_res && sl + 2 < nslab) {
                            bar_expect_tx(rfull0 + 8 * (sl & 1), a_bytes);
                            tma4<false>(rbox, &tmRes, rfull0 + 8 * (sl & 1), col0 + (sl + 2) * 64, x0, y0, n0);
                        }
                    }
                    ++osl;
                }
            }
        }
        if (lead) bulk_wait_read<0>();
        if (lead) TR(9);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_any<false>(tmem_base, uint32_t(2 * p.tmem_cols));
    }
#ifdef SW_TRACE
    if (threadIdx.x == 0) {
        TR(10);
        TRG(12);
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        if (blockIdx.x < unsigned(kTraceCtas)) {
            g_sw_trace[blockIdx.x * kTraceSlots + 13] = smid;
            g_sw_trace[blockIdx.x * kTraceSlots + 14] = (unsigned long long)p.bn | ((unsigned long long)p.stages << 16) |
                                                        (1ull << 24) | (1ull << 35) | ((unsigned long long)p.tma_epi << 34);
            g_sw_trace[blockIdx.x * kTraceSlots + 15] = (unsigned long long)gridDim.x | (1ull << 20) | (1ull << 40);
        }
    }
#endif
}

// Choose the pixel box (bw x bh x bi, <= 128 pixels) that wastes the fewest MMA rows.
void pick_box(int n, int h, int w, int* bw, int* bh, int* bi) {
    double best = -1.0;
    int b0 = 1, b1 = 1, b2 = 1;
    for (int cw = 1; cw <= w && cw <= 128; ++cw) {
        const int tiles_x = (w + cw - 1) / cw;
        const int max_h = 128 / cw;
        for (int ch = 1; ch <= h && ch <= max_h; ++ch) {
            int ci = 1;
            if (cw == w && ch == h) {
                ci = 128 / (cw * ch);
                if (ci > n) ci = n;
                if (ci < 1) ci = 1;
            }
            const int tiles_y = (h + ch - 1) / ch;
            const int tiles_n = (n + ci - 1) / ci;
            const double useful = double(n) * h * w;
            const double spent = double(tiles_x) * tiles_y * tiles_n * 128.0;
            double score = useful / spent;
            // prefer wide boxes (longer contiguous TMA rows) on ties
            score += 1e-6 * cw;
            if (score > best) {
                best = score;
                b0 = cw;
                b1 = ch;
                b2 = ci;
            }
        }
    }
    *bw = b0;
    *bh = b1;
    *bi = b2;
}

// dynamic shared memory: operand ring (at least the four 16 KB epilogue staging boxes) + alignment + barriers
int smem_bytes(const IGemmArgs& a, bool pair) {
    const int stage_bytes = A_STAGE_BYTES + (pair ? a.bn / 2 : a.bn) * BK * 2;
    int ring = a.stages * stage_bytes;
    if (a.tma_epi && ring < 65536) ring = 65536;
    return ring + 1024 + 256 + (a.cluster_k ? 128 * (a.bn / 2) * 4 : 0);
}

int launch(const IGemmArgs& a, const CUtensorMap& tA1, const CUtensorMap& tA2, const CUtensorMap& tB,
           const CUtensorMap& tOut, const CUtensorMap& tRes, const CUtensorMap& tOut32, int grid_x, bool pair,
           cudaStream_t st) {
    const int smem = smem_bytes(a, pair);
    static bool configured[64] = {false};
    int dev = 0;
    SW_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 64 && !configured[dev]) {
        SW_CUDA_OK(cudaFuncSetAttribute(k_igemm, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
        SW_CUDA_OK(cudaFuncSetAttribute(k_igemm2, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
        configured[dev] = true;
    }
    const int ntile = (a.cout + a.bn - 1) / a.bn;
    if (!pair) {
        if (a.cluster_k) {
            SW_CUDA_OK(swh::launch_pdl_z(k_igemm, dim3(grid_x, ntile, 2), dim3(kThreads), size_t(smem), st, 2, tA1, tA2,
                                         tB, tOut, tRes, tOut32, a));
            return 0;
        }
        SW_CUDA_OK(swh::launch_pdl(k_igemm, dim3(grid_x, ntile, a.splits), dim3(kThreads), size_t(smem), st, 1, tA1,
                                   tA2, tB, tOut, tRes, tOut32, a));
        return 0;
    }
    SW_CUDA_OK(swh::launch_pdl(k_igemm2, dim3((grid_x + 1) / 2 * 2, ntile, a.splits), dim3(kThreads), size_t(smem),
                               st, 2, tA1, tA2, tB, tOut, tRes, tOut32, a));
    return 0;
}

constexpr int kMaxSplitTiles = 4096;

}  // namespace

static int igemm_impl(const sw_igemm_params* p, const sw_peers* peers, int mode, int m_local,
                      unsigned int* ticket, void* stream) {
    if (!p || !p->wgt || (mode != 1 && !p->a1) || (mode != 2 && !p->out))
        return swh::fail(-1, "sw_igemm: null argument");
    if (mode != 0) {
        if (!peers || peers->world < 1 || peers->world > 8) return swh::fail(-1, "sw_igemm: bad peer description");
        if (p->kh != 1 || p->kw != 1 || p->n != 1 || p->h != 1 || p->c2 != 0 || p->act == SW_ACT_GEGLU)
            return swh::fail(-1, "sw_igemm: fused collectives are defined for plain GEMMs");
        if (m_local <= 0 || m_local % 128 || p->w != m_local * peers->world)
            return swh::fail(-1, "sw_igemm: m_local=%d must be a multiple of 128 and w == world*m_local", m_local);
        if (mode == 2 && (p->bias || p->row_bias || p->residual || p->act != SW_ACT_NONE || !ticket))
            return swh::fail(-1, "sw_igemm_rs: the push epilogue carries raw partial sums only");
    }
    if (p->c1 <= 0 || p->c1 % 8 || p->c2 % 8 || p->c2 < 0 || (p->c2 > 0 && !p->a2))
        return swh::fail(-1, "sw_igemm: channel counts must be positive multiples of 8 (c1=%d c2=%d)",
                         p->c1, p->c2);
    if (p->cout <= 0 || p->cout % 8) return swh::fail(-1, "sw_igemm: cout=%d not a multiple of 8", p->cout);
    if (p->n <= 0 || p->h <= 0 || p->w <= 0 || p->kh <= 0 || p->kw <= 0)
        return swh::fail(-1, "sw_igemm: bad shape");
    if (p->a_h_total < 0 || (p->a_h_total > 0 && (mode != 0 || p->a_y_off < 0 ||
                                                  p->a_y_off + (p->in_h > 0 ? p->in_h : p->h) > p->a_h_total)))
        return swh::fail(-1, "sw_igemm: a_h_total=%d / a_y_off=%d do not contain the %d output rows", p->a_h_total,
                         p->a_y_off, p->h);
    const bool ln_fused = p->row_stats_out != nullptr || p->row_stats_in != nullptr;
    if (ln_fused) {
        if (mode != 0) return swh::fail(-1, "sw_igemm: folded LayerNorm is defined for local GEMMs");
        if (p->row_stats_out && p->act == SW_ACT_GEGLU)
            return swh::fail(-1, "sw_igemm: row statistics cannot be emitted from a GEGLU epilogue");
        if (p->row_stats_in && (!p->col_sums || p->stats_tiles <= 0 || p->ln_width <= 0))
            return swh::fail(-1, "sw_igemm: row_stats_in needs col_sums, stats_tiles and ln_width");
    }
    if (p->gn_stats) return swh::fail(-1, "sw_igemm: gn_stats epilogue not available in this build");
    if (p->bf16) return swh::fail(-1, "sw_igemm: bf16 operands not enabled in this build (fp16 only)");
    if ((p->a_ld && (p->a_ld % 8 || p->a_ld < p->c1)) || (p->w_ld % 8))
        return swh::fail(-1, "sw_igemm: a_ld/w_ld must be multiples of 8 (a_ld >= c1)");
    if (p->ldc % 8 || (p->residual && p->ldr % 8) || (p->row_bias && p->row_bias_ld % 8))
        return swh::fail(-1, "sw_igemm: leading dimensions must be multiples of 8");

    const int cstride = p->stride > 1 ? p->stride : 1;
    const int in_h = p->in_h > 0 ? p->in_h : p->h, in_w = p->in_w > 0 ? p->in_w : p->w;
    const int osx = p->out_sx > 1 ? p->out_sx : 1, osy = p->out_sy > 1 ? p->out_sy : 1;
    const bool out_strided = osx > 1 || osy > 1;
    if (cstride > 2 || (cstride == 1 && (in_h != p->h || in_w != p->w) && p->a_h_total == 0))
        return swh::fail(-1, "sw_igemm: stride must be 1 or 2 (in_h / in_w only with stride 2)");
    if ((cstride > 1 || out_strided) && (mode != 0 || ln_fused || p->a_ld))
        return swh::fail(-1, "sw_igemm: strided convolutions / strided outputs are plain local convolutions");
    if (out_strided && (p->residual || p->act == SW_ACT_GEGLU))
        return swh::fail(-1, "sw_igemm: a strided output takes no residual and no GEGLU");
    const bool phases = p->out_phases != 0;
    if (phases && (p->kh != 2 || p->kw != 2 || p->pad != 1 || p->pad_w1 != 0 || osx != 2 || osy != 2 || p->c2 != 0 ||
                   p->cout % 64 || p->act != SW_ACT_NONE || p->row_bias || cstride != 1 || p->w_kmajor || p->w_ld ||
                   (p->out_scale != 0.0f && p->out_scale != 1.0f)))
        return swh::fail(-1, "sw_igemm: out_phases is the folded 2x upsample: 2x2 taps, pad 1, out strides 2, bias only, cout %% 64 == 0");
    IGemmArgs a;
    memset(&a, 0, sizeof(a));
    a.sx = cstride;
    a.phases = phases ? 1 : 0;
    a.c1 = p->c1;
    a.c2 = p->c2;
    a.n = p->n;
    a.h = p->h;
    a.w = p->w;
    a.kh = p->kh;
    a.kw = p->kw;
    a.pad = p->pad;
    a.pad_w = p->pad_w1 > 0 ? p->pad_w1 - 1 : p->pad;
    a.row_stats_out = p->row_stats_out;
    a.row_stats_in = p->row_stats_in;
    a.col_sums = p->col_sums;
    a.stats_tiles = p->stats_tiles;
    a.ln_width = p->ln_width;
    a.ln_eps = p->ln_eps;
    a.pad_y1 = p->pad - p->a_y_off;  // a1 may be a taller tensor (halo rows / a window of a full image)
    a.cout = p->cout;
    pick_box(p->n, p->h, p->w, &a.bw, &a.bh, &a.bi);
    a.tiles_x = (p->w + a.bw - 1) / a.bw;
    a.tiles_y = (p->h + a.bh - 1) / a.bh;
    const int tiles_n = (p->n + a.bi - 1) / a.bi;
    const int grid_x = a.tiles_x * a.tiles_y * tiles_n;
    a.bias = static_cast<const __half*>(p->bias);
    a.row_bias = static_cast<const __half*>(p->row_bias);
    a.row_bias_ld = p->row_bias_ld;
    a.residual = static_cast<const __half*>(p->residual);
    a.ldr = p->ldr;
    a.out = static_cast<__half*>(p->out);
    a.ldc = p->ldc;
    a.act = p->act;
    a.out_scale = p->out_scale == 0.0f ? 1.0f : p->out_scale;
    a.bf16 = p->bf16;
    a.mode = mode;
    if (mode != 0) {
        a.world = peers->world;
        a.rank = peers->rank;
        a.epoch = peers->epoch;
        a.m_local = m_local;
        a.rs_ticket = ticket;
        for (int d = 0; d < peers->world; ++d) {
            a.flags[d] = peers->flags[d];
            if (mode == 2) a.rs_out[d] = static_cast<__half*>(peers->peer[d]);
        }
        if (mode == 2) a.out = a.rs_out[peers->rank];
    }

    const int ktotal = p->kh * p->kw * (p->c1 + p->c2);
    const int kblocks = p->kh * p->kw * ((p->c1 + 63) / 64 + (p->c2 + 63) / 64);
    const int sms = swh::sm_count();

    // ---- N tile.  The kernel is bound by L2 -> SM operand traffic, so wide tiles win: the
    // activation tile (re-read once per filter tap) is amortised over more output channels.
    // Layers with few pixel tiles are weight-streaming bound instead and want many CTAs.
    int bn = p->tile_n;
    if (p->act == SW_ACT_GEGLU) {
        // value / gate halves are interleaved per tile by the weight packer: tile_n names that granule
        bn = p->tile_n == 256 ? 256 : 128;
        if (p->cout % bn) return swh::fail(-1, "sw_igemm: GEGLU needs cout %% tile_n == 0");
    } else if (bn == 0) {
        // graph-timed sweeps (scripts/bench_gemm_graph.py, bench_stream.py): 128 columns is within a few
        // percent of the best everywhere except (a) layers big enough to fill the chip several times over
        // with 128-wide tiles, which want the widest tile (least operand re-fetch per flop), and (b) grids
        // so small that even split-K cannot spread them, which want many narrow tiles
        const long ctas128 = long(grid_x) * ((p->cout + 127) / 128) * (p->out_phases ? 4 : 1);
        if (2 * ctas128 >= 5L * sms) {
            const int ntiles = (p->cout + 255) / 256;
            bn = ((p->cout + ntiles - 1) / ntiles + 31) / 32 * 32;
        } else if (p->cout <= 64) {
            bn = p->cout <= 32 ? 32 : 64;
        } else if (2 * ctas128 <= sms && kblocks < 16) {
            bn = 64;
        } else {
            bn = p->cout < 128 ? (p->cout + 31) / 32 * 32 : 128;
        }
    }
    if (bn < 32 || bn > 256 || bn % 32) return swh::fail(-1, "sw_igemm: tile_n=%d must be a multiple of 32 in [32,256]", bn);
    // phase mode has no 32-column tail map (its slot carries a phase's output map): whole 64-column slabs only
    if (phases && bn % 64) bn = (bn + 63) / 64 * 64;
    const int tiles_ncol = (p->cout + bn - 1) / bn;
    a.bn = bn;
    a.tmem_cols = bn <= 32 ? 32 : bn <= 64 ? 64 : bn <= 128 ? 128 : 256;

    // ---- CTA pairs: with >= 2 pixel tiles per weight tile the layer runs as 256 x BN tiles on
    // tcgen05 cta_group::2 (halves every SM's shared-memory reads of the weight operand).
    // measured: pays off for wide tiles with many pixel tiles (fused collectives: shards must hold whole pairs)
    bool pair = grid_x >= 32 && bn >= 192 && (mode == 0 || m_local % 256 == 0);
    if (p->cta_pair == 1) pair = false;
    if (p->cta_pair == 2) pair = grid_x >= 2 && bn >= 64 && mode == 0;
    if (ln_fused) pair = false;  // folded LayerNorm lives in the single-CTA epilogues only

    // ---- pipeline depth: two CTAs per SM (epilogue of one overlaps the main loop of the other)
    // when three stages fit in ~110 KB, else one CTA with a deeper ring.
    const int stage_bytes = A_STAGE_BYTES + (pair ? bn / 2 : bn) * BK * 2;
    int stages = (3 * stage_bytes <= 110 * 1024) ? 3 : 4;
    if (stage_bytes <= 24 * 1024) stages = 4;
    // one or two pixel tiles: pure weight streaming, latency-bound -> deepest ring that fits (one CTA per SM)
    if (grid_x <= 2 && !pair) stages = stage_bytes <= 24 * 1024 ? 8 : (200 * 1024) / stage_bytes;
    {
        static int env_forced = -1;  // tuning hook: SW_IGEMM_STAGES overrides the ring depth
        if (env_forced < 0) {
            const char* e = getenv("SW_IGEMM_STAGES");
            env_forced = e ? atoi(e) : 0;
        }
        int forced = env_forced;
        if (p->stages > 0) forced = p->stages;
        if (forced > 0) {
            stages = forced;
            while (stages > 2 && stages * stage_bytes > 220 * 1024) --stages;
        }
    }
    if (kblocks < stages) stages = kblocks;
    a.stages = stages;
    a.idesc = sw::umma_idesc(pair ? 256 : 128, bn, 0, 0, 0);

    // ---- split-K.  A CTA ingests operands from L2 at a bounded rate (~40 B/clk/SM measured), so a
    // layer finishes when its busiest SM has pulled its share: with fewer CTAs than SMs, several
    // CTAs share each tile's K range (deterministic reduction by the last arrival).
    a.splits = 1;
    a.kb_per_split = kblocks;
    const long ctas = long(pair ? (grid_x + 1) / 2 * 2 : grid_x) * tiles_ncol * (phases ? 4 : 1);
    // ---- split-K over a cluster of two (DSMEM exchange, see the epilogue): small GEMMs whose tiles do not fill
    // the chip, so that every SM streams half of its activation slab
    {
        static int env_ck = -1;  // SW_IGEMM_CLUSTERK: 0 never, 1 when requested (default), 2 whenever eligible
        if (env_ck < 0) {
            const char* e = getenv("SW_IGEMM_CLUSTERK");
            env_ck = e ? atoi(e) : 1;
        }
        const bool eligible = mode == 0 && !pair && !ln_fused && !out_strided && p->act != SW_ACT_GEGLU && bn % 64 == 0 && kblocks >= 4 &&
                              2 * ctas <= 2L * sms;
        const bool want = env_ck == 2 || (env_ck == 1 && p->cluster_k == 2);
        if (eligible && want && p->cluster_k != 1) {
            a.cluster_k = 1;
            a.splits = 2;
            a.kb_per_split = (kblocks + 1) / 2;
            if (a.kb_per_split < a.stages) a.stages = a.kb_per_split;
            while (a.stages > 2 && a.stages * stage_bytes + 128 * (bn / 2) * 4 > 220 * 1024) --a.stages;
        }
    }
    const int min_kb = p->splitk_min_kb > 0 ? p->splitk_min_kb : 16;
    if (mode == 0 && !a.cluster_k && !ln_fused && !out_strided && p->splitk_scratch && p->act != SW_ACT_GEGLU && 2 * ctas <= 3L * sms / 2 && kblocks >= 2 * min_kb &&
        ctas <= kMaxSplitTiles) {
        const int slots = (2 * (stages * stage_bytes + 2048) <= 226 * 1024) ? 2 * sms : sms;
        int want = int(slots / ctas);
        int max_by_k = kblocks / min_kb;
        int splits = want < max_by_k ? want : max_by_k;
        if (splits > 32) splits = 32;
        const long long per_tile = (long long)(bn / 32) * 4096 * sizeof(float);
        const long long avail = p->splitk_scratch_bytes - (long long)kMaxSplitTiles * 8;
        while (splits > 1 && per_tile * splits * ctas > avail) --splits;
        if (splits > 1) {
            a.kb_per_split = (kblocks + splits - 1) / splits;
            a.splits = (kblocks + a.kb_per_split - 1) / a.kb_per_split;
            a.tickets = static_cast<unsigned int*>(p->splitk_scratch);
            a.done = a.tickets + kMaxSplitTiles;
            a.partials = reinterpret_cast<float*>(static_cast<char*>(p->splitk_scratch) + kMaxSplitTiles * 8);
            if (a.kb_per_split < a.stages) a.stages = a.kb_per_split;
        }
    }

    CUtensorMap tA1, tA2, tB;
    AgMaps ag;
    if (mode == 1) {
        for (int d = 0; d < peers->world; ++d) {
            const uint64_t dims[4] = {uint64_t(p->c1), uint64_t(m_local), 1, 1};
            const uint64_t lda = p->a_ld > 0 ? uint64_t(p->a_ld) : uint64_t(p->c1);
            const uint64_t str[3] = {lda * 2, uint64_t(m_local) * lda * 2, uint64_t(m_local) * lda * 2};
            const uint32_t box[4] = {64, uint32_t(a.bw), uint32_t(a.bh), uint32_t(a.bi)};
            int rc = swh::encode_tmap_f16(&ag.m[d], peers->peer[d], 4, dims, str, box);
            if (rc) return rc;
        }
        for (int d = peers->world; d < 8; ++d) ag.m[d] = ag.m[0];
        tA1 = ag.m[0];
    } else {
        const uint64_t lda = p->a_ld > 0 ? uint64_t(p->a_ld) : uint64_t(p->c1);
        const uint64_t ah = p->a_h_total > 0 ? uint64_t(p->a_h_total) : uint64_t(in_h);
        const uint64_t dims[4] = {uint64_t(p->c1), uint64_t(in_w), ah, uint64_t(p->n)};
        const uint64_t str[3] = {lda * 2, uint64_t(in_w) * lda * 2, ah * in_w * lda * 2};
        // stride 2: the box spans 2x the pixels and the TMA unit keeps every other one
        const uint32_t box[4] = {64, uint32_t(a.bw * cstride), uint32_t(a.bh * cstride), uint32_t(a.bi)};
        const uint32_t es[4] = {1, uint32_t(cstride), uint32_t(cstride), 1};
        int rc = swh::encode_tmap_f16_strided(&tA1, p->a1, 4, dims, str, box, es);
        if (rc) return rc;
    }
    if (p->c2 > 0) {
        const uint64_t dims[4] = {uint64_t(p->c2), uint64_t(in_w), uint64_t(in_h), uint64_t(p->n)};
        const uint64_t str[3] = {uint64_t(p->c2) * 2, uint64_t(in_w) * p->c2 * 2,
                                 uint64_t(in_h) * in_w * p->c2 * 2};
        const uint32_t box[4] = {64, uint32_t(a.bw * cstride), uint32_t(a.bh * cstride), uint32_t(a.bi)};
        const uint32_t es[4] = {1, uint32_t(cstride), uint32_t(cstride), 1};
        int rc = swh::encode_tmap_f16_strided(&tA2, p->a2, 4, dims, str, box, es);
        if (rc) return rc;
    } else {
        tA2 = tA1;
    }
    a.w_kmajor = p->w_kmajor ? 1 : 0;
    if (a.w_kmajor) {
        if (p->w_ld > 0) return swh::fail(-1, "sw_igemm: k-block-major weights are dense (w_ld must be 0)");
        const uint64_t dims[3] = {64, uint64_t(p->cout), uint64_t(kblocks)};
        const uint64_t str[2] = {128, uint64_t(p->cout) * 128};
        const uint32_t box[3] = {64, uint32_t(pair ? bn / 2 : bn), 1};
        int rc = swh::encode_tmap_f16(&tB, p->wgt, 3, dims, str, box);
        if (rc) return rc;
    } else {
        const uint64_t dims[2] = {uint64_t(ktotal), uint64_t(p->cout) * (phases ? 4 : 1)};
        const uint64_t str[1] = {(p->w_ld > 0 ? uint64_t(p->w_ld) : uint64_t(ktotal)) * 2};
        const uint32_t box[2] = {64, uint32_t(pair ? bn / 2 : bn)};
        int rc = swh::encode_tmap_f16(&tB, p->wgt, 2, dims, str, box);
        if (rc) return rc;
    }
    // ---- epilogue through TMA stores (not for the reduce-scatter push or split-K partial tiles)
    CUtensorMap tOut, tRes, tOut32;
    // (GEGLU tiles produce a single 64-column slab: nothing to overlap the store's round trip with, the
    // direct register -> global path measured 30-50 % faster there)
    a.tma_epi = (mode != 2 && a.splits == 1 && !a.cluster_k && p->act != SW_ACT_GEGLU) ? 1 : 0;
    {
        static int env_epi = -1;  // SW_IGEMM_TMA_EPI=0 keeps the register -> global store epilogue (A/B runs)
        if (env_epi < 0) {
            const char* e = getenv("SW_IGEMM_TMA_EPI");
            env_epi = e ? atoi(e) : 1;
        }
        if (!env_epi && !ln_fused && !out_strided) a.tma_epi = 0;
    }
    if (out_strided && !a.tma_epi) return swh::fail(-1, "sw_igemm: a strided output needs the TMA-store epilogue");
    if (a.tma_epi) {
        const uint64_t nc = p->act == SW_ACT_GEGLU ? uint64_t(p->cout / 2) : uint64_t(p->cout);
        const uint64_t dims[4] = {nc, uint64_t(p->w), uint64_t(p->h), uint64_t(p->n)};
        // (strided output: pixel (y, x) lands in row y*osy, column x*osx of a [h*osy, w*osx] image)
        const uint64_t so[3] = {uint64_t(p->ldc) * 2 * osx, uint64_t(p->w) * osx * p->ldc * 2 * osy,
                                uint64_t(p->h) * osy * p->w * osx * p->ldc * 2};
        const uint32_t box[4] = {64, uint32_t(a.bw), uint32_t(a.bh), uint32_t(a.bi)};
        const uint32_t box32[4] = {32, uint32_t(a.bw), uint32_t(a.bh), uint32_t(a.bi)};
        int rc = swh::encode_tmap_f16(&tOut, p->out, 4, dims, so, box);
        if (rc) return rc;
        rc = swh::encode_tmap_f16_ex(&tOut32, p->out, 4, dims, so, box32, false);
        if (rc) return rc;
        tRes = tOut;
        if (p->residual) {
            const uint64_t sr[3] = {uint64_t(p->ldr) * 2, uint64_t(p->w) * p->ldr * 2,
                                    uint64_t(p->h) * p->w * p->ldr * 2};
            rc = swh::encode_tmap_f16(&tRes, p->residual, 4, dims, sr, box);
            if (rc) return rc;
        }
    } else {
        tOut = tRes = tOut32 = tB;
    }
    if (phases) {
        // the maps of phases 1..3 ride in the slots of tmRes, tmOut32 and tmA2 (unused by such a convolution)
        if (!a.tma_epi) return swh::fail(-1, "sw_igemm: out_phases needs the TMA-store epilogue");
        CUtensorMap* slot[4] = {&tOut, &tRes, &tOut32, &tA2};
        const uint64_t dims[4] = {uint64_t(p->cout), uint64_t(p->w), uint64_t(p->h), uint64_t(p->n)};
        const uint64_t so[3] = {uint64_t(p->ldc) * 4, uint64_t(p->w) * p->ldc * 8, uint64_t(p->h) * p->w * p->ldc * 8};
        const uint32_t box[4] = {64, uint32_t(a.bw), uint32_t(a.bh), uint32_t(a.bi)};
        for (int ph = 0; ph < 4; ++ph) {
            __half* base = static_cast<__half*>(p->out) + (size_t(ph >> 1) * 2 * p->w + (ph & 1)) * p->ldc;
            if (int rc = swh::encode_tmap_f16(slot[ph], base, 4, dims, so, box)) return rc;
        }
    }
    // ---- lean kernel (see igemm_lean_body): plain bias / row bias / residual epilogue, no split, no folded
    // LayerNorm -- most layers of a step
    {
        static int env_lean = -1;  // SW_IGEMM_LEAN=0 keeps the general kernel (A/B runs)
        if (env_lean < 0) {
            const char* e = getenv("SW_IGEMM_LEAN");
            env_lean = e ? atoi(e) : 1;
        }
        const bool lean_ok = mode == 0 && a.splits == 1 && !a.cluster_k && a.tma_epi && p->act == SW_ACT_NONE &&
                             a.out_scale == 1.0f && (!p->row_bias || a.bi == 1) && !ln_fused && p->persist != 2;
        if (env_lean && lean_ok) {
            const int boxes = p->residual ? 65536 : 32768;
            // a grid of at most one CTA per SM gains nothing from a second resident CTA: give the epilogue its own
            // staging boxes (residual prefetched under the main loop) and the ring every remaining byte
            a.epi_alias = ctas > long(sms) ? 1 : 0;
            int st = a.stages;
            if (!a.epi_alias) {
                st = (225 * 1024 - 1024 - 512 - boxes) / stage_bytes;
                if (st > 8) st = 8;
                if (p->stages > 0 && p->stages < st) st = p->stages;
                if (st > kblocks) st = kblocks;
                if (st < 1) st = 1;
            }
            a.stages = st;
            const int ring = a.epi_alias ? (st * stage_bytes < 65536 ? 65536 : st * stage_bytes) : st * stage_bytes + boxes;
            const int smem = ring + 1024 + 512;
            static bool configured[64] = {false};
            int dev = 0;
            SW_CUDA_OK(cudaGetDevice(&dev));
            if (dev < 64 && !configured[dev]) {
                SW_CUDA_OK(cudaFuncSetAttribute(k_igemm_lean, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
                SW_CUDA_OK(cudaFuncSetAttribute(k_igemm_lean_r, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
                SW_CUDA_OK(cudaFuncSetAttribute(k_igemm2_lean, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
                SW_CUDA_OK(cudaFuncSetAttribute(k_igemm2_lean_r, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
                configured[dev] = true;
            }
            const int ntile = (a.cout + a.bn - 1) / a.bn;
            cudaStream_t st_ = static_cast<cudaStream_t>(stream);
            if (!pair) {
                auto kern = p->residual ? k_igemm_lean_r : k_igemm_lean;
                SW_CUDA_OK(swh::launch_pdl(kern, dim3(grid_x, ntile, phases ? 4 : 1), dim3(kThreadsLean), size_t(smem), st_, 1,
                                           tA1, tA2, tB, tOut, tRes, tOut32, a));
            } else {
                auto kern = p->residual ? k_igemm2_lean_r : k_igemm2_lean;
                SW_CUDA_OK(swh::launch_pdl(kern, dim3((grid_x + 1) / 2 * 2, ntile, phases ? 4 : 1), dim3(kThreadsLean),
                                           size_t(smem), st_, 2, tA1, tA2, tB, tOut, tRes, tOut32, a));
            }
            return 0;
        }
    }
    if (phases) return swh::fail(-1, "sw_igemm: out_phases runs on the lean kernel only (SW_IGEMM_LEAN=0?)");
    // ---- persistent kernel (see k_igemm_p): worth it once a CTA would get more than one tile
    if (mode == 0 && !pair && !out_strided && a.splits == 1 && (!ln_fused || (p->act == SW_ACT_GEGLU && !p->row_stats_out))) {
        static int env_persist = -1;  // SW_IGEMM_PERSIST: 0 never, 1 auto (default), 2 whenever possible
        if (env_persist < 0) {
            const char* e = getenv("SW_IGEMM_PERSIST");
            env_persist = e ? atoi(e) : 1;
        }
        const long total = long(grid_x) * tiles_ncol;
        int want = p->persist ? p->persist : (env_persist == 0 ? 1 : env_persist == 2 ? 2 : 0);
        // measured: the GELU-gated epilogue is the one heavy enough to win from the overlap by default
        // (25-30 %); other shapes are within +-10 % either way and are left to the tuning table
        bool persist = want == 2 || (want == 0 && p->act == SW_ACT_GEGLU && total > 2L * sms);
        if (persist && total < (1L << 30)) {
            // tiles up to 128 columns wide: register -> global epilogue, which leaves room for two CTAs
            // (= eight epilogue warps, two MMA issuers) per SM; wider tiles stage through TMA stores
            if (bn <= 128) a.tma_epi = 0;
            const int stage_area = a.tma_epi ? 65536 : 0;
            // GEGLU (no staging boxes, epilogue-heavy): two persistent CTAs per SM = eight epilogue warps
            const int occ = (!a.tma_epi && 2 * a.tmem_cols * 2 <= 512 && 2 * (3 * stage_bytes + 4096) <= 227 * 1024) ? 2 : 1;
            int st = ((occ == 2 ? 110 : 222) * 1024 - 1280 - stage_area) / stage_bytes;
            if (st > 8) st = 8;
            if (st > kblocks * 2) st = kblocks * 2;
            if (st >= 2) {
                a.stages = st;
                a.ptiles = grid_x;
                a.total_tiles = int(total);
                const int smem = st * stage_bytes + stage_area + 1024 + 256;
                static bool configured[64] = {false};
                int dev = 0;
                SW_CUDA_OK(cudaGetDevice(&dev));
                if (dev < 64 && !configured[dev]) {
                    SW_CUDA_OK(cudaFuncSetAttribute(k_igemm_p, cudaFuncAttributeMaxDynamicSharedMemorySize, 222 * 1024));
                    configured[dev] = true;
                }
                const int grid = total < long(occ) * sms ? int(total) : occ * sms;
                SW_CUDA_OK(swh::launch_pdl(k_igemm_p, dim3(grid), dim3(kThreads), size_t(smem),
                                           static_cast<cudaStream_t>(stream), 1, tA1, tA2, tB, tOut, tRes, tOut32, a));
                return 0;
            }
        }
    }
    if (mode == 1) {
        const int smem = smem_bytes(a, pair);
        static bool configured[64] = {false};
        int dev = 0;
        SW_CUDA_OK(cudaGetDevice(&dev));
        if (dev < 64 && !configured[dev]) {
            SW_CUDA_OK(cudaFuncSetAttribute(k_igemm_ag, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
            SW_CUDA_OK(cudaFuncSetAttribute(k_igemm2_ag, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
            configured[dev] = true;
        }
        const int ntile = (a.cout + a.bn - 1) / a.bn;
        if (!pair) {
            k_igemm_ag<<<dim3(grid_x, ntile, 1), kThreads, smem, static_cast<cudaStream_t>(stream)>>>(ag, tB, tOut, tRes,
                                                                                                      tOut32, a);
            SW_CUDA_OK(cudaGetLastError());
            return 0;
        }
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3(grid_x, ntile, 1);  // even: m_local % 256 == 0
        cfg.blockDim = dim3(kThreads, 1, 1);
        cfg.dynamicSmemBytes = size_t(smem);
        cfg.stream = static_cast<cudaStream_t>(stream);
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        SW_CUDA_OK(cudaLaunchKernelEx(&cfg, k_igemm2_ag, ag, tB, tOut, tRes, tOut32, a));
        return 0;
    }
    return launch(a, tA1, tA2, tB, tOut, tRes, tOut32, grid_x, pair, static_cast<cudaStream_t>(stream));
}

// ---- fused collectives on the lean kernel (copy / reduce roles inside the GEMM grid, per-tile flags)
static int igemm_fused_lean(const sw_igemm_params* p, const sw_peers* peers, int mode, int m_local, void* final_out,
                            int ld_final, void* stream) {
    if (!p || !p->wgt || !peers || peers->world < 1 || peers->world > 8 || peers->rank < 0 || peers->rank >= peers->world)
        return swh::fail(-1, "sw_igemm fused: null argument or bad peer description");
    const int W = peers->world, R = peers->rank;
    if (p->kh != 1 || p->kw != 1 || p->n != 1 || p->h != 1 || p->c2 != 0 || p->act != SW_ACT_NONE || p->row_bias ||
        p->row_stats_in || p->row_stats_out || p->bf16 || p->w_kmajor)
        return swh::fail(-1, "sw_igemm fused: defined for plain fp16 GEMMs (no activation, no row bias)");
    if (m_local <= 0 || m_local % BM || p->w != m_local * W)
        return swh::fail(-1, "sw_igemm fused: m_local=%d must be a multiple of 128 and w == world*m_local", m_local);
    if (p->c1 <= 0 || p->c1 % 8 || p->cout <= 0 || p->cout % 8 || p->ldc % 8 || (p->a_ld && (p->a_ld % 8 || p->a_ld < p->c1)))
        return swh::fail(-1, "sw_igemm fused: channel counts / leading dimensions must be multiples of 8");
    if (mode == 1 && (!p->out || p->residual)) return swh::fail(-1, "sw_igemm_allgather: needs out, takes no residual");
    if (mode == 2 && (!p->a1 || !final_out || p->cout % 64 || ld_final % 8 || (p->residual && p->ldr % 8)))
        return swh::fail(-1, "sw_igemm_reducescatter: needs a1 and out, cout %% 64 == 0");
    for (int d = 0; d < W; ++d)
        if (!peers->peer[d] || !peers->flags[d]) return swh::fail(-1, "sw_igemm fused: peer %d has no buffer / flags", d);

    IGemmArgs a;
    memset(&a, 0, sizeof(a));
    a.c1 = p->c1;
    a.sx = 1;
    a.n = a.h = 1;
    a.w = p->w;
    a.kh = a.kw = 1;
    a.cout = p->cout;
    a.bw = BM;
    a.bh = a.bi = 1;
    a.tiles_x = p->w / BM;
    a.tiles_y = 1;
    a.bias = static_cast<const __half*>(p->bias);
    a.residual = static_cast<const __half*>(p->residual);
    a.ldr = p->ldr;
    a.out = static_cast<__half*>(p->out);
    a.ldc = p->ldc;
    a.out_scale = 1.0f;
    a.mode = mode;
    a.world = W;
    a.rank = R;
    a.epoch = peers->epoch;
    a.epoch_base = peers->epoch_base;
    a.epoch_mul = peers->epoch_mul;
    a.m_local = m_local;
    a.shard_tiles = m_local / BM;
    const int lda = p->a_ld > 0 ? p->a_ld : p->c1;
    a.a_ld = lda;
    for (int d = 0; d < W; ++d) {
        a.flags[d] = peers->flags[d];
        if (mode == 1) a.gathered[d] = static_cast<__half*>(peers->peer[d]);
    }
    if (mode == 2) {
        a.rs_stage = static_cast<const __half*>(peers->peer[R]);
        a.rs_final = static_cast<__half*>(final_out);
        a.ld_final = ld_final;
    }
    int bn = p->tile_n > 0 ? p->tile_n : 128;
    if (bn > p->cout) bn = (p->cout + 31) / 32 * 32;
    if (bn < 32 || bn > 256 || bn % 32 || (mode == 2 && bn % 64))
        return swh::fail(-1, "sw_igemm fused: tile_n=%d must be a multiple of 32 (64 for reduce-scatter) in [32,256]", bn);
    a.bn = bn;
    a.tmem_cols = bn <= 32 ? 32 : bn <= 64 ? 64 : bn <= 128 ? 128 : 256;
    const int ntile = (p->cout + bn - 1) / bn;
    const int kblocks = (p->c1 + 63) / 64;
    const int stage_bytes = A_STAGE_BYTES + bn * BK * 2;
    int stages = (3 * stage_bytes <= 110 * 1024) ? 3 : 4;
    if (stage_bytes <= 24 * 1024) stages = 4;
    if (p->stages > 0) stages = p->stages;
    while (stages > 2 && stages * stage_bytes > 200 * 1024) --stages;
    if (kblocks < stages) stages = kblocks;
    a.stages = stages;
    a.idesc = sw::umma_idesc(128, bn, 0, 0, 0);
    a.tma_epi = 1;
    a.epi_alias = 1;
    {
        const char* e = getenv("SW_FUSED_DBG");
        a.dbg = e ? atoi(e) : 0;
    }
    const int ring = stages * stage_bytes < 65536 ? 65536 : stages * stage_bytes;
    const int smem = ring + 1024 + 512;

    CUtensorMap tA, tB;
    {
        const void* abase = mode == 1 ? peers->peer[R] : p->a1;   // (all-gather: A is this rank's OWN gathered buffer)
        const uint64_t dims[4] = {uint64_t(p->c1), uint64_t(p->w), 1, 1};
        const uint64_t str[3] = {uint64_t(lda) * 2, uint64_t(p->w) * lda * 2, uint64_t(p->w) * lda * 2};
        const uint32_t box[4] = {64, BM, 1, 1};
        if (int rc = swh::encode_tmap_f16(&tA, abase, 4, dims, str, box)) return rc;
        const uint64_t bd[2] = {uint64_t(p->c1), uint64_t(p->cout)};
        const uint64_t bs[1] = {(p->w_ld > 0 ? uint64_t(p->w_ld) : uint64_t(p->c1)) * 2};
        const uint32_t bb[2] = {64, uint32_t(bn)};
        if (int rc = swh::encode_tmap_f16(&tB, p->wgt, 2, bd, bs, bb)) return rc;
    }
    static bool configured[64] = {false};
    int dev = 0;
    SW_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 64 && !configured[dev]) {
        SW_CUDA_OK(cudaFuncSetAttribute(k_igemm_lean_ag, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        SW_CUDA_OK(cudaFuncSetAttribute(k_igemm_lean_rs, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        configured[dev] = true;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // 1-D grid: [copy blocks |] compute blocks (row tile major) [| one reduce block per tile of this rank's rows]
    const dim3 grid(a.tiles_x * ntile + a.shard_tiles * (mode == 1 ? 1 : ntile), 1, 1);
    if (mode == 1) {
        CUtensorMap tOut, tOut32;
        const uint64_t dims[4] = {uint64_t(p->cout), uint64_t(p->w), 1, 1};
        const uint64_t so[3] = {uint64_t(p->ldc) * 2, uint64_t(p->w) * p->ldc * 2, uint64_t(p->w) * p->ldc * 2};
        const uint32_t box[4] = {64, BM, 1, 1}, box32[4] = {32, BM, 1, 1};
        if (int rc = swh::encode_tmap_f16(&tOut, p->out, 4, dims, so, box)) return rc;
        if (int rc = swh::encode_tmap_f16_ex(&tOut32, p->out, 4, dims, so, box32, false)) return rc;
        SW_CUDA_OK(swh::launch_pdl(k_igemm_lean_ag, grid, dim3(kThreadsLean), size_t(smem), st, 1, tA, tB, tOut, tOut32, a));
    } else {
        PeerMaps pm;
        for (int d = 0; d < W; ++d) {  // rank d's staging slot [R]: rows of d's shard, as this rank's partial
            __half* base = static_cast<__half*>(peers->peer[d]) + size_t(R) * m_local * p->ldc;
            const uint64_t dims[4] = {uint64_t(p->cout), uint64_t(m_local), 1, 1};
            const uint64_t so[3] = {uint64_t(p->ldc) * 2, uint64_t(m_local) * p->ldc * 2, uint64_t(m_local) * p->ldc * 2};
            const uint32_t box[4] = {64, BM, 1, 1};
            if (int rc = swh::encode_tmap_f16(&pm.m[d], base, 4, dims, so, box)) return rc;
        }
        for (int d = W; d < 8; ++d) pm.m[d] = pm.m[0];
        SW_CUDA_OK(swh::launch_pdl(k_igemm_lean_rs, grid, dim3(kThreadsLean), size_t(smem), st, 1, tA, tB, pm, a));
    }
    SW_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int sw_igemm_allgather(const sw_igemm_params* p, const sw_peers* peers, int m_local, void* stream) {
    return igemm_fused_lean(p, peers, 1, m_local, nullptr, 0, stream);
}

extern "C" int sw_igemm_reducescatter(const sw_igemm_params* p, const sw_peers* peers, int m_local, void* out,
                                      int ldo, void* stream) {
    return igemm_fused_lean(p, peers, 2, m_local, out, ldo, stream);
}

extern "C" int sw_igemm(const sw_igemm_params* p, void* stream) {
    return igemm_impl(p, nullptr, 0, 0, nullptr, stream);
}

extern "C" int sw_igemm_ag(const sw_igemm_params* p, const sw_peers* peers, int m_local, void* stream) {
    return igemm_impl(p, peers, 1, m_local, nullptr, stream);
}

extern "C" int sw_igemm_rs(const sw_igemm_params* p, const sw_peers* peers, int m_local,
                           unsigned int* ticket, void* stream) {
    return igemm_impl(p, peers, 2, m_local, ticket, stream);
}

// Probe builds (-DSW_TRACE) only: copies the per-CTA phase stamps of the last launches to the host.
SW_TRACE_DUMP_FN(sw_trace_dump)
