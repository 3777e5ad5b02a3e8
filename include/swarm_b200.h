/* swarm_b200.h -- C ABI of libswarm_sm100.so (hand-written sm_100a kernels for the
 * chiaswarm_b200 diffusion worker).
 *
 * Conventions
 *  - every pointer is a device pointer unless its name ends in _host;
 *  - activations are fp16, channels-last ("NHWC" == tokens [N*H*W, C]); accumulation is fp32;
 *  - `stream` is a cudaStream_t passed as void*; all entry points are asynchronous on it;
 *  - return value 0 == success; otherwise sw_last_error() describes the failure.
 *
 * The reference (Jsewill/chiaSWARM) has no native code at all: each entry point below names the
 * library kernel family it replaces and the reference call
 * site that triggers that family.
 */
#ifndef SWARM_B200_H_
#define SWARM_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* sw_last_error(void);

/* Peer description for the fused compute+collective entry points (one process per GPU, buffers
 * in symmetric memory mapped on every rank; see chiaswarm_b200/parallel/runtime.py).
 * peer[d]  : rank d's copy of the buffer the operation communicates through
 * flags[d] : rank d's int32[world] epoch flags; slot [src] is written by rank src
 * epoch    : strictly increasing per use of a flag set. */
typedef struct {
    int32_t world, rank, epoch;
    void* peer[8];
    int32_t* flags[8];
    /* Optional device-resident epoch offset: the effective epoch is epoch + *epoch_base * epoch_mul.
     * A captured CUDA graph bakes `epoch` in; the counter (advanced once per replay) keeps the
     * flags monotonic across replays.  NULL = use `epoch` as is. */
    const int32_t* epoch_base;
    int32_t epoch_mul;
} sw_peers;

/* Epilogue activations for sw_igemm. */
enum { SW_ACT_NONE = 0, SW_ACT_SILU = 1, SW_ACT_GELU = 2, SW_ACT_GEGLU = 3 };

/* Implicit-GEMM convolution / linear layer on tcgen05 tensor cores (TMA-fed, TMEM accumulator).
 * Replaces cuDNN conv2d (K4) and cuBLAS hgemm (K3) reached from the reference's
 * `pipeline(**kwargs)` call (swarm/diffusion/diffusion_func.py:96).
 *
 *   out[n,y,x,:] = act( sum_{ky,kx,c} in[n, y+ky-pad, x+kx-pad, c] * wgt[:, (ky*kw+kx)*C + c]
 *                       + bias + row_bias[n,:] ) * out_scale + residual[n,y,x,:]
 *
 * `in` is the channel-concatenation of a1 (c1 channels) and a2 (c2 channels, may be NULL): the
 * UNet skip-concat is consumed as two K-slabs, never materialised.  A plain linear layer is the
 * case n=h=1, w=M, kh=kw=1, pad=0.  Stride 1 unless `stride` says 2 (see the end of the struct).
 * With act == SW_ACT_GEGLU the weight rows must be pre-interleaved in blocks of 64 value rows
 * followed by their 64 gate rows; `cout` counts weight rows and the output has cout/2 columns.
 */
typedef struct {
    const void* a1;      /* fp16 [n,h,w,c1] */
    const void* a2;      /* fp16 [n,h,w,c2] or NULL */
    int32_t c1, c2;      /* channel counts, multiples of 8 */
    int32_t n, h, w;
    int32_t kh, kw, pad;
    const void* wgt;     /* fp16 [cout, kh*kw*(c1+c2)], K contiguous */
    int32_t cout;        /* multiple of 8 */
    const void* bias;    /* fp16 [cout] or NULL */
    const void* row_bias;/* fp16 [n, row_bias_ld] or NULL (per-image bias, e.g. time embedding) */
    int32_t row_bias_ld;
    const void* residual;/* fp16 [n*h*w, ldr] or NULL */
    int32_t ldr;
    void* out;           /* fp16 [n*h*w, ldc] */
    int32_t ldc;
    int32_t act;
    float out_scale;     /* applied after the activation, before the residual add */
    int32_t bf16;        /* 0: fp16 operands, 1: bf16 operands */
    int32_t tile_n;      /* 0 = auto, else a multiple of 32 in [32, 256] */
    float* gn_stats;     /* optional fp32 [n, gn_groups, 2]: += (sum, sumsq) of the stored output */
    int32_t gn_groups;
    int32_t a_ld;        /* elements between consecutive pixels of a1 (0 = c1): strided-view operand */
    int32_t w_ld;        /* elements between consecutive weight rows (0 = kh*kw*(c1+c2)) */
    void* splitk_scratch;/* optional: enables split-K for layers with few output tiles.  Caller-owned,
                            zero-filled once before first use (the kernel leaves it reusable);
                            calls sharing one scratch must be stream-ordered. */
    int64_t splitk_scratch_bytes;
    int32_t splitk_min_kb; /* 0 = default (16): fewest 64-wide K blocks a split may get */
    int32_t pad_w1;      /* 0 = horizontal padding equals `pad`; else horizontal padding + 1 (temporal (3,1,1) convs) */
    int32_t stages;      /* 0 = auto, else operand ring depth (clamped to what fits in shared memory) */
    int32_t cta_pair;    /* 0 = auto, 1 = single-CTA tiles (128 x N), 2 = CTA-pair tiles (256 x N, cta_group::2) */
    /* a1 taller than the output: a1 is [n, a_h_total, w, c1] and output row y reads input rows
     * y + a_y_off + ky - pad (rows outside [0, a_h_total) are zero).  Patch-parallel convs read a
     * buffer with one halo row above and below (a_h_total = h + 2, a_y_off = 1) or a window of the
     * full image.  0 = a1 has exactly h rows. */
    int32_t a_h_total;
    int32_t a_y_off;
    /* LayerNorm folded into the GEMM that consumes it.  The consumer runs on the RAW activation with
     * weights pre-multiplied by gamma (bias' = W.beta + b, col_sums[c] = sum_k W'[c,k], host side) and
     * its epilogue computes rstd_r * (acc - mean_r * col_sums[c]) + bias'[c]; (mean, rstd) of row r come
     * from `stats_tiles` partial (sum, sum of squares) pairs that the GEMM PRODUCING the activation wrote
     * to row_stats_out = float [its column tiles][rows][2] in its own epilogue (pass tile_n so the tile
     * count is known).  Replaces the LayerNorm launch and the normalised copy of the activation. */
    float* row_stats_out;
    const float* row_stats_in;
    const float* col_sums;
    int32_t stats_tiles;
    int32_t ln_width;
    float ln_eps;
    int32_t cluster_k;   /* 0 / 1 = off, 2 = split K over a cluster of two CTAs (partials meet in distributed shared memory) when eligible */
    int32_t persist;     /* 0 = auto, 1 = one tile per CTA, 2 = persistent CTAs with double-buffered accumulators when possible */
    int32_t w_kmajor;    /* 0: wgt is [cout, kh*kw*(c1+c2)] row-major.  1: wgt is packed k-block-major,
                            [k-blocks][cout][64] fp16, k-block = (filter tap, 64-channel slab of a1, then of a2), slabs
                            zero-padded to 64 channels: every weight tile the kernel fetches is then one contiguous
                            run of memory (sequential DRAM bursts instead of one 128-byte segment per row), which is
                            what weight-streaming layers (few pixels, 1280 channels) are bound by */
    /* Strided convolution as implicit GEMM (no im2col buffer): output pixel (y, x) reads input pixels
     * (y*stride + ky - pad, x*stride + kx - pad) of an input of in_h x in_w pixels (h, w stay the OUTPUT size); the
     * TMA box simply skips every other pixel.  0 / 1 = stride 1 with in_h = h, in_w = w. */
    int32_t stride;
    int32_t in_h, in_w;
    /* Output pixel strides (0 = 1): pixel (n, y, x) is stored at out + ((n*h*out_sy + y*out_sy) * w*out_sx +
     * x*out_sx) * ldc, i.e. into every out_sy-th row / out_sx-th column of a [n, h*out_sy, w*out_sx, ldc] tensor whose
     * first written pixel `out` points at.  A nearest-2x upsample followed by a 3x3 conv is four such 2x2 convs on the
     * small map, one per output phase, with the filter taps pre-summed (ops/packing.py::fold_upsample2x).  Needs the
     * TMA-store epilogue: no residual, no split-K, no GEGLU. */
    int32_t out_sx, out_sy;
    /* 1: all four phases of such a folded upsample in ONE launch (grid z = phase a*2+b).  kh = kw = 2, pad = 1 (the
     * kernel takes 1-a / 1-b per phase), out_sx = out_sy = 2, `out` = the first pixel of the [n, 2h, 2w, ldc] tensor,
     * wgt = the four phase matrices stacked, [4*cout, 4*c1]; bias only (no residual / activation / second input),
     * cout % 64 == 0. */
    int32_t out_phases;
} sw_igemm_params;

int sw_igemm(const sw_igemm_params* p, void* stream);

/* All-gather fused into the GEMM (collective call site 2): the activation is
 * row-sharded, rank d holds X_d [m_local, c1] at peers->peer[d]; the kernel computes
 * out[world*m_local, cout] = concat_d(X_d) . wgt^T by TMA-loading every A tile straight from the
 * rank that owns it (NVLink peer mapping) -- no gathered copy exists.  The loader of a remote
 * tile first waits for that rank's epoch flag (raised by sw_peer_signal or a producer kernel).
 * p->a1 is ignored; p->w must equal world*m_local; m_local % 128 == 0.
 * Reuse rule for all peer-visible operands of the fused kernels (activation shards, K / V, staging, gathered
 * buffers): there is no consumer-done handshake, so a buffer a peer may still be reading must not be rewritten until
 * a sw_peer_exchange barrier or a flag round trip in the opposite direction has passed -- or alternate two buffers by
 * epoch parity (what parallel/collectives.py::FusedOps and the GroupNorm slots do). */
int sw_igemm_ag(const sw_igemm_params* p, const sw_peers* peers, int m_local, void* stream);

/* GEMM fused with reduce-scatter (collective call site 3): every rank multiplies
 * its K-shard, partial = A_r . W_r^T for all world*m_local rows, and the epilogue pushes each
 * partial row over NVLink into the staging slot [rank] of the rank that owns that row
 * (peers->peer[d]: fp16 [world][m_local][ldc] on rank d).  The CTA that finishes last raises this
 * rank's epoch flag on every peer.  sw_rs_reduce (owner side) waits for all flags and adds the
 * world slots in rank order (+ bias, + residual).  `ticket` is a zeroed uint32 in local memory. */
int sw_igemm_rs(const sw_igemm_params* p, const sw_peers* peers, int m_local, uint32_t* ticket, void* stream);
int sw_rs_reduce(const void* staging_local, void* out, const void* bias, const void* residual, int m_local,
                 int ncols, int ld_stage, int ldo, int ldr, const sw_peers* peers, void* stream);

/* The two halves of a GroupNorm as separate kernels (per-(image, group) sum / sum of squares into fp32
 * sums[n][groups][2]; apply given such sums over `count` elements per group): the like-for-like baseline of
 * sw_groupnorm_nhwc_dist with a library all-reduce in between.  Not used by any product path. */
int sw_gn_partial_sums(const void* x, float* sums, int n, int hw, int c, int groups, void* stream);
int sw_gn_apply_sums(const void* x, const float* sums, const void* gamma, const void* beta, void* y, int n, int hw,
                     int c, int groups, long long count, float eps, int silu, void* stream);

/* Second generation of the two fused GEMM collectives (what parallel/collectives.py and the tensor-parallel block use;
 * the first generation above is kept for A/B runs).  Both add a second role to the GEMM's own grid, tied to the
 * compute blocks by PER-TILE epoch flags, so data crosses NVLink once and reduction starts tile by tile:
 *
 * sw_igemm_allgather: peers->peer[d] = rank d's GATHERED activation [world*m_local, a_ld] (symmetric memory); this
 *   rank's shard already sits in rows [rank*m_local, (rank+1)*m_local) of its own buffer (the producing kernel wrote
 *   it there).  Copy blocks (dispatched first) push the shard tile by tile into the same rows of every peer's
 *   buffer and raise flag [rank*T + tile] there (T = m_local/128); a compute block whose rows came from rank s waits
 *   for flag [s*T + tile] and then TMA-loads LOCAL memory only.  p->a1 is ignored, p->a_ld is the pitch of the
 *   gathered buffer, p->w == world*m_local.  peers->flags[d]: int32 [world*T] on rank d.
 *   Reuse rule: a peer may still read its gathered buffer while this rank starts the next call -- alternate two
 *   gathered buffers (epoch parity) or put a collective of the opposite direction in between.
 *
 * sw_igemm_reducescatter: partial = a1 . wgt^T over this rank's K shard for all world*m_local rows; a compute block
 *   TMA-stores its fp16 tile into the staging slot [rank] of the rows' owner (peers->peer[d]: fp16
 *   [world][m_local][ldc] on rank d) and raises flag [(rank*T + tile_m)*NT + tile_n] there once the stores have
 *   completed (NT = column tiles); reduce blocks (dispatched last) wait for the `world` flags of a tile of this
 *   rank's rows and write out[m_local, ldo] = sum over ranks in rank order, fp32 (+ p->bias, + p->residual).
 *   p->out is ignored; p->cout % 64 == 0.  peers->flags[d]: int32 [world*T*NT].  Same reuse rule for `staging`.
 * Plain fp16 GEMMs only (no activation); m_local % 128 == 0; p->tile_n (default 128) names the column tile. */
int sw_igemm_allgather(const sw_igemm_params* p, const sw_peers* peers, int m_local, void* stream);
int sw_igemm_reducescatter(const sw_igemm_params* p, const sw_peers* peers, int m_local, void* out, int ldo,
                           void* stream);

/* Epoch-flag signal / wait between the ranks of a gang (st.release.sys / ld.acquire.sys on peer
 * mappings; bounded spin).  Used to publish a tensor to the peers and as the NCCL-free barrier. */
int sw_peer_signal(const sw_peers* peers, void* stream);
int sw_peer_wait(const sw_peers* peers, void* stream);

/* Pairwise push exchange over NVLink peer mappings: for each transfer, `nbytes` (multiple of 16)
 * go from local `src` to `dst_ptr` (an address inside rank `dst`'s mapping); the destination's
 * flag slot [rank] is then raised to the epoch and the kernel waits for the flag rank `dst` raises
 * after its own transfer to this rank.  Every rank of the exchange must issue the mirror call.
 * Used for conv halo rows between neighbouring patches and for the per-step
 * epsilon all-gather of the patch-parallel denoiser; the reference has no counterpart (single GPU
 * per job, swarm/worker.py:41-43).  At most 8 transfers, one per destination. */
typedef struct {
    int32_t dst;        /* index into peers->flags */
    const void* src;    /* local */
    void* dst_ptr;      /* peer-mapped */
    int64_t nbytes;
} sw_xfer;
int sw_peer_exchange(const sw_peers* peers, const sw_xfer* xfers, int count, void* stream);

/* Flash attention (no mask, or causal for short sequences) on tcgen05: S=QK^T and O+=PV accumulate in TMEM, online
 * softmax in registers.  Replaces SDPA / xformers memory_efficient_attention (K1, K2), reference
 * call sites swarm/diffusion/diffusion_func.py:86-87,96.
 * q/k/v are fp16 with layout [batch, seq, heads, head_dim] expressed through element strides
 * (head_dim contiguous); out is fp16 [batch, seq_q, heads*head_dim] with row stride ldo.
 * head_dim must be a multiple of 8, at most 160.
 */
typedef struct {
    const void* q; const void* k; const void* v;
    int64_t q_stride_b, q_stride_s, q_stride_h; /* elements */
    int64_t k_stride_b, k_stride_s, k_stride_h;
    int64_t v_stride_b, v_stride_s, v_stride_h;
    void* out; int64_t o_stride_b, o_stride_s;  /* elements; head h starts at column h*head_dim */
    int32_t batch, heads, seq_q, seq_kv, head_dim;
    float scale;                                 /* softmax scale, usually head_dim^-0.5 */
    int32_t causal;                              /* 1: query i attends to keys 0..i only (CLIP text encoder); needs
                                                    seq_q <= 128 (one query tile), sw_flash_attn only */
} sw_attn_params;

int sw_flash_attn(const sw_attn_params* p, void* stream);

/* Self-attention with the cross-patch K/V all-gather fused in (collective call
 * site 1): queries are this rank's rows, K/V are row-sharded over `world` ranks (s_local rows each,
 * same strides everywhere).  The KV loop starts on the local shard and then TMA-loads the peers'
 * tiles through their NVLink mappings -- K/V are never gathered into one buffer.
 * kpeers->peer[d] = rank d's K base, v_ptrs[d] = rank d's V base; p->k / p->v are ignored;
 * p->seq_kv == world * s_local, s_local % 128 == 0, head_dim <= 128. */
int sw_flash_attn_gather(const sw_attn_params* p, const sw_peers* kpeers, const void* const* v_ptrs,
                         int s_local, void* stream);

/* GroupNorm (+ optional SiLU) over channels-last fp16.  Replaces ATen GroupNorm + SiLU (K6, K8;
 * reference call site swarm/diffusion/diffusion_func.py:96).
 * The input is the channel-concatenation of x [n,hw,c1] and x2 [n,hw,c2] (x2 may be NULL with
 * c2 = 0): the UNet skip-concat is normalised without being materialised.  y: [n,hw,c1+c2];
 * gamma/beta fp16 [c1+c2].  Two paths, chosen per shape (sw_groupnorm_kernel_count tells which):
 *  1 launch : an image (or a slice of whole groups of it) fits the shared memory of a cluster of
 *             <= 16 CTAs: rows are parked in shared memory, partial sums meet through distributed
 *             shared memory behind one cluster barrier, the apply pass reads shared memory.  This
 *             path touches no scratch at all.
 *  2 launches: a statistics kernel writes per-CTA partial sums to `stats_scratch` and the last CTA
 *             of each image (ticket counter in the scratch, cleared by a memset node on `stream`
 *             before the launch) folds them in a fixed order; an apply kernel normalises.
 * Both are bit-reproducible (no floating-point atomics).
 * `stats_scratch` is caller-owned fp32 memory of sw_groupnorm_scratch_floats(n,hw,c,groups)
 * floats; its contents need no initialisation. */
int64_t sw_groupnorm_scratch_floats(int n, int hw, int c, int groups);
/* Kernels sw_groupnorm_nhwc launches for this shape: 1 (single-launch path) or 2 (statistics + apply). */
int sw_groupnorm_kernel_count(int n, int hw, int c, int groups);
int sw_groupnorm_nhwc(const void* x, const void* x2, void* y, const void* gamma, const void* beta,
                      int n, int hw, int c1, int c2, int groups, float eps, int silu,
                      float* stats_scratch, void* stream);
/* Alias of sw_groupnorm_nhwc kept for callers of the first release (its scratch-pre-zeroed fast
 * path was replaced by the scratch-free cluster kernel described above). */
int sw_groupnorm_nhwc_zeroed(const void* x, const void* x2, void* y, const void* gamma, const void* beta,
                             int n, int hw, int c1, int c2, int groups, float eps, int silu,
                             float* stats_scratch, void* stream);

/* GroupNorm over a feature map whose rows are sharded across `peers->world` GPUs (patch /
 * sequence parallel): the statistics all-reduce (SURVEY.md collective call site 4) is done by the
 * GroupNorm kernels themselves with peer stores + epoch flags over NVLink, no NCCL call.
 * hw is the LOCAL number of pixels; every rank must hold the same hw.
 * peers->peer[d]: fp32 [2][world][n][groups][2] on rank d. */
int sw_groupnorm_nhwc_dist(const void* x, const void* x2, void* y, const void* gamma, const void* beta,
                           int n, int hw, int c1, int c2, int groups, float eps, int silu,
                           float* stats_scratch, const sw_peers* peers, void* stream);

/* Same operation as sw_groupnorm_nhwc for callers that cannot allocate device memory (this entry
 * point has no counterpart in the reference, which has no native code).  Who needs it: a host
 * process that committed all of its device memory to its own arena at start-up and has nothing
 * left to carve a scratch buffer from, or a caller on a latency-critical path that may not call
 * cudaMalloc/cudaFree because they can stall the whole device.  The statistics scratch is a
 * fixed array of SW_GN_STATIC_MAX_FLOATS floats (1 MiB) compiled into the library; every GPU
 * that loads the library keeps its copy until the process exits.  The call itself performs no
 * allocation of any kind.
 *  - Limit: sw_groupnorm_scratch_floats(n, hw, c1+c2, groups) must not exceed
 *    SW_GN_STATIC_MAX_FLOATS, otherwise the call fails with a message saying so (for 32 groups
 *    that is up to 15 images per call at any resolution).
 *  - The first call on each device looks up the array's device address (a one-time, per-device
 *    set-up done with that device current); make that first call before starting any stream
 *    capture on the thread.
 *  - All calls on a device share the one array, which still holds the previous call's values;
 *    the call clears the few counters that must start at zero itself.  Two calls whose kernels
 *    can overlap on the GPU (different streams without an ordering between them) race on it:
 *    keep them on one stream or order them with events. */
#define SW_GN_STATIC_MAX_FLOATS (256 * 1024)
int sw_groupnorm_nhwc_static(const void* x, const void* x2, void* y, const void* gamma, const void* beta,
                             int n, int hw, int c1, int c2, int groups, float eps, int silu,
                             void* stream);

/* LayerNorm over the last dim of fp16 [rows, c] (K7). */
int sw_layernorm(const void* x, void* y, const void* gamma, const void* beta, int rows, int c,
                 float eps, void* stream);

/* Row softmax of fp16 [rows, cols]: y = softmax(x * scale) (K10: the VAE mid-block attention has
 * one 512-wide head, so its scores are materialised by sw_igemm and normalised here). */
int sw_softmax_rows(const void* x, void* y, int rows, int cols, float scale, void* stream);

/* Nearest-neighbour 2x upsample, channels-last fp16 (K8). */
int sw_upsample2x_nhwc(const void* x, void* y, int n, int h, int w, int c, void* stream);

/* im2col for strided / asymmetric-pad convolutions: out[n,oy,ox,(ky*kw+kx)*c + ch] (K4 helper). */
int sw_im2col_nhwc(const void* x, void* out, int n, int h, int w, int c, int kh, int kw, int stride,
                   int pad_top, int pad_left, int oh, int ow, void* stream);

/* Layout converters between the public NCHW tensors and the internal channels-last fp16 ones.
 * c_pad >= c: extra channels are written as zero (nchw->nhwc) / ignored (nhwc->nchw). */
int sw_nchw_to_nhwc(const void* x, void* y, int n, int c, int h, int w, int c_pad, int src_fp32,
                    void* stream);
int sw_nhwc_to_nchw(const void* x, void* y, int n, int c, int h, int w, int c_pad, int dst_fp32,
                    void* stream);

/* Sinusoidal timestep embedding (flip_sin_to_cos=True, freq_shift=0): out fp16 [n, dim] (K8). */
int sw_timestep_embedding(const float* t_table, const int32_t* step_idx, void* out, int n, int dim,
                          float max_period, int flip_sin_to_cos, float freq_shift, void* stream);

/* ++*counter on the device: lets one captured CUDA graph walk per-step tables on replay. */
int sw_advance_counter(int32_t* counter, void* stream);

/* y = a + b (fp16, n elements, multiple of 8) and friends (K8). */
int sw_add(const void* a, const void* b, void* y, int64_t n, void* stream);
int sw_silu(const void* x, void* y, int64_t n, void* stream);

/* Fused classifier-free guidance + scheduler update (K9; replaces the diffusers scheduler.step
 * python/elementwise chain reached from swarm/diffusion/diffusion_func.py:71-74,96).
 * Every supported scheduler (DPM-Solver++ 2M, Euler, Euler-ancestral, DDIM) is linear in
 * (x, x0, previous x0, noise), so the host precomputes per-step coefficients in fp64
 * (ops/schedulers.py) and the kernel evaluates, per latent element,
 *     eps    = CFG(eps_0 .. eps_{n_cond-1})
 *     x0     = e0*x + e1*eps
 *     x_new  = c0*x + c1*x0 + c2*x0_prev + c3*noise
 *     x0_prev = x0;  x = x_new;  model_in[k] = fp16(x_new * s_in) for k < n_cond
 * coef row = {e0, e1, c0, c1, c2, c3, s_in, 0}; row `*step_idx` of the table is used.
 * n_cond == 2: eps = eps_0 + g_text*(eps_1 - eps_0)                 (order [uncond, cond])
 * n_cond == 3: eps = eps_2 + g_text*(eps_0 - eps_1) + g_image*(eps_1 - eps_2)  (pix2pix order)
 */
typedef struct {
    const void* eps;        /* fp16 [n_cond, rows, eps_ld] model output, channels-last */
    int32_t n_cond;         /* 1, 2 or 3 */
    int32_t rows;           /* batch*h*w latent pixels */
    int32_t c;              /* latent channels (4) */
    int32_t eps_ld;         /* padded channel stride of eps */
    float g_text, g_image;
    float* x;               /* fp32 [rows, c] latent, updated in place */
    float* x0_prev;         /* fp32 [rows, c] multistep state, in/out */
    const float* noise;     /* fp32 [steps, rows, c] ancestral noise table or NULL */
    void* x_model_in;       /* fp16 [n_cond, rows, xin_ld]: columns [0,c) are rewritten */
    int32_t xin_ld;
    const float* coef;      /* fp32 [steps, 8] */
    const int32_t* step_idx;/* device step counter (NULL = row 0) */
    const float* guidance;  /* optional device fp32 [2] = {g_text, g_image}: overrides the scalars above, so one
                               captured CUDA graph serves every guidance value (NULL = use the scalars) */
} sw_sched_params;

int sw_sched_step(const sw_sched_params* p, void* stream);

/* VAE output head: clamp(x/2+0.5,0,1)*255 -> uint8 HWC, from channels-last fp16 (K10). */
int sw_image_to_u8(const void* x_nhwc, uint8_t* out, int64_t pixels, int c_pad, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SWARM_B200_H_ */
