"""Plan recorder and packed layers for the sm_100a execution engine.

The engine does not run nn.Modules.  It walks the module tree once (the modules are the
checkpoint-compatible *specification*, models/*.py), re-packs every weight into the layout the
kernels want, allocates every activation buffer up front and records a flat list of kernel
launches -- a static plan.  Executing the plan issues only our own kernels (ops/native.py), so
the whole thing can be captured into one CUDA graph and replayed per denoising step.

Activations are channels-last fp16 everywhere: a feature map [N,H,W,C] *is* the token matrix
[N*H*W, C], so convs, 1x1 projections and transformer blocks chain without a single permute.
Replaces the eager diffusers module graph reached at swarm/diffusion/diffusion_func.py:96.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import os

import torch

from ..ops import native as nat
from ..ops import tuning
from ..ops.packing import fold_layernorm, fold_upsample2x, interleave_geglu, pack_conv_weight, pad_vector


def _r8(c: int) -> int:
    return (c + 7) // 8 * 8


@dataclass
class FM:
    """Channels-last feature map view of a flat fp16 buffer."""

    t: torch.Tensor  # [n, h, w, c]
    n: int
    h: int
    w: int
    c: int
    # patch-parallel plans (engine/patch_engine.py): `sym` is the peer-visible [h+2, w, c] buffer whose
    # interior `t` views (rows 0 and h+1 are the halo rows the neighbours fill); `window` = (rows of the
    # full image `t` actually holds, first row of this rank's strip) for inputs every rank keeps whole
    sym: object = None
    window: Optional[tuple] = None

    @property
    def rows(self):
        return self.n * self.h * self.w

    def tokens(self):
        return self.t.view(self.rows, self.c)


class Recorder:
    """Records kernel launches with concrete buffers; `run()` replays them in order."""

    patch = None  # PatchContext when the plan runs patch-parallel (engine/patch_engine.py)

    _scratch = {}  # device -> (GroupNorm scratch, split-K scratch): ONE pair per device, shared by every plan

    def __init__(self, device):
        self.device = torch.device(device)
        self.ops = []
        self._free = {}
        self.bytes_allocated = 0
        # GroupNorm scratch and split-K scratch (tickets + fp32 partial tiles, zeroed once, left reusable by the
        # kernels).  Plans of a device run one after the other on one stream, so they share a single pair instead of
        # each cached graph (and each ControlNet job's embedder plan) allocating and clearing 68 MiB of its own.
        key = (self.device.type, self.device.index if self.device.index is not None else torch.cuda.current_device())
        pair = Recorder._scratch.get(key)
        if pair is None:
            pair = Recorder._scratch[key] = (
                torch.zeros(1 << 20, dtype=torch.float32, device=self.device),
                torch.zeros((64 << 20) + 16384, dtype=torch.uint8, device=self.device))
        self.gn_stats, self.splitk = pair

    # ---- buffers (explicit liveness: release() makes a buffer reusable by later ops)
    def buf(self, *shape, dtype=torch.float16):
        numel = 1
        for s in shape:
            numel *= s
        key = (numel, dtype)
        lst = self._free.get(key)
        if lst:
            return lst.pop().view(*shape)
        self.bytes_allocated += numel * torch.empty((), dtype=dtype).element_size()
        return torch.empty(*shape, dtype=dtype, device=self.device)

    def fm(self, n, h, w, c) -> FM:
        return FM(self.buf(n, h, w, c), n, h, w, c)

    def release(self, t):
        if isinstance(t, FM):
            t = t.t
        self._free.setdefault((t.numel(), t.dtype), []).append(t.view(-1))

    # ---- recording
    def add(self, fn, *args, **kw):
        self.ops.append((fn, args, kw))

    def run(self):
        for fn, args, kw in self.ops:
            fn(*args, **kw)

    def __len__(self):
        return len(self.ops)

    def profile(self, iters: int = 5):
        """Per-op device time (CUDA events, op by op, warm): list of (label, microseconds)."""
        out = []
        for fn, args, kw in self.ops:
            shapes = [tuple(a.shape) for a in args if isinstance(a, torch.Tensor)][:3]
            extra = {k: kw[k] for k in ("n", "h", "w", "kh", "act", "heads", "seq_q", "seq_kv", "head_dim", "hw", "groups")
                     if k in kw}
            label = f"{getattr(fn, '__name__', 'op')} {shapes} {extra}"
            fn(*args, **kw)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                fn(*args, **kw)
            e1.record()
            torch.cuda.synchronize()
            out.append((label, e0.elapsed_time(e1) / iters * 1e3))
        return out


# ------------------------------------------------------------------------------------------ layers
class Conv:
    """3x3 / 1x1 conv as implicit GEMM, stride 1 or 2 (no im2col, no materialised padding)."""

    def __init__(self, conv: torch.nn.Conv2d, cin_pad: Optional[int] = None, asym_pad: bool = False):
        w = conv.weight.detach().to(torch.float16)
        self.cout_real = w.shape[0]
        self.cout = _r8(w.shape[0])
        self.cin = cin_pad or _r8(w.shape[1])
        self.k = w.shape[2]
        self.stride = conv.stride[0]
        self.pad = conv.padding[0] if not asym_pad else 0
        self.asym = asym_pad
        self.wgt = pack_conv_weight(w, self.cin, self.cout)
        self.bias = None if conv.bias is None else pad_vector(conv.bias.detach().to(torch.float16), self.cout)

    def __call__(self, rec: Recorder, x: FM, x2: Optional[FM] = None, row_bias=None, residual: Optional[FM] = None,
                 act=nat.ACT_NONE, out_scale=1.0, out: Optional[FM] = None) -> FM:
        a, halo = x.t, {}
        if rec.patch is not None and rec.patch.rows and self.k == 3:  # rows above / below this strip live on the neighbours
            assert x2 is None and self.pad == 1 and not self.asym
            a, halo = rec.patch.conv_input(rec, x)
        if self.stride == 1:
            o = out or rec.fm(x.n, x.h, x.w, self.cout)
            rec.add(nat.igemm, a, self.wgt, o.t.view(-1, o.c), n=x.n, h=x.h, w=x.w, kh=self.k, kw=self.k, pad=self.pad,
                    a2=None if x2 is None else x2.t, bias=self.bias, row_bias=row_bias,
                    residual=None if residual is None else residual.tokens(), act=act, out_scale=out_scale,
                    splitk_scratch=rec.splitk, **halo)
            return o
        assert x2 is None and self.stride == 2
        # stride 2 as implicit GEMM: the TMA box skips every other pixel (sw_igemm `stride`), no im2col buffer.
        # UNet downsamplers pad 1 symmetrically; the VAE encoder pads (0,1,0,1) = no leading pad, and the trailing
        # pad row / column is the TMA unit's out-of-bounds zero fill either way.
        if halo:  # over the halo-padded strip: its top halo row is input row -1
            assert halo["a_y_off"] == 1 and x.h % 2 == 0
        if self.asym:
            oh, ow, pad = (x.h + 1 - self.k) // 2 + 1, (x.w + 1 - self.k) // 2 + 1, 0
        else:
            oh, ow, pad = (x.h + 2 * self.pad - self.k) // 2 + 1, (x.w + 2 * self.pad - self.k) // 2 + 1, self.pad
        o = out or rec.fm(x.n, oh, ow, self.cout)
        rec.add(nat.igemm, a, self.wgt, o.t.view(-1, o.c), n=x.n, h=oh, w=ow, kh=self.k, kw=self.k, pad=pad, stride=2,
                in_h=x.h, in_w=x.w, bias=self.bias, residual=None if residual is None else residual.tokens(), act=act,
                out_scale=out_scale, splitk_scratch=rec.splitk, **halo)
        return o


def patch_attention_mode() -> str:
    """How a row-split plan runs self-attention: "kv" = K/V of the other strips read inside the attention kernel
    (sw_flash_attn_gather, sequence parallel), "tp" = head parallel through the two fused GEMM collectives
    (TransformerBlock._self_attention_tp) where the shapes allow.  SW_PATCH_ATTN selects; default "kv"."""
    return os.environ.get("SW_PATCH_ATTN", "kv").lower()


STATS_TILE = 128  # column tile of GEMMs that emit row statistics (fixes the number of partial slots)


def fuse_layernorm(width: int = 0) -> bool:
    """Fold a transformer block's LayerNorms into the GEMMs around them (ops/packing.fold_layernorm)?
    SW_LN_FUSE = 0 never, 1 always, unset = blocks up to 640 channels.  Measured: all blocks fused gives SD1.5
    3.67 vs 3.62 img/s but SDXL 728 vs 722 ms (its sixty 1280-wide blocks lose: the GEMMs that carry the
    statistics give up CTA pairs / split-K and pay 1-2 us for the extra loads, about what a 4.5 us LayerNorm
    kernel costs); the 4-GPU patch-parallel step does not care (14.9 vs 14.7 ms with 210 fewer launches)."""
    mode = os.environ.get("SW_LN_FUSE", "auto")
    if mode in ("0", ""):
        return False
    if mode == "1":
        return True
    return 0 < width <= 640


class RowStats:
    """Per-row (sum, sum of squares) partials a producer GEMM left for the GEMM that consumes LayerNorm(y)."""

    def __init__(self, buf: torch.Tensor, tiles: int, width: int):
        self.buf, self.tiles, self.width = buf, tiles, width


class Linear:
    def __init__(self, weight, bias=None, geglu=False, ln: Optional[torch.nn.LayerNorm] = None):
        """`ln`: fold this LayerNorm (applied to the input) into the weights; calls then need `stats`."""
        w = weight.detach().to(torch.float16)
        if w.dim() == 4:
            w = w[:, :, 0, 0]
        b = None if bias is None else bias.detach().to(torch.float16)
        self.col_sums = None
        self.ln_eps = 0.0
        if ln is not None:
            w, b, cs = fold_layernorm(w, b, ln.weight, ln.bias)
            self.col_sums, self.ln_eps = cs, float(ln.eps)
        if w.shape[0] % 8 or w.shape[1] % 8:  # tiny projections (e.g. VAE post_quant 4->4): zero-pad
            assert ln is None
            wp = torch.zeros(_r8(w.shape[0]), _r8(w.shape[1]), dtype=w.dtype, device=w.device)
            wp[: w.shape[0], : w.shape[1]] = w
            w, b = wp, pad_vector(b, wp.shape[0])
        self.geglu = geglu
        self.tile_n = 0
        if geglu:
            # 128-wide tiles (64 values + 64 gates): two CTAs per SM overlap one tile's GELU epilogue with the
            # other's main loop; 256-wide tiles (half = 128) measured 30-40 % slower on the UNet shapes
            half = 64
            self.tile_n = 2 * half
            if self.col_sums is not None:  # the column sums ride through the same row interleave
                _, cs16 = interleave_geglu(w, self.col_sums, half)
                self.col_sums = cs16.float().contiguous() if cs16.dtype != torch.float32 else cs16
            w, b = interleave_geglu(w, b, half)
        self.nout = w.shape[0] // 2 if geglu else w.shape[0]
        self.wgt = w.contiguous()
        self.bias = None if b is None else b.contiguous()

    def _ln_kw(self, stats: Optional[RowStats], lo=None, hi=None):
        if self.col_sums is None:
            assert stats is None
            return {}
        assert stats is not None, "this projection has a LayerNorm folded in: it needs the producer's row statistics"
        cs = self.col_sums if lo is None else self.col_sums[lo:hi]
        return dict(row_stats_in=stats.buf, stats_tiles=stats.tiles, ln_width=stats.width, ln_eps=self.ln_eps, col_sums=cs)

    def __call__(self, rec: Recorder, x: torch.Tensor, residual=None, act=nat.ACT_NONE, out=None, out_scale=1.0,
                 stats: Optional[RowStats] = None, emit_stats: bool = False):
        """x: [rows, K] -> [rows, nout].  `stats`: row statistics of x (folded LayerNorm).  `emit_stats`: also
        return RowStats of the output for a downstream folded LayerNorm -> (out, RowStats)."""
        rows = x.shape[0]
        o = out if out is not None else rec.buf(rows, self.nout)
        kw = self._ln_kw(stats)
        tile_n = self.tile_n
        st = None
        act_code = nat.ACT_GEGLU if self.geglu else act
        if emit_stats or stats is not None:
            # GEMMs with LayerNorm plumbing bypass the binding's table lookup (pairs and split-K do not carry it):
            # take the measured tile width / persistence of the same shape when it is a single-CTA configuration
            cfg = tuning.lookup(1, 1, rows, 1, 1, x.shape[-1], 0, self.wgt.shape[0], act_code, residual is not None)
            if cfg is not None and cfg[1] != 2 and cfg[0] in (64, 128, 256) and not self.geglu:
                tile_n = cfg[0]
            if cfg is not None and stats is not None and not emit_stats:
                kw["persist"] = cfg[3]
        if emit_stats:
            assert not self.geglu
            tile = tile_n if tile_n in (64, 128, 256) else STATS_TILE
            tiles = (self.nout + tile - 1) // tile
            st = RowStats(rec.buf(tiles, rows, 2, dtype=torch.float32), tiles, self.nout)
            kw["row_stats_out"] = st.buf
            tile_n = tile
        rec.add(nat.igemm, x, self.wgt, o, n=1, h=1, w=rows, bias=self.bias, residual=residual,
                act=act_code, out_scale=out_scale, splitk_scratch=rec.splitk,
                tile_n=tile_n, **kw)
        return (o, st) if emit_stats else o

    def call_rows(self, rec: Recorder, x: torch.Tensor, lo: int, hi: int, out: torch.Tensor,
                  stats: Optional[RowStats] = None):
        """Only output columns [lo, hi) (a row block of the weight), e.g. Q or K/V of a fused QKV projection."""
        rec.add(nat.igemm, x, self.wgt[lo:hi], out, n=1, h=1, w=x.shape[0],
                bias=None if self.bias is None else self.bias[lo:hi], splitk_scratch=rec.splitk,
                **self._ln_kw(stats, lo, hi))
        return out


class GroupNorm:
    def __init__(self, gn: torch.nn.GroupNorm):
        self.groups, self.eps = gn.num_groups, gn.eps
        self.gamma = gn.weight.detach().to(torch.float16).contiguous()
        self.beta = gn.bias.detach().to(torch.float16).contiguous()

    def __call__(self, rec: Recorder, x: FM, silu: bool, x2: Optional[FM] = None, out: Optional[FM] = None,
                 dist: Optional[bool] = None) -> FM:
        """`dist`: the statistics span rows held by the other ranks of the strip group (default: every GroupNorm
        of a row-split plan; frame-split video plans say so for their temporal layers only)."""
        c = x.c + (x2.c if x2 is not None else 0)
        o = out or rec.fm(x.n, x.h, x.w, c)
        if dist is None:
            dist = rec.patch is not None and rec.patch.rows
        if dist and rec.patch is not None and rec.patch.P > 1:  # in-kernel all-reduce of the partial sums
            rec.add(nat.groupnorm_dist, x.t, self.gamma, self.beta, o.t, rec.patch.gn_peers(self.groups), n=x.n,
                    hw=x.h * x.w, groups=self.groups, eps=self.eps, silu=silu, stats=rec.gn_stats,
                    x2=None if x2 is None else x2.t)
            return o
        rec.add(nat.groupnorm, x.t, self.gamma, self.beta, o.t, n=x.n, hw=x.h * x.w, groups=self.groups, eps=self.eps,
                silu=silu, stats=rec.gn_stats, x2=None if x2 is None else x2.t, scratch_zeroed=True)
        return o


class LayerNorm:
    def __init__(self, ln: torch.nn.LayerNorm):
        self.eps = ln.eps
        self.gamma = ln.weight.detach().to(torch.float16).contiguous()
        self.beta = ln.bias.detach().to(torch.float16).contiguous()

    def __call__(self, rec: Recorder, x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        o = out if out is not None else rec.buf(*x.shape)
        rec.add(nat.layernorm, x, self.gamma, self.beta, o, self.eps)
        return o


class Resnet:
    """GN+SiLU -> conv3x3 (+time-embedding row bias) -> GN+SiLU -> conv3x3 (+shortcut residual).
    A skip tensor (UNet up path) is consumed as a second input, never concatenated in memory."""

    def __init__(self, m, temb_slice=None):
        self.norm1, self.norm2 = GroupNorm(m.norm1), GroupNorm(m.norm2)
        self.conv1, self.conv2 = Conv(m.conv1), Conv(m.conv2)
        self.shortcut = None
        if m.conv_shortcut is not None:
            self.shortcut = Linear(m.conv_shortcut.weight, m.conv_shortcut.bias)
        self.scale = 1.0 / m.output_scale_factor
        self.temb_slice = temb_slice  # (offset, width) into the fused time-embedding projection

    def __call__(self, rec: Recorder, x: FM, skip: Optional[FM] = None, temb_all=None) -> FM:
        h = self.norm1(rec, x, True, skip)
        rb = None
        if self.temb_slice is not None and temb_all is not None:
            off, width = self.temb_slice
            rb = temb_all[:, off:off + width]
        h2 = self.conv1(rec, h, row_bias=rb)
        rec.release(h)
        h3 = self.norm2(rec, h2, True)
        rec.release(h2)
        if self.shortcut is not None:
            sc = rec.fm(x.n, x.h, x.w, self.conv2.cout)
            rec.add(nat.igemm, x.t, self.shortcut.wgt, sc.tokens(), n=x.n, h=x.h, w=x.w, kh=1, kw=1, pad=0,
                    a2=None if skip is None else skip.t, bias=self.shortcut.bias, splitk_scratch=rec.splitk)
            res = sc
        else:
            assert skip is None
            res = x
        out = self.conv2(rec, h3, residual=res, out_scale=self.scale)
        # (x + h)/s == x/s + h/s: only scale != 1 needs the residual pre-scaled; SD uses s == 1.
        assert self.scale == 1.0, "output_scale_factor != 1 is not used by the supported checkpoints"
        rec.release(h3)
        if self.shortcut is not None:
            rec.release(sc)
        return out


class CrossKV:
    """Per-job precomputed K/V of one cross-attention layer (text is step-invariant)."""

    def __init__(self, attn):
        wk = attn.to_k.weight.detach().to(torch.float16)
        wv = attn.to_v.weight.detach().to(torch.float16)
        self.lin = Linear(torch.cat([wk, wv], 0))
        self.buffers = {}   # (n, s_text) -> [n, s_text, 2C]; captured graphs keep pointing at theirs
        self.kv = None      # buffer of the job being prepared


class TransformerBlock:
    """Self-attention, cross-attention, GEGLU feed-forward on the residual stream y [n*s, c].

    With `fuse_layernorm()` the three LayerNorms never run as kernels: each is folded into the projection that
    consumes it (ops/packing.fold_layernorm), and the GEMM that produces the residual stream (proj_in, the
    attention output projections, ff2) emits the row statistics in its epilogue."""

    def __init__(self, m):
        self.heads, self.dh = m.attn1.heads, m.attn1.dim_head
        c = self.heads * self.dh
        self.c = c
        a1, a2 = m.attn1, m.attn2
        self.self_is_cross = m.only_cross_attention
        self._tp = {}
        self.can_fuse = fuse_layernorm(c) and not self.self_is_cross
        self.ln1, self.ln2, self.ln3 = LayerNorm(m.norm1), LayerNorm(m.norm2), LayerNorm(m.norm3)
        if not self.self_is_cross:
            wqkv = torch.cat([a1.to_q.weight, a1.to_k.weight, a1.to_v.weight], 0)
            self.qkv = Linear(wqkv)
            if self.can_fuse:  # second set of the three consuming projections with their LayerNorm folded in
                self.qkv_f = Linear(wqkv, ln=m.norm1)
                self.q2_f = Linear(a2.to_q.weight, ln=m.norm2)
                self.ff1_f = Linear(m.ff.net[0].proj.weight, m.ff.net[0].proj.bias, geglu=True, ln=m.norm3)
        else:
            self.q1 = Linear(a1.to_q.weight)
            self.kv1 = CrossKV(a1)
        self.o1 = Linear(a1.to_out[0].weight, a1.to_out[0].bias)
        self.q2 = Linear(a2.to_q.weight)
        self.kv2 = CrossKV(a2)
        self.o2 = Linear(a2.to_out[0].weight, a2.to_out[0].bias)
        self.ff1 = Linear(m.ff.net[0].proj.weight, m.ff.net[0].proj.bias, geglu=True)
        self.ff2 = Linear(m.ff.net[2].weight, m.ff.net[2].bias)

    # ---- head-parallel (tensor-parallel) self-attention for row-split plans: the north star's "all-gather fused into
    # the QKV GEMM, out-projection GEMM fused with its reduce-scatter"  inside a model
    def tp_eligible(self, rec: Recorder, s: int) -> bool:
        ctx = rec.patch
        return (ctx is not None and ctx.rows and ctx.P > 1 and not self.self_is_cross and patch_attention_mode() == "tp"
                and self.heads % ctx.P == 0 and s % 128 == 0 and self.dh <= 128)

    def _tp_weights(self, P: int, p: int):
        key = (P, p)
        tw = self._tp.get(key)
        if tw is None:
            hp, c = self.heads // P, self.c
            lo, hi = p * hp * self.dh, (p + 1) * hp * self.dh
            w = self.qkv.wgt  # [3c, c]: q rows, k rows, v rows
            wqkv = torch.cat([w[lo:hi], w[c + lo:c + hi], w[2 * c + lo:2 * c + hi]], 0).contiguous()
            wo = self.o1.wgt[:, lo:hi].contiguous()  # this rank's K shard of the output projection
            tw = self._tp[key] = (wqkv, wo)
        return tw

    def _self_attention_tp(self, rec: Recorder, y: torch.Tensor, s: int) -> torch.Tensor:
        """y: this strip's rows [s, c] of the residual stream (not released here).  Returns y + out_proj(attention):
          LayerNorm(y) is written straight into this rank's rows of a peer-visible gathered buffer;
          sw_igemm_allgather pushes those rows to the group and computes Q/K/V of THIS rank's heads for ALL rows;
          flash attention runs locally (heads / P heads, the whole sequence);
          sw_igemm_reducescatter multiplies by this rank's K shard of W_o, pushes partial tiles to the rows' owners
          and reduces its own rows (+ bias + residual) -- no K/V gather, no separate collective."""
        ctx = rec.patch
        P, p, c = ctx.P, ctx.p, self.c
        hp = self.heads // P
        wqkv, wo = self._tp_weights(P, p)
        gathered = ctx.sym_buf(P * s, c)
        self.ln1(rec, y, out=gathered.local[p * s:(p + 1) * s])
        qkv = rec.buf(P * s, 3 * hp * self.dh)
        rec.add(nat.igemm_allgather, wqkv, qkv, ctx.tile_peers(gathered, P * (s // 128)), m_local=s, k=c, a_ld=c)
        v5 = qkv.view(1, P * s, 3, hp, self.dh)
        attn = rec.buf(P * s, hp * self.dh)
        rec.add(nat.flash_attn, v5[:, :, 0], v5[:, :, 1], v5[:, :, 2], attn.view(1, P * s, hp * self.dh), batch=1, heads=hp,
                seq_q=P * s, seq_kv=P * s, head_dim=self.dh)
        rec.release(qkv)
        staging = ctx.sym_buf(P * s, c)  # [P][s][c]: slot [src] holds rank src's partial of this rank's rows
        ntile = (c + 127) // 128
        y1 = rec.buf(s, c)
        rec.add(nat.igemm_reducescatter, attn, wo, ctx.tile_peers(staging, P * (s // 128) * ntile), y1, m_local=s, ldc=c,
                bias=self.o1.bias, residual=y)
        rec.release(attn)
        self._tp_debug = (gathered, qkv, attn, y1, wqkv, wo)  # (scripts/tp_block_check.py inspects the stages)
        return y1

    def use_fused(self, rows: int) -> bool:
        """Fold the LayerNorms for this call?  Big token counts keep the separate kernel: their GEMMs want CTA
        pairs, which the statistics plumbing rules out (ControlNet 768x768 batch 8 and text-to-video lost 3 %)."""
        return self.can_fuse and (rows <= 16384 or os.environ.get("SW_LN_FUSE") == "1")

    def cross_layers(self):
        return [self.kv2] + ([self.kv1] if self.self_is_cross else [])

    def _cross(self, rec, q, kv: CrossKV, n, s, out_lin, residual, emit):
        """q: projected queries [n*s, c] (released here); returns out_lin(attention) (+ RowStats if emit)."""
        skv = kv.kv.shape[1]
        kvv = kv.kv.view(n, skv, 2, self.heads, self.dh)
        attn = rec.buf(n, s, self.c)
        rec.add(nat.flash_attn, q.view(n, s, self.heads, self.dh), kvv[:, :, 0], kvv[:, :, 1], attn, batch=n,
                heads=self.heads, seq_q=s, seq_kv=skv, head_dim=self.dh)
        rec.release(q)
        y = out_lin(rec, attn.view(n * s, self.c), residual=residual, emit_stats=emit)
        rec.release(attn)
        return y

    def _self_attention(self, rec, x, st, n, s, qkv_lin=None):
        """x: input of the fused QKV projection (LayerNorm output, or the raw stream + its statistics)."""
        attn = rec.buf(n, s, self.c)
        lin = qkv_lin or self.qkv
        if rec.patch is not None and rec.patch.rows and rec.patch.P > 1:
            # sequence-parallel self-attention: K/V of the other strips are read from their owners inside the
            # kernel (TMA over the NVLink mappings), never gathered
            if s % 128 == 0 and self.dh <= 128:
                qs = rec.patch.sym_buf(n * s, 3 * self.c)
                lin(rec, x, out=qs.local, stats=st)
                v5 = qs.local.view(n, s, 3, self.heads, self.dh)
                kpeers, v_ptrs = rec.patch.kv_peers(qs, self.c)
                rec.add(nat.peer_signal, kpeers)
                rec.add(nat.flash_attn_gather, v5[:, :, 0], v5[:, :, 1], v5[:, :, 2], attn, kpeers, v_ptrs, batch=n,
                        heads=self.heads, seq_q=s, s_local=s, head_dim=self.dh)
            else:
                # strips too short for whole K/V tiles (deep levels): the K/V projection writes this strip's
                # rows of a full-sequence buffer and one push exchange completes it on every peer
                pp, P = rec.patch.p, rec.patch.P
                kvf = rec.patch.sym_buf(n * s * P, 2 * self.c)
                q = rec.buf(n * s, self.c)
                lin.call_rows(rec, x, 0, self.c, q, stats=st)
                mine = kvf.local[pp * s:(pp + 1) * s]
                lin.call_rows(rec, x, self.c, 3 * self.c, mine, stats=st)
                rec.patch.group_exchange(rec, mine, kvf, pp * s * 2 * self.c * 2)
                kv5 = kvf.local.view(n, s * P, 2, self.heads, self.dh)
                rec.add(nat.flash_attn, q.view(n, s, self.heads, self.dh), kv5[:, :, 0], kv5[:, :, 1], attn, batch=n,
                        heads=self.heads, seq_q=s, seq_kv=s * P, head_dim=self.dh)
                rec.release(q)
        else:
            qkv = lin(rec, x, stats=st)
            v5 = qkv.view(n, s, 3, self.heads, self.dh)
            rec.add(nat.flash_attn, v5[:, :, 0], v5[:, :, 1], v5[:, :, 2], attn, batch=n, heads=self.heads, seq_q=s,
                    seq_kv=s, head_dim=self.dh)
            rec.release(qkv)
        return attn
    def __call__(self, rec: Recorder, y: torch.Tensor, n: int, s: int, st=None, emit: bool = False):
        """y: [n*s, c] residual stream (released here).  Fused mode: `st` = RowStats of y, and with `emit` the
        new stream comes back as (y3, RowStats) for the next block.  Returns the new stream."""
        if st is not None:  # fused: y comes with its row statistics
            attn = self._self_attention(rec, y, st, n, s, self.qkv_f)
            rec.release(st.buf)
            y1, st1 = self.o1(rec, attn.view(n * s, self.c), residual=y, emit_stats=True)
            rec.release(attn)
            rec.release(y)
            q = self.q2_f(rec, y1, stats=st1)
            rec.release(st1.buf)
            y2, st2 = self._cross(rec, q, self.kv2, n, s, self.o2, y1, True)
            rec.release(y1)
            ff = self.ff1_f(rec, y2, stats=st2)
            rec.release(st2.buf)
            y3 = self.ff2(rec, ff, residual=y2, emit_stats=emit)
            rec.release(ff)
            rec.release(y2)
            return y3
        if n == 1 and self.tp_eligible(rec, s):
            y1 = self._self_attention_tp(rec, y, s)
            rec.release(y)
            x2 = self.ln2(rec, y1)
            y2 = self._cross(rec, self.q2(rec, x2), self.kv2, n, s, self.o2, y1, False)
            rec.release(x2)
            rec.release(y1)
            x3 = self.ln3(rec, y2)
            ff = self.ff1(rec, x3)
            rec.release(x3)
            y3 = self.ff2(rec, ff, residual=y2)
            rec.release(ff)
            rec.release(y2)
            return y3
        x1 = self.ln1(rec, y)
        if not self.self_is_cross:
            attn = self._self_attention(rec, x1, None, n, s)
            rec.release(x1)
            y1 = self.o1(rec, attn.view(n * s, self.c), residual=y)
            rec.release(attn)
        else:
            y1 = self._cross(rec, self.q1(rec, x1), self.kv1, n, s, self.o1, y, False)
            rec.release(x1)
        rec.release(y)
        x2 = self.ln2(rec, y1)
        y2 = self._cross(rec, self.q2(rec, x2), self.kv2, n, s, self.o2, y1, False)
        rec.release(x2)
        rec.release(y1)
        x3 = self.ln3(rec, y2)
        ff = self.ff1(rec, x3)
        rec.release(x3)
        y3 = self.ff2(rec, ff, residual=y2)
        rec.release(ff)
        rec.release(y2)
        return y3


class Transformer2D:
    def __init__(self, m):
        self.norm = GroupNorm(m.norm)
        self.proj_in = Linear(m.proj_in.weight, m.proj_in.bias)
        self.proj_out = Linear(m.proj_out.weight, m.proj_out.bias)
        self.blocks = [TransformerBlock(b) for b in m.transformer_blocks]

    def cross_layers(self):
        return [kv for b in self.blocks for kv in b.cross_layers()]

    def __call__(self, rec: Recorder, x: FM) -> FM:
        g = self.norm(rec, x, False)
        st = None
        s = x.h * x.w
        fused = all(b.use_fused(x.n * s) for b in self.blocks)
        if rec.patch is not None and os.environ.get("SW_LN_FUSE") != "1":
            fused = False  # patch-parallel plans: measured no gain there (14.9 vs 14.7 ms per step), keep the plain path
        if fused:
            y, st = self.proj_in(rec, g.tokens(), emit_stats=True)
        else:
            y = self.proj_in(rec, g.tokens())
        rec.release(g)
        for i, blk in enumerate(self.blocks):
            if fused:
                nxt = i + 1 < len(self.blocks)
                r = blk(rec, y, x.n, s, st, emit=nxt)
                y, st = r if nxt else (r, None)
            else:
                y = blk(rec, y, x.n, s)
        out = rec.fm(x.n, x.h, x.w, x.c)
        self.proj_out(rec, y, residual=x.tokens(), out=out.tokens())
        rec.release(y)
        return out


class Upsample:
    """Nearest 2x + conv3x3, folded: four 2x2 convs on the small map write the four output phases directly
    (ops/packing.fold_upsample2x) -- no upsampled buffer, 2.25x fewer multiply-adds."""

    def __init__(self, m):
        self.conv = Conv(m.conv)
        w = m.conv.weight.detach().to(torch.float16)
        self.phases = fold_upsample2x(w, self.conv.cin, self.conv.cout)
        # all four phases in one launch (grid z = phase) when the kernel's phase mode applies
        self.stacked = torch.cat([self.phases[(a, b)] for a in (0, 1) for b in (0, 1)], 0).contiguous() \
            if self.conv.cout % 64 == 0 else None

    def __call__(self, rec: Recorder, x: FM) -> FM:
        a, halo = x.t, {}
        if rec.patch is not None and rec.patch.rows:  # rows y-1 / y+1 of a strip's edge live on the neighbours
            a, halo = rec.patch.conv_input(rec, x)
        mode = os.environ.get("SW_UPSAMPLE", "fold")  # fold (default) | phases4 (one launch per phase) | materialise
        if mode == "materialise":  # A/B path: the nearest-2x copy in memory, then the 3x3 conv on it
            up = rec.fm(x.n, 2 * x.h, 2 * x.w, x.c)
            rec.add(nat.upsample2x, x.t, up.t, n=x.n, h=x.h, w=x.w)
            o = self.conv(rec, up)
            rec.release(up)
            return o
        o = rec.fm(x.n, 2 * x.h, 2 * x.w, self.conv.cout)
        if self.stacked is not None and mode != "phases4":
            rec.add(nat.igemm, a, self.stacked, o.t, n=x.n, h=x.h, w=x.w, kh=2, kw=2, pad=1, bias=self.conv.bias,
                    out_sx=2, out_sy=2, out_phases=1, ldc=o.c, **halo)
            return o
        for (pa, pb), wgt in self.phases.items():
            rec.add(nat.igemm, a, wgt, o.t[:, pa::2, pb::2, :], n=x.n, h=x.h, w=x.w, kh=2, kw=2, pad=1 - pa, pad_w=1 - pb,
                    bias=self.conv.bias, out_sx=2, out_sy=2, ldc=o.c, **halo)
        return o
