"""Patch-parallel denoiser: ONE image's UNet spread over 2 / 4 / 8 B200s (BASELINE config 3, SURVEY.md
section 2.3 "patch / sequence parallel", kernel rows W9 / W10).  The reference runs a job on a single
GPU (swarm/worker.py:41-43); this is the latency mode that replaces it for large single images.

Layout of the gang (world = G x P):
  * G = n_cond classifier-free-guidance branches run on disjoint GPU groups (branch g on ranks
    [g*P, (g+1)*P)); the branches only meet in the per-step epsilon exchange;
  * inside a group the latent is cut into P horizontal strips, rank p owns rows [p*H/P, (p+1)*H/P) of
    every feature map, so each rank always holds a batch of exactly one strip.

Every rank records the same static plan as the single-GPU engine (engine/unet_sm100.py) -- the layers
consult `rec.patch` at the four places where a strip needs its peers, and all of that traffic is issued
from inside kernels over the symmetric-memory arena (parallel/runtime.py), never through NCCL:

  conv 3x3     feature maps live in [h+2, w, c] buffers; `sw_peer_exchange` pushes the boundary rows into
               the neighbours' halo rows, the implicit-GEMM kernel then reads rows -1 / h through TMA
               (a_h_total / a_y_off); image borders keep their zero halo = the conv's zero padding
  GroupNorm    `sw_groupnorm_nhwc_dist`: partial sums pushed to every peer, folded in rank order
  self-attn    `sw_flash_attn_gather`: K/V tiles of the other strips TMA-loaded from their owners
  epsilon      conv_out writes this strip's slice of eps[n_cond, H, W, c]; one exchange makes it whole on
               every rank, and every rank runs the (tiny) fused CFG + scheduler kernel redundantly

Epochs: each channel (halo / GroupNorm / K-V / epsilon) owns a flag row; sync point i of a channel uses
epoch i + tick * K, K = sync points per step and `tick` a device counter bumped once per step, so the
captured CUDA graph replays with monotonically increasing epochs.  Peer-visible buffers are never
recycled inside a step; the epsilon exchange is the step barrier that makes their reuse safe.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from ..ops import native as nat
from ..parallel.runtime import Gang, SymmetricArena, SymTensor
from .denoiser import MAX_STEPS, NOISE_STEPS_MIN, SM100Denoiser, _StepGraph
from .layers import FM, Recorder, _r8


class _Channel:
    def __init__(self, flags: SymTensor):
        self.flags = flags
        self.structs: List[nat.Peers] = []


class PatchContext:
    def __init__(self, gang: Gang, n_cond: int, arena, rows: bool = True):
        """`arena`: a SymmetricArena or one recyclable ArenaSlice of the gang's arena (PatchDenoiser hands every
        cached plan its own slice, so the peer-visible memory of an evicted plan is reused).
        `rows`: True = every feature map is cut into horizontal strips (images); False = a video clip is cut
        along the FRAME axis -- spatial layers are then local to a rank and only the temporal layers
        (engine/unet3d_sm100.py) talk to the strip group."""
        self.rows = rows
        if gang.world % n_cond:
            raise ValueError(f"{gang.world} ranks cannot be split into {n_cond} guidance groups")
        self.gang = gang
        self.world, self.rank = gang.world, gang.rank
        self.G = n_cond
        self.P = gang.world // n_cond
        self.g, self.p = gang.rank // self.P, gang.rank % self.P
        self.arena = SymmetricArena(gang, arena) if isinstance(arena, int) else arena
        self.tick = torch.zeros(1, dtype=torch.int32, device=gang.device)
        self._channels: Dict[str, _Channel] = {}
        self._gn_stats: Dict[int, SymTensor] = {}
        self._tile_syncs = 0

    # ---- addressing
    def global_rank(self, d: int) -> int:
        return self.g * self.P + d

    def _channel(self, name: str) -> _Channel:
        ch = self._channels.get(name)
        if ch is None:
            ch = self._channels[name] = _Channel(self.arena.alloc_flags(8))
        return ch

    def _peers(self, name: str, peer_ptrs, scope_world: bool = False) -> nat.Peers:
        """Next sync point of channel `name`: group scope (P strips) or the whole gang."""
        ch = self._channel(name)
        n = self.world if scope_world else self.P
        me = self.rank if scope_world else self.p
        ranks = list(range(self.world)) if scope_world else [self.global_rank(d) for d in range(self.P)]
        flag_ptrs = [ch.flags.peer_ptr(r) for r in ranks]
        ps = nat.make_peers(n, me, len(ch.structs) + 1, peer_ptrs, flag_ptrs, epoch_base=self.tick, epoch_mul=0)
        ch.structs.append(ps)
        return ps

    def finalize(self):
        """All sync points are known: epochs advance by the per-step count of their channel."""
        for ch in self._channels.values():
            for ps in ch.structs:
                ps.epoch_mul = len(ch.structs)
        if self._tile_syncs:  # freshly zeroed per-tile flag arrays: nobody may push into one that is not cleared yet
            torch.cuda.synchronize()
            self.gang.barrier()

    # ---- buffers
    def sym_buf(self, rows: int, c: int) -> SymTensor:
        return self.arena.alloc(rows, c)

    def owns(self, t: torch.Tensor) -> bool:
        return self.arena.contains(t)

    # ---- the four communication points
    def conv_input(self, rec: Recorder, x: FM):
        if x.window is not None:
            total, first = x.window
            return x.t, dict(a_h_total=total, a_y_off=first)
        if x.sym is None or x.n != 1:
            raise ValueError("patch-parallel 3x3 convs need a halo-padded single-strip input")
        st, h = x.sym, x.h
        if self.P > 1:
            row = x.w * x.c * 2
            xfers = []
            if self.p > 0:  # my first row -> bottom halo of the strip above
                xfers.append((self.p - 1, st.local[1], st.peer_ptr(self.global_rank(self.p - 1)) + (h + 1) * row))
            if self.p < self.P - 1:  # my last row -> top halo of the strip below
                xfers.append((self.p + 1, st.local[h], st.peer_ptr(self.global_rank(self.p + 1))))
            rec.add(nat.peer_exchange, self._peers("halo", None), xfers)
        return st.local.view(1, h + 2, x.w, x.c), dict(a_h_total=h + 2, a_y_off=1)

    def gn_peers(self, groups: int) -> nat.Peers:
        st = self._gn_stats.get(groups)  # one double-buffered slot set per group count (32 spatial, 8 temporal conv)
        if st is None:
            st = self._gn_stats[groups] = self.arena.alloc(2 * self.P * groups * 2, dtype=torch.float32)
        return self._peers("gn", [st.peer_ptr(self.global_rank(d)) for d in range(self.P)])

    def tile_peers(self, buf: SymTensor, flag_count: int) -> nat.Peers:
        """Peers of one fused GEMM collective (sw_igemm_allgather / sw_igemm_reducescatter): `buf` on every rank of
        the strip group plus a per-tile flag array of its own, epoch = step counter + 1 (one use per step)."""
        flags = self.arena.alloc(flag_count, dtype=torch.int32)
        flags.local.zero_()  # (an evicted plan may have left large epochs here; finalize() barriers before any use)
        ranks = [self.global_rank(d) for d in range(self.P)]
        ps = nat.make_peers(self.P, self.p, 1, [buf.peer_ptr(r) for r in ranks], [flags.peer_ptr(r) for r in ranks],
                            epoch_base=self.tick, epoch_mul=1)
        self._tile_syncs += 1
        return ps

    def kv_peers(self, qkv: SymTensor, c: int):
        ranks = [self.global_rank(d) for d in range(self.P)]
        kp = self._peers("kv", [qkv.peer_ptr(r) + 2 * c for r in ranks])
        return kp, [qkv.peer_ptr(r) + 4 * c for r in ranks]

    def group_exchange(self, rec: Recorder, local: torch.Tensor, sym: SymTensor, offset: int):
        """All-gather inside the strip group: `local` lands at `offset` of `sym` on the other P-1 strips."""
        ps = self._peers("kvpush", None)
        xfers = [(d, local, sym.peer_ptr(self.global_rank(d)) + offset) for d in range(self.P) if d != self.p]
        rec.add(nat.peer_exchange, ps, xfers)

    def all_exchange(self, rec: Recorder, local: Optional[torch.Tensor], sym: Optional[SymTensor], offset: int):
        """Every rank sends `local` (or nothing: a pure barrier) to the same offset of `sym` on all others."""
        ps = self._peers("eps", None, scope_world=True)
        xfers = []
        for r in range(self.world):
            if r == self.rank:
                continue
            xfers.append((r, None, 0) if local is None else (r, local, sym.peer_ptr(r) + offset))
        rec.add(nat.peer_exchange, ps, xfers)


class PatchRecorder(Recorder):
    def __init__(self, device, ctx: PatchContext):
        super().__init__(device)
        self.patch = ctx

    def fm(self, n, h, w, c) -> FM:
        if n != 1:
            raise ValueError("patch-parallel plans hold one strip per rank")
        st = self.patch.arena.alloc(h + 2, w, c)
        return FM(st.local[1:h + 1].view(1, h, w, c), 1, h, w, c, sym=st)

    def release(self, t):
        if isinstance(t, FM):
            if t.sym is not None or t.window is not None:
                return
            t = t.t
        if self.patch.owns(t):
            return  # peers may still be reading it: peer-visible memory is only reused across steps
        super().release(t)


class FrameRecorder(Recorder):
    """Recorder of a frame-split video plan: feature maps are ordinary local buffers (whole frames live on one
    rank); only what the temporal layers exchange is carved from the symmetric arena."""

    def __init__(self, device, ctx: PatchContext):
        super().__init__(device)
        self.patch = ctx

    def halo_fm(self, nf: int, hw: int, c: int) -> FM:
        """[nf + 2, hw, c] peer-visible buffer viewed as the (1, nf, hw, c) map a (3,1,1) conv reads: frames -1 and
        nf are filled by the neighbouring ranks (zero at the ends of the clip = the conv's padding)."""
        st = self.patch.arena.alloc(nf + 2, hw, c)
        return FM(st.local[1:nf + 1].view(1, nf, hw, c), 1, nf, hw, c, sym=st)

    def release(self, t):
        if isinstance(t, FM):
            if t.sym is not None:
                return
            t = t.t
        if self.patch.owns(t):
            return
        super().release(t)


class _PatchStepGraph(_StepGraph):
    def __init__(self, eng: "PatchDenoiser", h, w, n_cond, extra_ch, has_noise, aug, arena):
        super().__init__(eng, 1, h, w, n_cond, extra_ch, False, has_noise, aug)
        self.arena_slice = arena
        self.ctx = PatchContext(eng.gang, n_cond, arena)
        if aug:
            self.aug_full = self.aug_emb                      # [n_cond, D]: _load_job fills all branches
            self.aug_emb = self.aug_full[self.ctx.g:self.ctx.g + 1]

    def build(self, max_noise_steps: int):
        eng, up, ctx = self.eng, self.eng.unet_plan, self.ctx
        H, W = self.h, self.w
        levels = len(up.boc)
        if H % ctx.P or (H // ctx.P) % (1 << (levels - 1)):
            raise ValueError(f"latent height {H} does not split into {ctx.P} strips at every UNet level")
        hl = H // ctx.P
        if self.has_noise:
            self.noise = torch.zeros(max_noise_steps, self.rows, self.lat_c, dtype=torch.float32, device=eng.device)
        rec = PatchRecorder(eng.device, ctx)
        x_in = FM(self.model_in[ctx.g], 1, hl, W, up.in_pad, window=(H, ctx.p * hl))
        temb = up.record_time_embedding(rec, 1, self.t_table, self.step, self.aug_emb)
        mid, skips = up.record_encoder(rec, x_in, temb)
        cpad = _r8(up.out_ch)
        self.eps_all = ctx.arena.alloc(self.n_cond, H, W, cpad)
        mine = self.eps_all.local[ctx.g, ctx.p * hl:(ctx.p + 1) * hl]
        up.record_decoder(rec, mid, skips, temb, out=FM(mine.view(1, hl, W, cpad), 1, hl, W, cpad))
        if ctx.world > 1:
            # barrier first: nobody overwrites a peer's eps while that peer still runs the previous
            # step's scheduler kernel on it; then the all-gather push
            ctx.all_exchange(rec, None, None, 0)
            off = ((ctx.g * H + ctx.p * hl) * W * cpad) * 2
            ctx.all_exchange(rec, mine, self.eps_all, off)
        rec.add(self._sched, FM(self.eps_all.local, self.n_cond, H, W, cpad))
        rec.add(nat.advance_counter, self.step)
        rec.add(nat.advance_counter, ctx.tick)
        ctx.finalize()
        self.rec = rec
        return self


class _FrameStepGraph(_StepGraph):
    """Text-to-video over a gang (BASELINE config 5; reference call site swarm/video/tx2vid.py:55): guidance branch
    g on ranks [g*P, (g+1)*P), and inside a branch rank p owns frames [p*F/P, (p+1)*F/P) of the clip.  Spatial
    convs / GroupNorms / attention see whole frames and never communicate; the temporal layers do:
      temporal conv (3,1,1)   one halo FRAME from each neighbour (the same push + flag as an image strip's halo row)
      its GroupNorm, and the temporal transformer's: statistics over the whole clip -> sw_groupnorm_nhwc_dist
      temporal attention      K/V of this rank's frames pushed into every peer's full-clip K/V buffer
    followed by the same epsilon exchange + redundant scheduler step as the image plan."""

    def __init__(self, eng: "PatchDenoiser", frames, h, w, n_cond, has_noise, arena):
        super().__init__(eng, frames, h, w, n_cond, 0, False, has_noise, False)
        self.arena_slice = arena
        self.ctx = PatchContext(eng.gang, n_cond, arena, rows=False)

    def build(self, max_noise_steps: int):
        eng, up, ctx = self.eng, self.eng.unet_plan, self.ctx
        F, H, W = self.batch, self.h, self.w
        if F % ctx.P:
            raise ValueError(f"{F} frames do not split over {ctx.P} ranks")
        fl = F // ctx.P
        if self.has_noise:
            self.noise = torch.zeros(max_noise_steps, self.rows, self.lat_c, dtype=torch.float32, device=eng.device)
        rec = FrameRecorder(eng.device, ctx)
        up.num_frames, up.total_frames = fl, F
        lo = ctx.g * F + ctx.p * fl
        x_in = FM(self.model_in[lo:lo + fl], fl, H, W, up.in_pad)
        temb = up.record_time_embedding(rec, fl, self.t_table, self.step, None)
        mid, skips = up.record_encoder(rec, x_in, temb)
        cpad = _r8(up.out_ch)
        self.eps_all = ctx.arena.alloc(self.n_cond, F, H, W, cpad)
        mine = self.eps_all.local[ctx.g, ctx.p * fl:(ctx.p + 1) * fl]
        up.record_decoder(rec, mid, skips, temb, out=FM(mine, fl, H, W, cpad))
        if ctx.world > 1:
            ctx.all_exchange(rec, None, None, 0)
            off = ((ctx.g * F + ctx.p * fl) * H * W * cpad) * 2
            ctx.all_exchange(rec, mine, self.eps_all, off)
        rec.add(self._sched, FM(self.eps_all.local.view(self.n_cond * F, H, W, cpad), self.n_cond * F, H, W, cpad))
        rec.add(nat.advance_counter, self.step)
        rec.add(nat.advance_counter, ctx.tick)
        ctx.finalize()
        self.rec = rec
        return self


class PatchDenoiser(SM100Denoiser):
    """SM100Denoiser whose single-image jobs run patch-parallel over `gang`; every rank of the gang
    must make the same calls with the same seeds (SPMD).  Other jobs fall back to the local engine."""

    def __init__(self, pipe, gang: Gang, arena_bytes: int = 16 << 30, max_patch_graphs: int = 2):
        super().__init__(pipe)
        self.gang = gang
        self.arena_bytes = arena_bytes
        # ONE symmetric arena per gang, cut into recyclable slices (one per cached patch plan, least recently used
        # evicted): a stream of jobs with new shapes can neither exhaust HBM nor pay a rendezvous per plan
        self.max_patch_graphs = max_patch_graphs
        self._slices = None

    def _take_slice(self):
        if self._slices is None:
            self._slices = SymmetricArena(self.gang, self.arena_bytes).slices(self.max_patch_graphs)
        free = [s for s in self._slices if not s.in_use]
        if not free:
            victim = next(k for k in self._graphs if isinstance(k, tuple) and k and k[0] == "patch")
            self._evict(victim)
            free = [s for s in self._slices if not s.in_use]
        sl = free[0]
        sl.reset()
        sl.in_use = True
        return sl

    def _evict(self, key):
        sg = self._graphs.get(key)
        if sg is not None and getattr(sg, "arena_slice", None) is not None:
            sg.arena_slice.in_use = False
        super()._evict(key)

    def supports(self, x, n_cond, control, inpaint) -> bool:
        return (self.gang.world > 1 and x.shape[0] == 1 and control is None and inpaint is None
                and self.gang.world % n_cond == 0)

    @torch.no_grad()
    def run(self, x, embeds, plan, n_cond, g_text, g_image, extra, control, control_scale, added, noise, inpaint,
            num_frames: int = 1):
        if num_frames > 1 and self.supports_frames(x, n_cond, num_frames, extra, control, added, inpaint):
            return self._run_frames(x, embeds, plan, n_cond, g_text, noise, num_frames)
        if not self.supports(x, n_cond, control, inpaint) or num_frames != 1:
            return super().run(x, embeds, plan, n_cond, g_text, g_image, extra, control, control_scale, added, noise,
                               inpaint, num_frames)
        batch, c, h, w = x.shape
        steps = plan.num_steps
        if steps > MAX_STEPS:
            raise ValueError(f"at most {MAX_STEPS} denoising steps are supported")
        extra_ch = 0 if extra is None else extra.shape[1]
        # shapes only (guidance is a device slot, see SM100Denoiser.run)
        key = ("patch", h, w, n_cond, extra_ch, noise is not None, added is not None)
        sg = self._graphs.pop(key, None)
        fresh = sg is None
        if fresh:
            sg = _PatchStepGraph(self, h, w, n_cond, extra_ch, noise is not None, added is not None, self._take_slice())
        self._graphs[key] = sg  # most recently used goes last
        g = sg.ctx.g
        self.unet_plan.set_context(embeds[g:g + 1].to(torch.float16))
        if fresh:
            sg.build(max(steps, NOISE_STEPS_MIN) if noise is not None else 0)
            self._load_patch_job(sg, x, plan, extra, added, noise, g_text, g_image)
            sg.capture()
        self._load_patch_job(sg, x, plan, extra, added, noise, g_text, g_image)
        for _ in range(steps):
            sg.replay()
        self.last_launches = sg.launches_per_step * steps
        return sg.x.view(batch, h, w, c).permute(0, 3, 1, 2).contiguous()

    def supports_frames(self, x, n_cond, num_frames, extra, control, added, inpaint) -> bool:
        w = self.gang.world
        return (w > 1 and hasattr(self.unet_plan, "num_frames") and x.shape[0] == num_frames and w % n_cond == 0
                and num_frames % (w // n_cond) == 0 and extra is None and control is None and added is None
                and inpaint is None)

    @torch.no_grad()
    def _run_frames(self, x, embeds, plan, n_cond, g_text, noise, num_frames):
        """One clip ([F, c, h, w] latents, `embeds` one row per frame, branch-major) frame-split over the gang."""
        F, c, h, w = x.shape
        steps = plan.num_steps
        if steps > MAX_STEPS:
            raise ValueError(f"at most {MAX_STEPS} denoising steps are supported")
        key = ("patch", "frames", F, h, w, n_cond, noise is not None)
        sg = self._graphs.pop(key, None)
        fresh = sg is None
        if fresh:
            sg = _FrameStepGraph(self, F, h, w, n_cond, noise is not None, self._take_slice())
        self._graphs[key] = sg
        ctx = sg.ctx
        fl = F // ctx.P
        lo = ctx.g * F + ctx.p * fl
        self.unet_plan.set_context(embeds[lo:lo + fl].to(torch.float16))  # this rank's frames of its branch
        if fresh:
            sg.build(max(steps, NOISE_STEPS_MIN) if noise is not None else 0)
            self._load_job(sg, x, plan, None, None, None, noise, g_text, 0.0)
            sg.capture()
        self._load_job(sg, x, plan, None, None, None, noise, g_text, 0.0)
        for _ in range(steps):
            sg.replay()
        self.last_launches = sg.launches_per_step * steps
        return sg.x.view(F, h, w, c).permute(0, 3, 1, 2).contiguous()

    def _load_patch_job(self, sg: _PatchStepGraph, x, plan, extra, added, noise, g_text=7.5, g_image=1.5):
        local_aug = sg.aug_emb
        if added is not None:
            sg.aug_emb = sg.aug_full
        try:
            self._load_job(sg, x, plan, extra, None, added, noise, g_text, g_image)
        finally:
            sg.aug_emb = local_aug
