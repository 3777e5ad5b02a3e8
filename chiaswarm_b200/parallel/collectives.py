"""The four communication-bound paths, each in two forms (SURVEY.md section 2.4 "collective call
sites"; the reference has none -- grep over /root/reference finds no torch.distributed):

  * `*_nccl`  -- NCCL collective + local kernel: the baseline, not the product;
  * `*_fused` -- ONE sm_100a kernel (plus at most a tiny owner-side reduce) that moves the data
                 itself over NVLink peer mappings while it computes.

    kv_gather_attention   W9   cross-patch K/V all-gather inside the flash-attention KV loop
    groupnorm_allreduce   W10  GroupNorm statistics all-reduce inside the GroupNorm kernels
    allgather_gemm        W11  all-gather inside the A-operand TMA loads of the GEMM
    gemm_reducescatter    W12  GEMM epilogue pushes partial rows to their owners

All tensors that peers touch live in a `SymmetricArena`; epoch flags order producers/consumers.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from ..ops import native as nat
from .runtime import Gang, SymmetricArena, SymTensor


class FusedOps:
    """Per-gang context: flag sets, tickets and scratch shared by the fused kernels."""

    def __init__(self, arena: SymmetricArena):
        self.arena = arena
        self.gang = arena.gang
        self.world, self.rank = arena.world, arena.rank
        self.flags = {name: arena.alloc_flags() for name in ("ag", "rs", "kv", "gn")}
        self._tile_flags = {}   # (kind, count) -> per-tile flag arrays of the second-generation GEMM collectives
        self._gather = {}       # (m_local, k) -> two gathered buffers, alternated by call parity
        self._stage = {}        # (m_local, n) -> two staging buffers, alternated by call parity
        self._calls = {"ag": 0, "rs": 0}
        self.ticket = torch.zeros(4, dtype=torch.int32, device=self.gang.device)
        self.gn_scratch = torch.zeros(1 << 20, dtype=torch.float32, device=self.gang.device)
        self._gn_slots = {}

    def peers(self, buf: Optional[SymTensor], flags: SymTensor, epoch: int) -> nat.Peers:
        return nat.make_peers(self.world, self.rank, epoch, None if buf is None else buf.peer_ptrs(),
                              flags.peer_ptrs())

    def _flags_for(self, kind: str, count: int) -> SymTensor:
        fl = self._tile_flags.get((kind, count))
        if fl is None:
            fl = self._tile_flags[(kind, count)] = self.arena.alloc_flags(count)
        return fl

    # ------------------------------------------------------------------ W11, second generation
    def gather_buffer(self, m_local: int, k: int) -> SymTensor:
        """The gathered activation [world*m_local, k] the NEXT allgather_gemm call will use.  The producer of the
        shard writes rows [rank*m_local, (rank+1)*m_local) of `.local` directly (no staging copy); two buffers
        alternate so a peer still reading call n cannot be overwritten by a fast rank's call n+1."""
        pair = self._gather.get((m_local, k))
        if pair is None:
            pair = self._gather[(m_local, k)] = [self.arena.alloc(self.world * m_local, k) for _ in range(2)]
        return pair[self._calls["ag"] & 1]

    def allgather_gemm(self, gathered: SymTensor, wgt, out, bias=None, tile_n=0):
        """ONE kernel: copy role pushes this rank's rows of `gathered` to every peer, compute blocks wait per tile."""
        rows, k = gathered.local.shape
        m_local = rows // self.world
        tiles = m_local // 128
        fl = self._flags_for("ag", self.world * tiles)
        self._calls["ag"] += 1
        pf = self.peers(gathered, fl, self.arena.next_epoch())
        nat.igemm_allgather(wgt, out, pf, m_local=m_local, k=k, a_ld=gathered.local.stride(0), bias=bias, tile_n=tile_n)
        return out

    # ------------------------------------------------------------------ W12, second generation
    def gemm_reducescatter(self, x_kshard, wgt_kshard, out, bias=None, residual=None, tile_n=0):
        """ONE kernel: partial tiles TMA-stored to their owners, reduce role folds them tile by tile."""
        m_local, n = out.shape
        pair = self._stage.get((m_local, n))
        if pair is None:
            pair = self._stage[(m_local, n)] = [self.arena.alloc(self.world, m_local, n) for _ in range(2)]
        staging = pair[self._calls["rs"] & 1]
        self._calls["rs"] += 1
        bn = tile_n or 128
        ntile = (n + bn - 1) // bn
        fl = self._flags_for("rs", self.world * (m_local // 128) * ntile)
        pf = self.peers(staging, fl, self.arena.next_epoch())
        nat.igemm_reducescatter(x_kshard, wgt_kshard, pf, out, m_local=m_local, ldc=n, bias=bias, residual=residual,
                                tile_n=tile_n)
        return out

    # ------------------------------------------------------------------ W11 (first generation: pull through peer TMA)
    def allgather_gemm_fused(self, x: SymTensor, wgt, out, bias=None, act=nat.ACT_NONE):
        """x: this rank's [m_local, K] shard (symmetric). out: [world*m_local, N]."""
        m_local, k = x.local.shape
        e = self.arena.next_epoch()
        pf = self.peers(x, self.flags["ag"], e)
        nat.peer_signal(pf)  # publish the shard (orders all earlier writes on this stream)
        nat.igemm_ag(wgt, out, pf, m_local=m_local, k=k, bias=bias, act=act)
        return out

    def allgather_gemm_nccl(self, x_local, wgt, out, bias=None, act=nat.ACT_NONE):
        m_local, k = x_local.shape
        full = torch.empty(self.world * m_local, k, dtype=x_local.dtype, device=x_local.device)
        dist.all_gather_into_tensor(full, x_local.contiguous(), group=self.gang.group)
        nat.linear(full, wgt, out=out, bias=bias, act=act)
        return out

    # ------------------------------------------------------------------ W12
    def gemm_reducescatter_fused(self, x_kshard, wgt_kshard, staging: SymTensor, out, bias=None, residual=None):
        """x_kshard: [world*m_local, K/world]; wgt_kshard: [N, K/world]; staging: symmetric
        [world, m_local, N]; out: this rank's [m_local, N] rows."""
        _, m_local, n = staging.local.shape
        e = self.arena.next_epoch()
        pf = self.peers(staging, self.flags["rs"], e)
        nat.igemm_rs(x_kshard, wgt_kshard, pf, self.ticket[0:1], m_local=m_local, ldc=n)
        nat.rs_reduce(staging.local.view(self.world * m_local, n), out, pf, m_local=m_local, ncols=n, bias=bias,
                      residual=residual)
        return out

    def gemm_reducescatter_nccl(self, x_kshard, wgt_kshard, out, bias=None, residual=None):
        partial = nat.linear(x_kshard, wgt_kshard)
        dist.reduce_scatter_tensor(out, partial, group=self.gang.group)
        if bias is not None:
            out += bias
        if residual is not None:
            out += residual
        return out

    # ------------------------------------------------------------------ W9
    def kv_gather_attention_fused(self, q, k: SymTensor, v: SymTensor, out, *, batch, heads, head_dim):
        """q: [b, sq_local, h, d]; k, v: symmetric [b, s_local, h, d] shards."""
        s_local = k.local.shape[1]
        e = self.arena.next_epoch()
        pf = self.peers(k, self.flags["kv"], e)
        nat.peer_signal(pf)  # K and V of this rank are complete
        nat.flash_attn_gather(q, k.local, v.local, out, pf, v.peer_ptrs(), batch=batch, heads=heads,
                              seq_q=q.shape[1], s_local=s_local, head_dim=head_dim)
        return out

    def kv_gather_attention_nccl(self, q, k_local, v_local, out, *, batch, heads, head_dim):
        s_local = k_local.shape[1]
        ks = [torch.empty_like(k_local) for _ in range(self.world)]
        vs = [torch.empty_like(v_local) for _ in range(self.world)]
        dist.all_gather(ks, k_local.contiguous(), group=self.gang.group)
        dist.all_gather(vs, v_local.contiguous(), group=self.gang.group)
        kf, vf = torch.cat(ks, dim=1), torch.cat(vs, dim=1)
        nat.flash_attn(q, kf, vf, out, batch=batch, heads=heads, seq_q=q.shape[1], seq_kv=s_local * self.world,
                       head_dim=head_dim)
        return out

    # ------------------------------------------------------------------ W10
    def groupnorm_fused(self, x, gamma, beta, out, *, n, hw_local, groups, eps, silu, x2=None):
        key = (n, groups)
        slot = self._gn_slots.get(key)
        if slot is None:
            slot = self._gn_slots[key] = self.arena.alloc(2, self.world, n, groups, 2, dtype=torch.float32)
        e = self.arena.next_epoch()
        pf = self.peers(slot, self.flags["gn"], e)
        nat.groupnorm_dist(x, gamma, beta, out, pf, n=n, hw=hw_local, groups=groups, eps=eps, silu=silu,
                           stats=self.gn_scratch, x2=x2)
        return out

    def groupnorm_nccl_native(self, x, gamma, beta, out, *, n, hw_local, groups, eps, silu):
        """Like-for-like baseline: native partial-sum kernel, NCCL all-reduce of the sums, native apply kernel."""
        c = x.shape[-1]
        sums = nat.gn_partial_sums(x, n=n, hw=hw_local, groups=groups)        # fp32 [n, groups, 2]
        dist.all_reduce(sums, group=self.gang.group)
        nat.gn_apply_sums(x, sums, gamma, beta, out, n=n, hw=hw_local, groups=groups, count=hw_local * self.world * (c // groups),
                          eps=eps, silu=silu)
        return out

    def groupnorm_nccl(self, x, gamma, beta, out, *, n, hw_local, groups, eps, silu):
        """Baseline: local partial sums (torch), NCCL all-reduce, local apply (torch)."""
        c = x.shape[-1]
        xf = x.view(n, hw_local, groups, c // groups).float()
        stats = torch.stack([xf.sum(dim=(1, 3)), (xf * xf).sum(dim=(1, 3))], dim=-1)
        dist.all_reduce(stats, group=self.gang.group)
        cnt = hw_local * self.world * (c // groups)
        mean = stats[..., 0] / cnt
        var = (stats[..., 1] / cnt - mean * mean).clamp_min(0)
        y = (xf - mean[:, None, :, None]) * torch.rsqrt(var + eps)[:, None, :, None]
        y = y.view(n, hw_local, c) * gamma.float() + beta.float()
        if silu:
            y = torch.nn.functional.silu(y)
        out.copy_(y.view_as(out).to(out.dtype))
        return out
