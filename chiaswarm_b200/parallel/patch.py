"""Patch (row) parallel UNet -- the NCCL-only baseline path.

Every rank owns H/world rows of the latent at every UNet level.  What has to cross ranks:
  * conv3x3: one halo row from each neighbour (send/recv);      stride-2 conv: the top halo only
  * GroupNorm: (sum, sumsq) per (image, group)                  (all-reduce)
  * self-attention: every rank needs all K/V                    (all-gather)
everything else (1x1 convs, linears, cross-attention to the replicated text, LayerNorm, GEGLU,
nearest upsample, time embedding) is local.  This module expresses exactly that algebra with
torch ops + torch.distributed collectives (NCCL on GPUs, gloo in the CPU tests); the fused
sm_100a kernels in `collectives.py` replace the three communication sites one by one.  It is the
correctness reference for the partitioning and the "baseline, not the product" timing.

`PatchParallelVAEDecoder` is the same algebra for the image decoder (collective call site 6): conv halos,
GroupNorm statistics, and the single-head mid-block attention with local query rows against gathered K / V; every rank
decodes its strip of the image and the strips are gathered once, as pixels.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn.functional as F

from .runtime import Gang


class PatchOps:
    def __init__(self, gang: Gang):
        self.gang = gang
        self.rank, self.world, self.group = gang.rank, gang.world, gang.group

    # ---- halo exchange: returns x with `top` rows from the previous rank and `bottom` rows from the
    # next rank attached (zeros at the outer image border, which is the conv's zero padding)
    def with_halo(self, x, top=1, bottom=1):
        n, c, h, w = x.shape
        up = torch.zeros(n, c, top, w, dtype=x.dtype, device=x.device)
        down = torch.zeros(n, c, bottom, w, dtype=x.dtype, device=x.device)
        if self.world > 1:
            ops = []
            first, last = x[:, :, :bottom].contiguous(), x[:, :, h - top:].contiguous()
            if self.rank > 0:
                if bottom:
                    ops.append(dist.P2POp(dist.isend, first, self._peer(self.rank - 1), self.group))
                if top:
                    ops.append(dist.P2POp(dist.irecv, up, self._peer(self.rank - 1), self.group))
            if self.rank < self.world - 1:
                if top:
                    ops.append(dist.P2POp(dist.isend, last, self._peer(self.rank + 1), self.group))
                if bottom:
                    ops.append(dist.P2POp(dist.irecv, down, self._peer(self.rank + 1), self.group))
            if ops:
                for req in dist.batch_isend_irecv(ops):
                    req.wait()
        parts = ([up] if top else []) + [x] + ([down] if bottom else [])
        return torch.cat(parts, dim=2)

    def _peer(self, r):
        return dist.get_global_rank(self.group, r) if self.group is not None and self.group is not dist.group.WORLD else r

    def conv(self, conv: torch.nn.Conv2d, x):
        k, s = conv.kernel_size[0], conv.stride[0]
        if k == 1:
            return conv(x)
        if s == 1:
            xp = self.with_halo(x, 1, 1)
        else:  # stride-2, pad 1: output row y reads input rows 2y-1 .. 2y+1 -> only the row above
            xp = self.with_halo(x, 1, 0)
            xp = F.pad(xp, (0, 0, 0, 1))  # the (unused for even heights) bottom pad row
        return F.conv2d(xp, conv.weight, conv.bias, stride=s, padding=(0, conv.padding[1]))

    def group_norm(self, gn: torch.nn.GroupNorm, x):
        n, c, h, w = x.shape
        g = gn.num_groups
        xf = x.float().view(n, g, c // g, h * w)
        stats = torch.stack([xf.sum(dim=(2, 3)), (xf * xf).sum(dim=(2, 3))], dim=-1)
        if self.world > 1:
            dist.all_reduce(stats, group=self.group)
        cnt = float(c // g * h * w * self.world)
        mean = stats[..., 0] / cnt
        var = (stats[..., 1] / cnt - mean * mean).clamp_min(0)
        y = (xf - mean[:, :, None, None]) * torch.rsqrt(var + gn.eps)[:, :, None, None]
        y = y.view(n, c, h, w) * gn.weight.float()[None, :, None, None] + gn.bias.float()[None, :, None, None]
        return y.to(x.dtype)

    def gather_tokens(self, t):
        """[b, s_local, ...] -> [b, world*s_local, ...] in rank order."""
        if self.world == 1:
            return t
        parts = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(parts, t.contiguous(), group=self.group)
        return torch.cat(parts, dim=1)

    def gather_rows(self, x):
        """[n, c, h_local, w] -> full feature map."""
        if self.world == 1:
            return x
        parts = [torch.empty_like(x) for _ in range(self.world)]
        dist.all_gather(parts, x.contiguous(), group=self.group)
        return torch.cat(parts, dim=2)


class PatchParallelUNet:
    """Runs a UNet2DConditionModel with row-sharded activations. Every rank holds all weights."""

    def __init__(self, unet, gang: Gang):
        self.unet, self.ops = unet, PatchOps(gang)

    def split(self, sample):
        h = sample.shape[2]
        w = self.ops.world
        depth = len(self.unet.config["block_out_channels"]) - 1
        if h % (w * 2 ** depth):
            raise ValueError(f"latent height {h} cannot be split over {w} ranks at every UNet level")
        hl = h // w
        return sample[:, :, self.ops.rank * hl:(self.ops.rank + 1) * hl].contiguous()

    # ---- building blocks mirroring models/unet2d.py
    def _resnet(self, m, x, temb):
        o = self.ops
        h = o.conv(m.conv1, F.silu(o.group_norm(m.norm1, x)))
        if m.time_emb_proj is not None and temb is not None:
            h = h + m.time_emb_proj(F.silu(temb))[:, :, None, None]
        h = o.conv(m.conv2, F.silu(o.group_norm(m.norm2, h)))
        if m.conv_shortcut is not None:
            x = m.conv_shortcut(x)
        return (x + h) / m.output_scale_factor

    def _self_attention(self, attn, x):
        b, s, _ = x.shape
        sh = lambda t: t.view(b, -1, attn.heads, attn.dim_head).transpose(1, 2)
        q = sh(attn.to_q(x))
        k = sh(self.ops.gather_tokens(attn.to_k(x)))
        v = sh(self.ops.gather_tokens(attn.to_v(x)))
        o = F.scaled_dot_product_attention(q, k, v, scale=attn.scale)
        return attn.to_out[0](o.transpose(1, 2).reshape(b, s, attn.heads * attn.dim_head))

    def _transformer(self, m, x, context):
        b, c, h, w = x.shape
        res = x
        y = self.ops.group_norm(m.norm, x)
        if m.use_linear_projection:
            y = m.proj_in(y.permute(0, 2, 3, 1).reshape(b, h * w, c))
        else:
            y = m.proj_in(y).permute(0, 2, 3, 1).reshape(b, h * w, -1)
        for blk in m.transformer_blocks:
            if blk.only_cross_attention:
                y = y + blk.attn1(blk.norm1(y), context)
            else:
                y = y + self._self_attention(blk.attn1, blk.norm1(y))
            y = y + blk.attn2(blk.norm2(y), context)  # text K/V are replicated: local
            y = y + blk.ff(blk.norm3(y))
        if m.use_linear_projection:
            y = m.proj_out(y).reshape(b, h, w, c).permute(0, 3, 1, 2)
        else:
            y = m.proj_out(y.reshape(b, h, w, -1).permute(0, 3, 1, 2))
        return y + res

    @torch.no_grad()
    def forward_local(self, sample_local, timestep, context, added_cond=None):
        """sample_local: this rank's rows. Returns this rank's rows of the prediction."""
        u, o = self.unet, self.ops
        t = torch.as_tensor(timestep, device=sample_local.device)
        emb = u.time_embed(t, sample_local.shape[0], sample_local.dtype, added_cond)
        x = o.conv(u.conv_in, sample_local)
        skips = [x]
        for blk in u.down_blocks:
            for i, r in enumerate(blk.resnets):
                x = self._resnet(r, x, emb)
                if blk.attentions is not None:
                    x = self._transformer(blk.attentions[i], x, context)
                skips.append(x)
            if blk.downsamplers is not None:
                x = o.conv(blk.downsamplers[0].conv, x)
                skips.append(x)
        x = self._resnet(u.mid_block.resnets[0], x, emb)
        if u.mid_block.attentions is not None:
            x = self._transformer(u.mid_block.attentions[0], x, context)
        x = self._resnet(u.mid_block.resnets[1], x, emb)
        for blk in u.up_blocks:
            for i, r in enumerate(blk.resnets):
                x = self._resnet(r, torch.cat([x, skips.pop()], dim=1), emb)
                if blk.attentions is not None:
                    x = self._transformer(blk.attentions[i], x, context)
            if blk.upsamplers is not None:
                x = o.conv(blk.upsamplers[0].conv, F.interpolate(x, scale_factor=2.0, mode="nearest"))
        return o.conv(u.conv_out, F.silu(o.group_norm(u.conv_norm_out, x)))

    @torch.no_grad()
    def __call__(self, sample, timestep, context, added_cond=None):
        """Full-tensor convenience wrapper: split, run, gather."""
        return self.ops.gather_rows(self.forward_local(self.split(sample), timestep, context, added_cond))


class PatchParallelVAEDecoder:
    """`AutoencoderKL.decode` with row-sharded activations: [n, 4, h, w] latents -> this rank's strip of the image."""

    def __init__(self, vae, gang: Gang):
        self.vae, self.ops = vae, PatchOps(gang)

    def split(self, z):
        h, w = z.shape[2], self.ops.world
        if h % w:
            raise ValueError(f"latent height {h} cannot be split over {w} ranks")
        hl = h // w
        return z[:, :, self.ops.rank * hl:(self.ops.rank + 1) * hl].contiguous()

    def _resnet(self, m, x):
        o = self.ops
        h = o.conv(m.conv1, F.silu(o.group_norm(m.norm1, x)))
        h = o.conv(m.conv2, F.silu(o.group_norm(m.norm2, h)))
        if m.conv_shortcut is not None:
            x = m.conv_shortcut(x)
        return (x + h) / m.output_scale_factor

    def _attention(self, a, x):
        b, c, h, w = x.shape
        y = self.ops.group_norm(a.group_norm, x).reshape(b, c, h * w).transpose(1, 2)
        q = a.to_q(y)
        k, v = self.ops.gather_tokens(a.to_k(y)), self.ops.gather_tokens(a.to_v(y))
        o = F.scaled_dot_product_attention(q[:, None], k[:, None], v[:, None])[:, 0]
        return x + a.to_out[0](o).transpose(1, 2).reshape(b, c, h, w)

    @torch.no_grad()
    def decode_local(self, z_local):
        d, o = self.vae.decoder, self.ops
        x = o.conv(d.conv_in, self.vae.post_quant_conv(z_local))
        x = self._resnet(d.mid_block.resnets[0], x)
        x = self._attention(d.mid_block.attentions[0], x)
        x = self._resnet(d.mid_block.resnets[1], x)
        for blk in d.up_blocks:
            for r in blk.resnets:
                x = self._resnet(r, x)
            if blk.upsamplers is not None:
                x = o.conv(blk.upsamplers[0].conv, F.interpolate(x, scale_factor=2.0, mode="nearest"))
        return o.conv(d.conv_out, F.silu(o.group_norm(d.conv_norm_out, x)))

    @torch.no_grad()
    def __call__(self, z):
        return self.ops.gather_rows(self.decode_local(self.split(z)))
