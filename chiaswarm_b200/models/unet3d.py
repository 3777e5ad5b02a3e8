"""UNet3DConditionModel (zeroscope / modelscope text-to-video; reached by the reference through
`pipeline(**kwargs)` at swarm/video/tx2vid.py:55; structure in SURVEY.md section 2.5 and row K5).

SD-2-style 2-D skeleton applied per frame, plus after every resnet a `TemporalConvLayer`
(4 x [GroupNorm -> SiLU -> Conv3d (3,1,1)], residual) and after every spatial transformer a
`TransformerTemporalModel` (attention over the frame axis), plus `transformer_in` after conv_in.
Parameter names follow the diffusers checkpoint (`temp_convs.{i}.conv{1..4}.{0|2|3}`,
`temp_attentions.{i}.transformer_blocks.0.attn1.to_q`, `transformer_in...`).

Tensors are kept as (B*F, C, H, W); temporal layers reshape internally.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .unet2d import (Attention, Downsample2D, FeedForward, ResnetBlock2D, TimestepEmbedding, Transformer2DModel,
                     Upsample2D, _as_list, timestep_embedding)

ZEROSCOPE_CONFIG = dict(
    in_channels=4, out_channels=4, block_out_channels=(320, 640, 1280, 1280), layers_per_block=2,
    down_block_types=("CrossAttnDownBlock3D", "CrossAttnDownBlock3D", "CrossAttnDownBlock3D", "DownBlock3D"),
    up_block_types=("UpBlock3D", "CrossAttnUpBlock3D", "CrossAttnUpBlock3D", "CrossAttnUpBlock3D"),
    attention_head_dim=64, cross_attention_dim=1024, norm_num_groups=32, norm_eps=1e-5, sample_size=32)


class TemporalConvLayer(nn.Module):
    def __init__(self, dim, groups=32):
        super().__init__()

        def conv():
            return nn.Conv3d(dim, dim, (3, 1, 1), padding=(1, 0, 0))

        self.conv1 = nn.Sequential(nn.GroupNorm(groups, dim), nn.SiLU(), conv())
        self.conv2 = nn.Sequential(nn.GroupNorm(groups, dim), nn.SiLU(), nn.Identity(), conv())
        self.conv3 = nn.Sequential(nn.GroupNorm(groups, dim), nn.SiLU(), nn.Identity(), conv())
        self.conv4 = nn.Sequential(nn.GroupNorm(groups, dim), nn.SiLU(), nn.Identity(), conv())

    def forward(self, x, num_frames):
        bf, c, h, w = x.shape
        y = x.reshape(bf // num_frames, num_frames, c, h, w).permute(0, 2, 1, 3, 4)
        ident = y
        y = self.conv4(self.conv3(self.conv2(self.conv1(y))))
        y = ident + y
        return y.permute(0, 2, 1, 3, 4).reshape(bf, c, h, w)


class TemporalTransformerBlock(nn.Module):
    """BasicTransformerBlock with double self-attention (no text conditioning on the frame axis)."""

    def __init__(self, dim, heads, dim_head):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = Attention(dim, heads, dim_head)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = Attention(dim, heads, dim_head)
        self.norm3 = nn.LayerNorm(dim)
        self.ff = FeedForward(dim)

    def forward(self, x):
        x = x + self.attn1(self.norm1(x))
        x = x + self.attn2(self.norm2(x))
        return x + self.ff(self.norm3(x))


class TransformerTemporalModel(nn.Module):
    def __init__(self, channels, heads, dim_head, groups=32):
        super().__init__()
        inner = heads * dim_head
        self.norm = nn.GroupNorm(groups, channels, eps=1e-6)
        self.proj_in = nn.Linear(channels, inner)
        self.transformer_blocks = nn.ModuleList([TemporalTransformerBlock(inner, heads, dim_head)])
        self.proj_out = nn.Linear(inner, channels)

    def forward(self, x, num_frames):
        bf, c, h, w = x.shape
        b = bf // num_frames
        res = x
        y = x.reshape(b, num_frames, c, h, w).permute(0, 2, 1, 3, 4)  # b c f h w (GroupNorm over c, f*h*w)
        y = self.norm(y)
        y = y.permute(0, 3, 4, 2, 1).reshape(b * h * w, num_frames, c)
        y = self.proj_in(y)
        for blk in self.transformer_blocks:
            y = blk(y)
        y = self.proj_out(y)
        y = y.reshape(b, h, w, num_frames, c).permute(0, 3, 4, 1, 2).reshape(bf, c, h, w)
        return y + res


class DownBlock3D(nn.Module):
    def __init__(self, cin, cout, temb_dim, layers, groups, eps, has_attn, heads, dim_head, cross_dim, add_down):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(cin if i == 0 else cout, cout, temb_dim, groups, eps) for i in range(layers)])
        self.temp_convs = nn.ModuleList([TemporalConvLayer(cout, groups) for _ in range(layers)])
        self.attentions = self.temp_attentions = None
        if has_attn:
            self.attentions = nn.ModuleList([Transformer2DModel(cout, heads, dim_head, 1, cross_dim, groups, True) for _ in range(layers)])
            self.temp_attentions = nn.ModuleList([TransformerTemporalModel(cout, heads, dim_head, groups) for _ in range(layers)])
        self.downsamplers = nn.ModuleList([Downsample2D(cout)]) if add_down else None

    def forward(self, x, temb, context, nf):
        outs = []
        for i, res in enumerate(self.resnets):
            x = self.temp_convs[i](res(x, temb), nf)
            if self.attentions is not None:
                x = self.temp_attentions[i](self.attentions[i](x, context), nf)
            outs.append(x)
        if self.downsamplers is not None:
            x = self.downsamplers[0](x)
            outs.append(x)
        return x, outs


class MidBlock3D(nn.Module):
    def __init__(self, ch, temb_dim, groups, eps, heads, dim_head, cross_dim):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(ch, ch, temb_dim, groups, eps) for _ in range(2)])
        self.temp_convs = nn.ModuleList([TemporalConvLayer(ch, groups) for _ in range(2)])
        self.attentions = nn.ModuleList([Transformer2DModel(ch, heads, dim_head, 1, cross_dim, groups, True)])
        self.temp_attentions = nn.ModuleList([TransformerTemporalModel(ch, heads, dim_head, groups)])

    def forward(self, x, temb, context, nf):
        x = self.temp_convs[0](self.resnets[0](x, temb), nf)
        x = self.temp_attentions[0](self.attentions[0](x, context), nf)
        return self.temp_convs[1](self.resnets[1](x, temb), nf)


class UpBlock3D(nn.Module):
    def __init__(self, cin, cout, prev_out, temb_dim, layers, groups, eps, has_attn, heads, dim_head, cross_dim, add_up):
        super().__init__()
        resnets = []
        for i in range(layers):
            skip = cin if i == layers - 1 else cout
            rin = prev_out if i == 0 else cout
            resnets.append(ResnetBlock2D(rin + skip, cout, temb_dim, groups, eps))
        self.resnets = nn.ModuleList(resnets)
        self.temp_convs = nn.ModuleList([TemporalConvLayer(cout, groups) for _ in range(layers)])
        self.attentions = self.temp_attentions = None
        if has_attn:
            self.attentions = nn.ModuleList([Transformer2DModel(cout, heads, dim_head, 1, cross_dim, groups, True) for _ in range(layers)])
            self.temp_attentions = nn.ModuleList([TransformerTemporalModel(cout, heads, dim_head, groups) for _ in range(layers)])
        self.upsamplers = nn.ModuleList([Upsample2D(cout)]) if add_up else None

    def forward(self, x, skips, temb, context, nf):
        for i, res in enumerate(self.resnets):
            x = self.temp_convs[i](res(torch.cat([x, skips.pop()], dim=1), temb), nf)
            if self.attentions is not None:
                x = self.temp_attentions[i](self.attentions[i](x, context), nf)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x)
        return x


class UNet3DConditionModel(nn.Module):
    def __init__(self, **cfg):
        super().__init__()
        c = dict(ZEROSCOPE_CONFIG)
        c.update(cfg)
        self.config = c
        boc = list(c["block_out_channels"])
        n = len(boc)
        groups, eps = c["norm_num_groups"], c["norm_eps"]
        temb_dim = boc[0] * 4
        layers = _as_list(c["layers_per_block"], n)
        cross = c["cross_attention_dim"]
        hd = _as_list(c["attention_head_dim"], n)  # true head dim for the 3-D UNet

        self.conv_in = nn.Conv2d(c["in_channels"], boc[0], 3, padding=1)
        self.time_embedding = TimestepEmbedding(boc[0], temb_dim)
        self.transformer_in = TransformerTemporalModel(boc[0], 8, hd[0], groups)
        self.down_blocks = nn.ModuleList()
        ch = boc[0]
        for i, typ in enumerate(c["down_block_types"]):
            cin, ch = ch, boc[i]
            self.down_blocks.append(DownBlock3D(cin, ch, temb_dim, layers[i], groups, eps, "CrossAttn" in typ,
                                                max(1, ch // hd[i]), hd[i], cross, i < n - 1))
        self.mid_block = MidBlock3D(boc[-1], temb_dim, groups, eps, max(1, boc[-1] // hd[-1]), hd[-1], cross)
        self.up_blocks = nn.ModuleList()
        rev, rhd, rlayers = boc[::-1], hd[::-1], layers[::-1]
        ch = rev[0]
        for i, typ in enumerate(c["up_block_types"]):
            prev, ch = ch, rev[i]
            cin = rev[min(i + 1, n - 1)]
            self.up_blocks.append(UpBlock3D(cin, ch, prev, temb_dim, rlayers[i] + 1, groups, eps, "CrossAttn" in typ,
                                            max(1, ch // rhd[i]), rhd[i], cross, i < n - 1))
        self.conv_norm_out = nn.GroupNorm(groups, boc[0], eps=eps)
        self.conv_out = nn.Conv2d(boc[0], c["out_channels"], 3, padding=1)

    def forward(self, sample, timestep, encoder_hidden_states):
        """sample: (B, C, F, H, W); encoder_hidden_states: (B, S, D). Returns (B, C, F, H, W)."""
        b, c, nf, h, w = sample.shape
        t = torch.as_tensor(timestep, device=sample.device)
        if t.dim() == 0:
            t = t[None].expand(b)
        emb = self.time_embedding(timestep_embedding(t, self.config["block_out_channels"][0]).to(sample.dtype))
        emb = emb.repeat_interleave(nf, dim=0)
        ctx = encoder_hidden_states.repeat_interleave(nf, dim=0)
        x = sample.permute(0, 2, 1, 3, 4).reshape(b * nf, c, h, w)
        x = self.transformer_in(self.conv_in(x), nf)
        skips = [x]
        for blk in self.down_blocks:
            x, outs = blk(x, emb, ctx, nf)
            skips.extend(outs)
        x = self.mid_block(x, emb, ctx, nf)
        for blk in self.up_blocks:
            x = blk(x, skips, emb, ctx, nf)
        x = self.conv_out(F.silu(self.conv_norm_out(x)))
        return x.reshape(b, nf, -1, h, w).permute(0, 2, 1, 3, 4)
