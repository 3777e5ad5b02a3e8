"""UNet2DConditionModel -- own implementation of the denoiser every Stable-Diffusion-family job
runs (the reference reaches it through `pipeline(**kwargs)`, swarm/diffusion/diffusion_func.py:96;
architecture summary in SURVEY.md section 2.5, config keys in appendix A.7).

Module / parameter names follow the diffusers checkpoint layout exactly (`conv_in`,
`time_embedding.linear_1`, `down_blocks.{i}.resnets.{j}.norm1`, `...attentions.{j}.
transformer_blocks.{k}.attn1.to_q`, `ff.net.0.proj`, `downsamplers.0.conv`, ...), so
`load_state_dict` consumes `unet/diffusion_pytorch_model.safetensors` unchanged.

This file is the *specification and oracle*: a plain eager-PyTorch forward (NCHW, cuDNN/cuBLAS/
SDPA -- the ops diffusers would issue).  The production path is `engine/unet_sm100.py`, which
walks the same module tree, re-packs the weights and runs hand-written sm_100a kernels.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

SD15_CONFIG = dict(
    in_channels=4, out_channels=4, block_out_channels=(320, 640, 1280, 1280), layers_per_block=2,
    down_block_types=("CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "CrossAttnDownBlock2D", "DownBlock2D"),
    up_block_types=("UpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D", "CrossAttnUpBlock2D"),
    attention_head_dim=8, cross_attention_dim=768, norm_num_groups=32, norm_eps=1e-5,
    use_linear_projection=False, sample_size=64,
)


def timestep_embedding(t: torch.Tensor, dim: int, flip_sin_to_cos=True, freq_shift=0.0, max_period=10000.0):
    half = dim // 2
    exponent = -math.log(max_period) * torch.arange(half, dtype=torch.float32, device=t.device)
    freqs = torch.exp(exponent / (half - freq_shift))
    args = t.float()[:, None] * freqs[None]
    emb = torch.cat([args.sin(), args.cos()], dim=-1)
    if flip_sin_to_cos:
        emb = torch.cat([emb[:, half:], emb[:, :half]], dim=-1)
    return emb


class TimestepEmbedding(nn.Module):
    def __init__(self, in_dim, dim):
        super().__init__()
        self.linear_1 = nn.Linear(in_dim, dim)
        self.linear_2 = nn.Linear(dim, dim)

    def forward(self, x):
        return self.linear_2(F.silu(self.linear_1(x)))


class ResnetBlock2D(nn.Module):
    def __init__(self, cin, cout, temb_dim, groups=32, eps=1e-5, output_scale_factor=1.0):
        super().__init__()
        self.in_channels, self.out_channels = cin, cout
        self.norm1 = nn.GroupNorm(groups, cin, eps=eps)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.time_emb_proj = nn.Linear(temb_dim, cout) if temb_dim else None
        self.norm2 = nn.GroupNorm(groups, cout, eps=eps)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None
        self.output_scale_factor = output_scale_factor

    def forward(self, x, temb=None):
        h = self.conv1(F.silu(self.norm1(x)))
        if self.time_emb_proj is not None and temb is not None:
            h = h + self.time_emb_proj(F.silu(temb))[:, :, None, None]
        h = self.conv2(F.silu(self.norm2(h)))
        if self.conv_shortcut is not None:
            x = self.conv_shortcut(x)
        return (x + h) / self.output_scale_factor


class Attention(nn.Module):
    def __init__(self, query_dim, heads, dim_head, cross_dim=None, bias=False, out_bias=True):
        super().__init__()
        inner = heads * dim_head
        self.heads, self.dim_head = heads, dim_head
        self.scale = dim_head ** -0.5
        cross_dim = cross_dim or query_dim
        self.to_q = nn.Linear(query_dim, inner, bias=bias)
        self.to_k = nn.Linear(cross_dim, inner, bias=bias)
        self.to_v = nn.Linear(cross_dim, inner, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim, bias=out_bias), nn.Identity()])
        # optional LoRA deltas (reference: `pipeline.unet.load_attn_procs(lora)`,
        # swarm/diffusion/diffusion_func.py:58-68); merged by utils/lora.py
        self.lora_scale = 1.0

    def forward(self, x, context=None):
        b, s, _ = x.shape
        ctx = x if context is None else context
        q = self.to_q(x).view(b, s, self.heads, self.dim_head).transpose(1, 2)
        k = self.to_k(ctx).view(b, -1, self.heads, self.dim_head).transpose(1, 2)
        v = self.to_v(ctx).view(b, -1, self.heads, self.dim_head).transpose(1, 2)
        o = F.scaled_dot_product_attention(q, k, v, scale=self.scale)
        o = o.transpose(1, 2).reshape(b, s, self.heads * self.dim_head)
        return self.to_out[0](o)


class GEGLU(nn.Module):
    def __init__(self, dim, inner):
        super().__init__()
        self.proj = nn.Linear(dim, inner * 2)

    def forward(self, x):
        v, g = self.proj(x).chunk(2, dim=-1)
        return v * F.gelu(g)


class FeedForward(nn.Module):
    def __init__(self, dim, mult=4):
        super().__init__()
        self.net = nn.ModuleList([GEGLU(dim, dim * mult), nn.Identity(), nn.Linear(dim * mult, dim)])

    def forward(self, x):
        return self.net[2](self.net[0](x))


class BasicTransformerBlock(nn.Module):
    def __init__(self, dim, heads, dim_head, cross_dim, only_cross_attention=False):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = Attention(dim, heads, dim_head, cross_dim if only_cross_attention else None)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = Attention(dim, heads, dim_head, cross_dim)
        self.norm3 = nn.LayerNorm(dim)
        self.ff = FeedForward(dim)
        self.only_cross_attention = only_cross_attention

    def forward(self, x, context):
        x = x + self.attn1(self.norm1(x), context if self.only_cross_attention else None)
        x = x + self.attn2(self.norm2(x), context)
        return x + self.ff(self.norm3(x))


class Transformer2DModel(nn.Module):
    def __init__(self, channels, heads, dim_head, depth, cross_dim, groups=32, use_linear_projection=False,
                 only_cross_attention=False):
        super().__init__()
        inner = heads * dim_head
        self.use_linear_projection = use_linear_projection
        self.norm = nn.GroupNorm(groups, channels, eps=1e-6)
        if use_linear_projection:
            self.proj_in = nn.Linear(channels, inner)
            self.proj_out = nn.Linear(inner, channels)
        else:
            self.proj_in = nn.Conv2d(channels, inner, 1)
            self.proj_out = nn.Conv2d(inner, channels, 1)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(inner, heads, dim_head, cross_dim, only_cross_attention) for _ in range(depth)])

    def forward(self, x, context):
        b, c, h, w = x.shape
        res = x
        y = self.norm(x)
        if self.use_linear_projection:
            y = self.proj_in(y.permute(0, 2, 3, 1).reshape(b, h * w, c))
        else:
            y = self.proj_in(y).permute(0, 2, 3, 1).reshape(b, h * w, -1)
        for blk in self.transformer_blocks:
            y = blk(y, context)
        if self.use_linear_projection:
            y = self.proj_out(y).reshape(b, h, w, c).permute(0, 3, 1, 2)
        else:
            y = self.proj_out(y.reshape(b, h, w, -1).permute(0, 3, 1, 2))
        return y + res


class Downsample2D(nn.Module):
    def __init__(self, channels, padding=1):
        super().__init__()
        self.padding = padding
        self.conv = nn.Conv2d(channels, channels, 3, stride=2, padding=padding)

    def forward(self, x):
        if self.padding == 0:  # VAE encoder: asymmetric (0,1,0,1) pad
            x = F.pad(x, (0, 1, 0, 1))
        return self.conv(x)


class Upsample2D(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.conv = nn.Conv2d(channels, channels, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


def _as_list(v, n):
    return list(v) if isinstance(v, (list, tuple)) else [v] * n


class DownBlock(nn.Module):
    """DownBlock2D / CrossAttnDownBlock2D."""

    def __init__(self, cin, cout, temb_dim, layers, groups, eps, has_attn, heads, dim_head, depth, cross_dim,
                 add_downsample, use_linear_projection, only_cross_attention=False):
        super().__init__()
        self.resnets = nn.ModuleList(
            [ResnetBlock2D(cin if i == 0 else cout, cout, temb_dim, groups, eps) for i in range(layers)])
        self.attentions = nn.ModuleList(
            [Transformer2DModel(cout, heads, dim_head, depth, cross_dim, groups, use_linear_projection, only_cross_attention)
             for _ in range(layers)]) if has_attn else None
        self.downsamplers = nn.ModuleList([Downsample2D(cout)]) if add_downsample else None

    def forward(self, x, temb, context):
        outs = []
        for i, res in enumerate(self.resnets):
            x = res(x, temb)
            if self.attentions is not None:
                x = self.attentions[i](x, context)
            outs.append(x)
        if self.downsamplers is not None:
            x = self.downsamplers[0](x)
            outs.append(x)
        return x, outs


class MidBlock(nn.Module):
    def __init__(self, ch, temb_dim, groups, eps, heads, dim_head, depth, cross_dim, use_linear_projection,
                 has_attn=True, only_cross_attention=False):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(ch, ch, temb_dim, groups, eps) for _ in range(2)])
        self.attentions = nn.ModuleList(
            [Transformer2DModel(ch, heads, dim_head, depth, cross_dim, groups, use_linear_projection,
                                only_cross_attention)]) if has_attn else None

    def forward(self, x, temb, context):
        x = self.resnets[0](x, temb)
        if self.attentions is not None:
            x = self.attentions[0](x, context)
        return self.resnets[1](x, temb)


class UpBlock(nn.Module):
    """UpBlock2D / CrossAttnUpBlock2D."""

    def __init__(self, cin, cout, prev_out, temb_dim, layers, groups, eps, has_attn, heads, dim_head, depth,
                 cross_dim, add_upsample, use_linear_projection, only_cross_attention=False):
        super().__init__()
        resnets = []
        for i in range(layers):
            skip = cin if i == layers - 1 else cout
            rin = prev_out if i == 0 else cout
            resnets.append(ResnetBlock2D(rin + skip, cout, temb_dim, groups, eps))
        self.resnets = nn.ModuleList(resnets)
        self.attentions = nn.ModuleList(
            [Transformer2DModel(cout, heads, dim_head, depth, cross_dim, groups, use_linear_projection, only_cross_attention)
             for _ in range(layers)]) if has_attn else None
        self.upsamplers = nn.ModuleList([Upsample2D(cout)]) if add_upsample else None

    def forward(self, x, skips, temb, context):
        for i, res in enumerate(self.resnets):
            x = res(torch.cat([x, skips.pop()], dim=1), temb)
            if self.attentions is not None:
                x = self.attentions[i](x, context)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x)
        return x


class UNet2DConditionModel(nn.Module):
    def __init__(self, **cfg):
        super().__init__()
        c = dict(SD15_CONFIG)
        c.update(cfg)
        self.config = c
        boc = list(c["block_out_channels"])
        n = len(boc)
        groups, eps = c["norm_num_groups"], c["norm_eps"]
        temb_dim = boc[0] * 4
        layers = _as_list(c["layers_per_block"], n)
        cross = c["cross_attention_dim"]
        if isinstance(cross, (list, tuple)):  # (per-block lists appear in some configs; this UNet takes one width)
            if len({x for x in cross}) > 1:
                raise ValueError(f"per-block cross_attention_dim {cross} is not supported")
            cross = cross[0]
        depth = _as_list(c.get("transformer_layers_per_block", 1), n)
        ulp = c.get("use_linear_projection", False)
        # `attention_head_dim` is the number of heads for SD1.x/2.x checkpoints (appendix A.7);
        # `num_attention_heads` wins when present.
        heads_cfg = c.get("num_attention_heads") or c["attention_head_dim"]
        heads = _as_list(heads_cfg, n)
        # per-level flag: "self"-attention of these blocks attends to the text instead (SD x4 upscaler)
        oca = [bool(v) for v in _as_list(c.get("only_cross_attention", False), n)]

        self.conv_in = nn.Conv2d(c["in_channels"], boc[0], 3, padding=1)
        self.time_embedding = TimestepEmbedding(boc[0], temb_dim)
        self.add_embedding = None
        if c.get("addition_embed_type") == "text_time":
            self.add_embedding = TimestepEmbedding(c["projection_class_embeddings_input_dim"], temb_dim)
        self.class_embedding = None
        if c.get("num_class_embeds"):
            self.class_embedding = nn.Embedding(c["num_class_embeds"], temb_dim)
        elif c.get("class_embed_type") == "timestep":
            self.class_embedding = TimestepEmbedding(boc[0], temb_dim)
        elif c.get("class_embed_type") == "simple_projection":  # AudioLDM: a Linear of the CLAP text embedding
            self.class_embedding = nn.Linear(c["projection_class_embeddings_input_dim"], temb_dim)
        elif c.get("class_embed_type") is not None:
            raise ValueError(f"class_embed_type {c['class_embed_type']} is not supported")
        # AudioLDM concatenates the class embedding to the time embedding instead of adding it: every resnet then
        # projects a vector twice as wide
        self.class_embeddings_concat = bool(c.get("class_embeddings_concat", False))
        temb_dim = temb_dim * 2 if self.class_embeddings_concat else temb_dim

        self.down_blocks = nn.ModuleList()
        ch = boc[0]
        for i, typ in enumerate(c["down_block_types"]):
            cin, ch = ch, boc[i]
            self.down_blocks.append(DownBlock(cin, ch, temb_dim, layers[i], groups, eps, "CrossAttn" in typ,
                                              heads[i], ch // heads[i], depth[i], cross, i < n - 1, ulp, oca[i]))
        self.mid_block = MidBlock(boc[-1], temb_dim, groups, eps, heads[-1], boc[-1] // heads[-1], depth[-1], cross,
                                  ulp, has_attn=c.get("mid_block_type", "UNetMidBlock2DCrossAttn") is not None,
                                  only_cross_attention=c.get("mid_block_only_cross_attention", oca[-1]) or False)
        self.up_blocks = nn.ModuleList()
        rev = boc[::-1]
        rheads, rdepth, rlayers, roca = heads[::-1], depth[::-1], layers[::-1], oca[::-1]
        ch = rev[0]
        for i, typ in enumerate(c["up_block_types"]):
            prev, ch = ch, rev[i]
            cin = rev[min(i + 1, n - 1)]
            self.up_blocks.append(UpBlock(cin, ch, prev, temb_dim, rlayers[i] + 1, groups, eps, "CrossAttn" in typ,
                                          rheads[i], ch // rheads[i], rdepth[i], cross, i < n - 1, ulp, roca[i]))
        self.conv_norm_out = nn.GroupNorm(groups, boc[0], eps=eps)
        self.conv_out = nn.Conv2d(boc[0], c["out_channels"], 3, padding=1)

    def time_embed(self, t, batch, dtype, added_cond=None, class_labels=None):
        if t.dim() == 0:
            t = t[None].expand(batch)
        c = self.config
        temb = timestep_embedding(t, c["block_out_channels"][0], c.get("flip_sin_to_cos", True),
                                  c.get("freq_shift", 0)).to(dtype)
        emb = self.time_embedding(temb)
        if self.add_embedding is not None and added_cond is not None and "time_ids" in added_cond:
            tids = added_cond["time_ids"]
            tid_emb = timestep_embedding(tids.flatten(), c["addition_time_embed_dim"], c.get("flip_sin_to_cos", True),
                                         c.get("freq_shift", 0)).reshape(tids.shape[0], -1)
            add = torch.cat([added_cond["text_embeds"].float(), tid_emb], dim=-1).to(dtype)
            emb = emb + self.add_embedding(add)
        if class_labels is None and added_cond is not None:
            class_labels = added_cond.get("class_labels")
        if self.class_embedding is not None and class_labels is not None:
            if isinstance(self.class_embedding, TimestepEmbedding):
                class_labels = timestep_embedding(class_labels, c["block_out_channels"][0]).to(dtype)
            class_emb = self.class_embedding(class_labels.to(dtype) if class_labels.is_floating_point() else class_labels)
            emb = torch.cat([emb, class_emb], dim=-1) if self.class_embeddings_concat else emb + class_emb
        elif self.class_embeddings_concat:
            raise ValueError("this UNet concatenates a class embedding to the time embedding: class_labels are required")
        return emb

    def forward(self, sample, timestep, encoder_hidden_states, down_residuals: Optional[Sequence] = None,
                mid_residual=None, added_cond=None, class_labels=None):
        t = torch.as_tensor(timestep, device=sample.device)
        emb = self.time_embed(t, sample.shape[0], sample.dtype, added_cond, class_labels)
        x = self.conv_in(sample)
        skips = [x]
        for blk in self.down_blocks:
            x, outs = blk(x, emb, encoder_hidden_states)
            skips.extend(outs)
        if down_residuals is not None:  # ControlNet residuals
            skips = [s + r for s, r in zip(skips, down_residuals)]
        x = self.mid_block(x, emb, encoder_hidden_states)
        if mid_residual is not None:
            x = x + mid_residual
        for blk in self.up_blocks:
            x = blk(x, skips, emb, encoder_hidden_states)
        return self.conv_out(F.silu(self.conv_norm_out(x)))
