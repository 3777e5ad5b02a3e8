"""K-diffusion latent UNet of `stabilityai/sd-x2-latent-upscaler` (the network behind the reference's `upscale` job flag:
swarm/diffusion/upscale.py:6-32 loads it through `StableDiffusionLatentUpscalePipeline.from_pretrained`).

The checkpoint is a diffusers `UNet2DConditionModel` built only from block types the latent-diffusion UNet
(models/unet2d.py) does not have.  This file implements those with the checkpoint's parameter names, so that
`unet/diffusion_pytorch_model.safetensors` loads with strict key matching (a wrong name is a loud load error):

  KDownBlock2D / KUpBlock2D                   resnets only
  KCrossAttnDownBlock2D / KCrossAttnUpBlock2D resnets interleaved with KAttentionBlocks
  ResnetBlockCondNorm2D ("ada_group")         both norms are AdaGroupNorm: an affine-free GroupNorm (group SIZE 32, so the
                                              group count follows the width) whose scale / shift are a Linear of the time
                                              embedding; the 1x1 shortcut has no bias
  KAttentionBlock                             optional self-attention and a cross-attention, each behind an AdaGroupNorm
                                              in NCHW, residual in NCHW; the text is LayerNorm'ed inside the cross-attention
  KDownsample2D / KUpsample2D                 parameter-free: a fixed separable [1 3 3 1]/8 filter, reflect padding,
                                              stride-2 conv / transposed conv (a depthwise conv here; diffusers builds the
                                              same thing as a dense conv with a diagonal weight)
  time_embedding_type = "fourier"             cos / sin of 2 pi * t * W with a stored random W (`time_proj.weight`);
                                              `time_cond_proj_dim`: the pipeline's [noise-level embedding ; CLIP pooled]
                                              vector enters through `time_embedding.cond_proj`
  no mid block, no output norm, 1x1 conv_in / conv_out

The skip wiring is the K-UNet's own: a down block hands its LAST resnet output (before down-sampling) to the mirrored up
block, which concatenates it once at its input; the deepest down block hands over nothing.

Eager PyTorch only: a 20-step post stage off the named hot path.  Its GEMM-shaped work still lands
on cuBLAS / SDPA in fp16 on the GPU; the reference runs the same network with sequential CPU offload."""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

GROUP_SIZE = 32

_DOWN = {"KDownBlock2D": False, "KCrossAttnDownBlock2D": True}
_UP = {"KUpBlock2D": False, "KCrossAttnUpBlock2D": True}


def _act(name):
    if name is None:
        return None
    return {"silu": F.silu, "swish": F.silu, "gelu": F.gelu, "mish": F.mish}[name]


class GaussianFourierProjection(nn.Module):
    def __init__(self, embedding_size, scale=1.0, flip_sin_to_cos=True):
        super().__init__()
        self.weight = nn.Parameter(torch.randn(embedding_size) * scale, requires_grad=False)
        self.flip_sin_to_cos = flip_sin_to_cos

    def forward(self, x):
        p = x[:, None].float() * self.weight[None, :].float() * (2 * math.pi)
        return torch.cat([p.cos(), p.sin()] if self.flip_sin_to_cos else [p.sin(), p.cos()], dim=-1)


class TimestepEmbedding(nn.Module):
    def __init__(self, in_dim, dim, act, post_act=None, cond_proj_dim=None):
        super().__init__()
        self.linear_1 = nn.Linear(in_dim, dim)
        self.cond_proj = nn.Linear(cond_proj_dim, in_dim, bias=False) if cond_proj_dim else None
        self.linear_2 = nn.Linear(dim, dim)
        self.act, self.post_act = _act(act), _act(post_act)

    def forward(self, x, cond=None):
        if self.cond_proj is not None:
            if cond is None:
                raise ValueError("this UNet's time embedding is conditioned (time_cond_proj_dim): timestep_cond is required")
            x = x + self.cond_proj(cond)
        x = self.linear_2(self.act(self.linear_1(x)))
        return x if self.post_act is None else self.post_act(x)


class AdaGroupNorm(nn.Module):
    def __init__(self, emb_dim, channels, groups, eps=1e-5):
        super().__init__()
        self.groups, self.eps = groups, eps
        self.linear = nn.Linear(emb_dim, channels * 2)

    def forward(self, x, emb):
        scale, shift = self.linear(emb)[:, :, None, None].chunk(2, dim=1)
        return F.group_norm(x, self.groups, eps=self.eps) * (1 + scale) + shift


class ResnetBlockCondNorm2D(nn.Module):
    def __init__(self, cin, cout, temb_dim, eps, act, conv2_out=None):
        super().__init__()
        conv2_out = conv2_out or cout
        self.norm1 = AdaGroupNorm(temb_dim, cin, cin // GROUP_SIZE, eps)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.norm2 = AdaGroupNorm(temb_dim, cout, cout // GROUP_SIZE, eps)
        self.conv2 = nn.Conv2d(cout, conv2_out, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(cin, conv2_out, 1, bias=False) if cin != conv2_out else None
        self.act = _act(act)

    def forward(self, x, temb):
        h = self.conv1(self.act(self.norm1(x, temb)))
        h = self.conv2(self.act(self.norm2(h, temb)))
        return (x if self.conv_shortcut is None else self.conv_shortcut(x)) + h


class Attention(nn.Module):
    """diffusers `Attention` as KAttentionBlock configures it: biased projections, optional LayerNorm on the context."""

    def __init__(self, dim, heads, dim_head, cross_dim=None, norm_cross=False, upcast=False):
        super().__init__()
        inner = heads * dim_head
        self.heads, self.upcast = heads, upcast
        self.norm_cross = nn.LayerNorm(cross_dim) if (norm_cross and cross_dim) else None
        self.to_q = nn.Linear(dim, inner)
        self.to_k = nn.Linear(cross_dim or dim, inner)
        self.to_v = nn.Linear(cross_dim or dim, inner)
        self.to_out = nn.ModuleList([nn.Linear(inner, dim), nn.Identity()])

    def forward(self, x, context=None):
        b, n, _ = x.shape
        if context is None:
            context = x
        elif self.norm_cross is not None:
            context = self.norm_cross(context)

        def heads(t):
            return t.view(b, -1, self.heads, t.shape[-1] // self.heads).transpose(1, 2)

        q, k, v = heads(self.to_q(x)), heads(self.to_k(context)), heads(self.to_v(context))
        if self.upcast:
            q, k = q.float(), k.float()
            v = v.float()
        o = F.scaled_dot_product_attention(q, k, v).to(x.dtype)
        return self.to_out[0](o.transpose(1, 2).reshape(b, n, -1))


class KAttentionBlock(nn.Module):
    def __init__(self, dim, heads, dim_head, cross_dim, temb_dim, self_attention, upcast=False):
        super().__init__()
        groups = max(1, dim // GROUP_SIZE)
        self.add_self_attention = self_attention
        if self_attention:
            self.norm1 = AdaGroupNorm(temb_dim, dim, groups)
            self.attn1 = Attention(dim, heads, dim_head)
        self.norm2 = AdaGroupNorm(temb_dim, dim, groups)
        self.attn2 = Attention(dim, heads, dim_head, cross_dim, norm_cross=True, upcast=upcast)

    @staticmethod
    def _tokens(x):
        return x.flatten(2).transpose(1, 2)

    def forward(self, x, context, temb):
        shape = x.shape
        if self.add_self_attention:
            x = x + self.attn1(self._tokens(self.norm1(x, temb))).transpose(1, 2).reshape(shape)
        return x + self.attn2(self._tokens(self.norm2(x, temb)), context).transpose(1, 2).reshape(shape)


def _blur_kernel(gain=1.0):
    k = torch.tensor([1 / 8, 3 / 8, 3 / 8, 1 / 8]) * gain
    return torch.outer(k, k)


class KDownsample2D(nn.Module):
    def __init__(self):
        super().__init__()
        self.register_buffer("kernel", _blur_kernel(), persistent=False)

    def forward(self, x):
        c = x.shape[1]
        x = F.pad(x, (1, 1, 1, 1), "reflect")
        return F.conv2d(x, self.kernel.to(x.dtype)[None, None].expand(c, 1, 4, 4), stride=2, groups=c)


class KUpsample2D(nn.Module):
    def __init__(self):
        super().__init__()
        self.register_buffer("kernel", _blur_kernel(2.0), persistent=False)

    def forward(self, x):
        c = x.shape[1]
        x = F.pad(x, (1, 1, 1, 1), "reflect")
        return F.conv_transpose2d(x, self.kernel.to(x.dtype)[None, None].expand(c, 1, 4, 4), stride=2, padding=3, groups=c)


class KDownBlock(nn.Module):
    """KDownBlock2D / KCrossAttnDownBlock2D.  Returns (output, skip): the skip is the last resnet/attention output before
    down-sampling, or None for the deepest block (which has no down-sampler)."""

    def __init__(self, cin, cout, temb_dim, layers, eps, act, add_down, attn=None):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlockCondNorm2D(cin if i == 0 else cout, cout, temb_dim, eps, act)
                                      for i in range(layers)])
        self.attentions = None
        if attn is not None:
            head_dim, cross_dim, self_attention = attn
            self.attentions = nn.ModuleList([KAttentionBlock(cout, cout // head_dim, head_dim, cross_dim, temb_dim,
                                                             self_attention) for _ in range(layers)])
        self.downsamplers = nn.ModuleList([KDownsample2D()]) if add_down else None

    def forward(self, x, temb, context):
        for i, res in enumerate(self.resnets):
            x = res(x, temb)
            if self.attentions is not None:
                x = self.attentions[i](x, context, temb)
        if self.downsamplers is None:
            return x, (None if self.attentions is not None else x)
        return self.downsamplers[0](x), x


class KUpBlock(nn.Module):
    """KUpBlock2D / KCrossAttnUpBlock2D.  `cin` is the width this block hands on (diffusers' k_out_channels), `cout` the
    width it works at; `layers` counts resnets (diffusers passes layers_per_block + 1 and builds one fewer)."""

    def __init__(self, cin, cout, temb_dim, layers, eps, act, add_up, attn=None, upcast=False):
        super().__init__()
        has_attn = attn is not None
        first = has_attn and cin == cout == temb_dim       # the deepest block: no skip to concatenate, self-attention on
        middle = cin != cout
        k_in = cout if first else 2 * cout
        res, att = [], []
        for i in range(layers):
            last = i == layers - 1
            if has_attn:
                res.append(ResnetBlockCondNorm2D(k_in if i == 0 else cout, cout, temb_dim, eps, act,
                                                 conv2_out=cin if (middle and last) else None))
                head_dim, cross_dim = attn
                width = cin if last else cout
                att.append(KAttentionBlock(width, width // head_dim, head_dim, cross_dim, temb_dim, first, upcast))
            else:
                res.append(ResnetBlockCondNorm2D(k_in if i == 0 else cout, cin if last else cout, temb_dim, eps, act))
        self.resnets = nn.ModuleList(res)
        self.attentions = nn.ModuleList(att) if has_attn else None
        self.upsamplers = nn.ModuleList([KUpsample2D()]) if add_up else None

    def forward(self, x, skip, temb, context):
        if skip is not None:
            x = torch.cat([x, skip], dim=1)
        for i, res in enumerate(self.resnets):
            x = res(x, temb)
            if self.attentions is not None:
                x = self.attentions[i](x, context, temb)
        return x if self.upsamplers is None else self.upsamplers[0](x)


K_TINY_CONFIG = dict(
    in_channels=8, out_channels=5, block_out_channels=(32, 32, 64, 64), layers_per_block=(1, 1, 2, 2),
    down_block_types=("KDownBlock2D", "KCrossAttnDownBlock2D", "KCrossAttnDownBlock2D", "KCrossAttnDownBlock2D"),
    up_block_types=("KCrossAttnUpBlock2D", "KCrossAttnUpBlock2D", "KCrossAttnUpBlock2D", "KUpBlock2D"),
    mid_block_type=None, attention_head_dim=16, cross_attention_dim=32, only_cross_attention=(True, True, True, False),
    act_fn="gelu", norm_num_groups=None, norm_eps=1e-5, resnet_time_scale_shift="scale_shift",
    time_embedding_type="fourier", timestep_post_act="gelu", time_cond_proj_dim=48, conv_in_kernel=1, conv_out_kernel=1,
    flip_sin_to_cos=True, freq_shift=0)


class KUNet(nn.Module):
    """forward(sample [b, in, h, w], timestep, encoder_hidden_states [b, s, cross], timestep_cond [b, cond]) -> [b, out, h, w]."""

    def __init__(self, **cfg):
        super().__init__()
        c = dict(cfg)
        for key, default in (("norm_eps", 1e-5), ("act_fn", "gelu"), ("layers_per_block", 2), ("flip_sin_to_cos", True),
                             ("conv_in_kernel", 1), ("conv_out_kernel", 1), ("only_cross_attention", False)):
            c.setdefault(key, default)
        self.config = c
        bad = [t for t in c["down_block_types"] if t not in _DOWN] + [t for t in c["up_block_types"] if t not in _UP]
        if (bad or c.get("mid_block_type") is not None or c.get("time_embedding_type") != "fourier"
                or c.get("norm_num_groups") is not None):
            raise ValueError(f"not a K-diffusion latent-upscaler UNet configuration: blocks {sorted(set(bad))}, mid "
                             f"{c.get('mid_block_type')}, time embedding {c.get('time_embedding_type')}, "
                             f"norm_num_groups {c.get('norm_num_groups')}")
        boc = list(c["block_out_channels"])
        n = len(boc)
        as_list = lambda v: list(v) if isinstance(v, (list, tuple)) else [v] * n  # noqa: E731
        layers, head_dim, oca = as_list(c["layers_per_block"]), as_list(c["attention_head_dim"]), as_list(c["only_cross_attention"])
        eps, act, cross = c["norm_eps"], c["act_fn"], c["cross_attention_dim"]
        temb_dim = boc[0] * 2
        kin, kout = c["conv_in_kernel"], c["conv_out_kernel"]
        self.conv_in = nn.Conv2d(c["in_channels"], boc[0], kin, padding=(kin - 1) // 2)
        self.time_proj = GaussianFourierProjection(temb_dim // 2, flip_sin_to_cos=c["flip_sin_to_cos"])
        self.time_embedding = TimestepEmbedding(temb_dim, temb_dim, act, c.get("timestep_post_act"), c.get("time_cond_proj_dim"))

        self.down_blocks = nn.ModuleList()
        ch = boc[0]
        for i, typ in enumerate(c["down_block_types"]):
            cin, ch = ch, boc[i]
            attn = (head_dim[i], cross, not oca[i]) if _DOWN[typ] else None
            self.down_blocks.append(KDownBlock(cin, ch, temb_dim, layers[i], eps, act, i < n - 1, attn))
        self.up_blocks = nn.ModuleList()
        rev, rlayers, rhead = boc[::-1], layers[::-1], head_dim[::-1]
        for i, typ in enumerate(c["up_block_types"]):
            attn = (rhead[i], cross) if _UP[typ] else None
            self.up_blocks.append(KUpBlock(rev[min(i + 1, n - 1)], rev[i], temb_dim, rlayers[i], eps, act, i < n - 1, attn,
                                           bool(c.get("upcast_attention", False))))
        self.conv_out = nn.Conv2d(boc[0], c["out_channels"], kout, padding=(kout - 1) // 2)

    def forward(self, sample, timestep, encoder_hidden_states, timestep_cond=None):
        t = torch.as_tensor(timestep, device=sample.device, dtype=torch.float32)
        if t.dim() == 0:
            t = t[None]
        t = t.expand(sample.shape[0])
        emb = self.time_embedding(self.time_proj(t).to(sample.dtype), timestep_cond)
        x = self.conv_in(sample)
        skips = []
        for blk in self.down_blocks:
            x, skip = blk(x, emb, encoder_hidden_states)
            skips.append(skip)
        for blk in self.up_blocks:
            x = blk(x, skips.pop(), emb, encoder_hidden_states)
        return self.conv_out(x)
