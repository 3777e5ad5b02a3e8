"""DeepFloyd-IF pixel-space UNet (stages I and II of the cascade the reference runs in
swarm/diffusion/diffusion_func_if.py:16-55 through `DiffusionPipeline.from_pretrained`).

The checkpoints are diffusers `UNet2DConditionModel`s assembled from block types the latent-diffusion UNet
(models/unet2d.py) does not have; this file implements exactly those, with the checkpoint's parameter names, so that
`unet/diffusion_pytorch_model.safetensors` loads with strict key matching (a name this file gets wrong is a loud load
error, never silently random weights):

  ResnetDownsampleBlock2D / ResnetUpsampleBlock2D     resnets + a resampling ResnetBlock2D (avg-pool / nearest 2x)
  SimpleCrossAttn{Down,Up}Block2D, UNetMidBlock2DSimpleCrossAttn
                                                      resnets interleaved with added-KV attention: spatial tokens attend
                                                      to [projected text tokens ; spatial tokens] (add_k_proj / add_v_proj),
                                                      GroupNorm inside the attention, optional GroupNorm on the text
  ResnetBlock2D, time_embedding_norm = "scale_shift"  the time embedding modulates norm2's output: h * (1 + scale) + shift
  addition_embed_type = "text"                        attention-pooled T5 states added to the time embedding
  encoder_hid_proj                                    Linear from the T5 width (4096) to cross_attention_dim
  class_embed_type = "timestep" (stage II)            the noise level of the low-resolution input

Eager PyTorch only: the 64 / 256 px pixel-space stages are off the named hot path; stage III, the x4
latent upscaler, runs on the engine (diffusion/pipelines.py mode sd_upscale)."""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .unet2d import timestep_embedding


def _act(name: str):
    return {"silu": F.silu, "swish": F.silu, "gelu": F.gelu, "mish": F.mish}[name]


class TimestepEmbedding(nn.Module):
    def __init__(self, in_dim, dim, act="silu"):
        super().__init__()
        self.linear_1 = nn.Linear(in_dim, dim)
        self.linear_2 = nn.Linear(dim, dim)
        self.act = _act(act)

    def forward(self, x):
        return self.linear_2(self.act(self.linear_1(x)))


class AttentionPooling(nn.Module):
    def __init__(self, num_heads, embed_dim):
        super().__init__()
        self.positional_embedding = nn.Parameter(torch.randn(1, embed_dim) / embed_dim ** 0.5)
        self.k_proj = nn.Linear(embed_dim, embed_dim)
        self.q_proj = nn.Linear(embed_dim, embed_dim)
        self.v_proj = nn.Linear(embed_dim, embed_dim)
        self.num_heads = num_heads
        self.dim_per_head = embed_dim // num_heads

    def forward(self, x):
        bs, length, width = x.shape

        def shape(t):  # [bs, n, width] -> [bs * heads, dim_per_head, n]
            return t.view(bs, -1, self.num_heads, self.dim_per_head).transpose(1, 2).reshape(
                bs * self.num_heads, -1, self.dim_per_head).transpose(1, 2)

        class_token = x.mean(dim=1, keepdim=True) + self.positional_embedding.to(x.dtype)
        x = torch.cat([class_token, x], dim=1)
        q, k, v = shape(self.q_proj(class_token)), shape(self.k_proj(x)), shape(self.v_proj(x))
        scale = 1 / math.sqrt(math.sqrt(self.dim_per_head))
        weight = torch.softmax(torch.einsum("bct,bcs->bts", q * scale, k * scale).float(), dim=-1).type(x.dtype)
        a = torch.einsum("bts,bcs->bct", weight, v)
        return a.reshape(bs, -1, 1).transpose(1, 2)[:, 0, :]


class TextTimeEmbedding(nn.Module):
    def __init__(self, encoder_dim, time_embed_dim, num_heads=64):
        super().__init__()
        self.norm1 = nn.LayerNorm(encoder_dim)
        self.pool = AttentionPooling(num_heads, encoder_dim)
        self.proj = nn.Linear(encoder_dim, time_embed_dim)
        self.norm2 = nn.LayerNorm(time_embed_dim)

    def forward(self, hidden_states):
        return self.norm2(self.proj(self.pool(self.norm1(hidden_states))))


class ResnetBlock2D(nn.Module):
    def __init__(self, cin, cout, temb_dim, groups, eps, act, time_norm="default", output_scale_factor=1.0,
                 skip_time_act=False, up=False, down=False):
        super().__init__()
        self.norm1 = nn.GroupNorm(groups, cin, eps=eps)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.time_norm = time_norm
        self.time_emb_proj = nn.Linear(temb_dim, 2 * cout if time_norm == "scale_shift" else cout) if temb_dim else None
        self.norm2 = nn.GroupNorm(groups, cout, eps=eps)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None
        self.act, self.skip_time_act = _act(act), skip_time_act
        self.output_scale_factor, self.up, self.down = output_scale_factor, up, down

    def forward(self, x, temb):
        h = self.act(self.norm1(x))
        if self.up:  # (parameter-free resampling of both branches, as diffusers' use_conv=False samplers)
            x, h = F.interpolate(x, scale_factor=2.0, mode="nearest"), F.interpolate(h, scale_factor=2.0, mode="nearest")
        elif self.down:
            x, h = F.avg_pool2d(x, 2, 2), F.avg_pool2d(h, 2, 2)
        h = self.conv1(h)
        t = None
        if self.time_emb_proj is not None and temb is not None:
            t = self.time_emb_proj(temb if self.skip_time_act else self.act(temb))[:, :, None, None]
        if t is not None and self.time_norm == "default":
            h = h + t
        h = self.norm2(h)
        if t is not None and self.time_norm == "scale_shift":
            scale, shift = t.chunk(2, dim=1)
            h = h * (1 + scale) + shift
        h = self.conv2(self.act(h))
        if self.conv_shortcut is not None:
            x = self.conv_shortcut(x)
        return (x + h) / self.output_scale_factor


class AddedKVAttention(nn.Module):
    """diffusers `Attention` as configured by the SimpleCrossAttn blocks, evaluated as AttnAddedKVProcessor does."""

    def __init__(self, query_dim, heads, dim_head, added_kv_dim, groups, only_cross_attention=False,
                 cross_attention_norm=None, cross_norm_groups=32):
        super().__init__()
        inner = heads * dim_head
        self.heads, self.dim_head, self.only_cross_attention = heads, dim_head, only_cross_attention
        self.group_norm = nn.GroupNorm(groups, query_dim, eps=1e-5)
        self.norm_cross = None
        if cross_attention_norm == "group_norm":
            self.norm_cross = nn.GroupNorm(cross_norm_groups, added_kv_dim, eps=1e-5)
        elif cross_attention_norm == "layer_norm":
            self.norm_cross = nn.LayerNorm(added_kv_dim)
        self.to_q = nn.Linear(query_dim, inner)
        if not only_cross_attention:
            self.to_k = nn.Linear(query_dim, inner)
            self.to_v = nn.Linear(query_dim, inner)
        self.add_k_proj = nn.Linear(added_kv_dim, inner)
        self.add_v_proj = nn.Linear(added_kv_dim, inner)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim), nn.Identity()])

    def forward(self, x, context):
        b, c, hh, ww = x.shape
        tokens = self.group_norm(x.view(b, c, hh * ww)).transpose(1, 2)
        if self.norm_cross is not None:
            context = (self.norm_cross(context.transpose(1, 2)).transpose(1, 2)
                       if isinstance(self.norm_cross, nn.GroupNorm) else self.norm_cross(context))

        def heads(t):
            return t.view(b, -1, self.heads, self.dim_head).transpose(1, 2)

        q = heads(self.to_q(tokens))
        k, v = heads(self.add_k_proj(context)), heads(self.add_v_proj(context))
        if not self.only_cross_attention:
            k = torch.cat([k, heads(self.to_k(tokens))], dim=2)
            v = torch.cat([v, heads(self.to_v(tokens))], dim=2)
        o = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(b, hh * ww, c)
        return self.to_out[0](o).transpose(1, 2).reshape(b, c, hh, ww) + x


class _Common:
    """Constructor arguments shared by every block of one UNet."""

    def __init__(self, cfg, temb_dim):
        self.temb_dim = temb_dim
        self.groups, self.eps = cfg["norm_num_groups"], cfg["norm_eps"]
        self.act = cfg["act_fn"]
        self.time_norm = cfg.get("resnet_time_scale_shift", "default")
        self.skip_time_act = cfg.get("resnet_skip_time_act", False)
        self.out_scale = cfg.get("resnet_out_scale_factor", 1.0)
        self.cross_dim = cfg["cross_attention_dim"]
        self.cross_norm = cfg.get("cross_attention_norm")

    def resnet(self, cin, cout, scale=None, **kw):
        return ResnetBlock2D(cin, cout, self.temb_dim, self.groups, self.eps, self.act, self.time_norm,
                             self.out_scale if scale is None else scale, self.skip_time_act, **kw)

    def attention(self, ch, head_dim, only_cross):
        return AddedKVAttention(ch, ch // head_dim, head_dim, self.cross_dim, self.groups, only_cross, self.cross_norm)


class DownBlock(nn.Module):
    """ResnetDownsampleBlock2D / SimpleCrossAttnDownBlock2D."""

    def __init__(self, k: _Common, cin, cout, layers, has_attn, head_dim, only_cross, add_down):
        super().__init__()
        self.resnets = nn.ModuleList([k.resnet(cin if i == 0 else cout, cout) for i in range(layers)])
        self.attentions = nn.ModuleList([k.attention(cout, head_dim, only_cross) for _ in range(layers)]) if has_attn else None
        self.downsamplers = nn.ModuleList([k.resnet(cout, cout, down=True)]) if add_down else None

    def forward(self, x, temb, context):
        outs = []
        for i, res in enumerate(self.resnets):
            x = res(x, temb)
            if self.attentions is not None:
                x = self.attentions[i](x, context)
            outs.append(x)
        if self.downsamplers is not None:
            x = self.downsamplers[0](x, temb)
            outs.append(x)
        return x, outs


class MidBlock(nn.Module):
    """UNetMidBlock2DSimpleCrossAttn."""

    def __init__(self, k: _Common, ch, head_dim, only_cross, scale):
        super().__init__()
        self.resnets = nn.ModuleList([k.resnet(ch, ch, scale), k.resnet(ch, ch, scale)])
        self.attentions = nn.ModuleList([k.attention(ch, head_dim, only_cross)])

    def forward(self, x, temb, context):
        x = self.resnets[0](x, temb)
        return self.resnets[1](self.attentions[0](x, context), temb)


class UpBlock(nn.Module):
    """ResnetUpsampleBlock2D / SimpleCrossAttnUpBlock2D."""

    def __init__(self, k: _Common, cin, cout, prev_out, layers, has_attn, head_dim, only_cross, add_up):
        super().__init__()
        self.resnets = nn.ModuleList([k.resnet((prev_out if i == 0 else cout) + (cin if i == layers - 1 else cout), cout)
                                      for i in range(layers)])
        self.attentions = nn.ModuleList([k.attention(cout, head_dim, only_cross) for _ in range(layers)]) if has_attn else None
        self.upsamplers = nn.ModuleList([k.resnet(cout, cout, up=True)]) if add_up else None

    def forward(self, x, skips, temb, context):
        for i, res in enumerate(self.resnets):
            x = res(torch.cat([x, skips.pop()], dim=1), temb)
            if self.attentions is not None:
                x = self.attentions[i](x, context)
        if self.upsamplers is not None:
            x = self.upsamplers[0](x, temb)
        return x


IF_TINY_CONFIG = dict(
    in_channels=3, out_channels=6, block_out_channels=(32, 64), layers_per_block=1, attention_head_dim=16,
    down_block_types=("ResnetDownsampleBlock2D", "SimpleCrossAttnDownBlock2D"),
    up_block_types=("SimpleCrossAttnUpBlock2D", "ResnetUpsampleBlock2D"), mid_block_type="UNetMidBlock2DSimpleCrossAttn",
    cross_attention_dim=32, encoder_hid_dim=24, addition_embed_type="text", addition_embed_type_num_heads=4,
    class_embed_type=None, act_fn="gelu", norm_num_groups=8, norm_eps=1e-5, resnet_time_scale_shift="scale_shift",
    resnet_skip_time_act=True, time_embedding_act_fn="gelu", cross_attention_norm="group_norm",
    only_cross_attention=False, mid_block_scale_factor=1.0, resnet_out_scale_factor=1.0, flip_sin_to_cos=True,
    freq_shift=0)

_DOWN = {"ResnetDownsampleBlock2D": False, "SimpleCrossAttnDownBlock2D": True}
_UP = {"ResnetUpsampleBlock2D": False, "SimpleCrossAttnUpBlock2D": True}


class IFUNet(nn.Module):
    """forward(sample, timestep, encoder_hidden_states [b, s, encoder_hid_dim], class_labels=None) -> [b, out, h, w]."""

    def __init__(self, **cfg):
        super().__init__()
        c = dict(cfg)
        for key, default in (("norm_num_groups", 32), ("norm_eps", 1e-5), ("act_fn", "silu"), ("layers_per_block", 2),
                             ("flip_sin_to_cos", True), ("freq_shift", 0), ("mid_block_scale_factor", 1.0)):
            c.setdefault(key, default)
        self.config = c
        bad = [t for t in c["down_block_types"] if t not in _DOWN] + [t for t in c["up_block_types"] if t not in _UP]
        if bad or c.get("mid_block_type", "UNetMidBlock2DSimpleCrossAttn") != "UNetMidBlock2DSimpleCrossAttn":
            raise ValueError(f"not a DeepFloyd-IF UNet configuration: {sorted(set(bad))} / {c.get('mid_block_type')}")
        boc = list(c["block_out_channels"])
        n = len(boc)
        temb_dim = boc[0] * 4
        k = _Common(c, temb_dim)
        as_list = lambda v: list(v) if isinstance(v, (list, tuple)) else [v] * n  # noqa: E731
        head_dim, layers, oca = as_list(c["attention_head_dim"]), as_list(c["layers_per_block"]), as_list(c.get("only_cross_attention", False))
        self.conv_in = nn.Conv2d(c["in_channels"], boc[0], 3, padding=1)
        tact = c.get("timestep_post_act")
        self.time_embedding = TimestepEmbedding(boc[0], temb_dim, c["act_fn"])
        if tact:
            raise ValueError("timestep_post_act is not used by the IF checkpoints this UNet implements")
        self.time_embed_act = _act(c["time_embedding_act_fn"]) if c.get("time_embedding_act_fn") else None
        self.encoder_hid_proj = nn.Linear(c["encoder_hid_dim"], c["cross_attention_dim"]) if c.get("encoder_hid_dim") else None
        self.add_embedding = None
        if c.get("addition_embed_type") == "text":
            self.add_embedding = TextTimeEmbedding(c.get("encoder_hid_dim") or c["cross_attention_dim"], temb_dim,
                                                   c.get("addition_embed_type_num_heads", 64))
        self.class_embedding = TimestepEmbedding(boc[0], temb_dim, c["act_fn"]) if c.get("class_embed_type") == "timestep" else None

        self.down_blocks = nn.ModuleList()
        ch = boc[0]
        for i, typ in enumerate(c["down_block_types"]):
            cin, ch = ch, boc[i]
            self.down_blocks.append(DownBlock(k, cin, ch, layers[i], _DOWN[typ], head_dim[i], bool(oca[i]), i < n - 1))
        self.mid_block = MidBlock(k, boc[-1], head_dim[-1], bool(c.get("mid_block_only_cross_attention", oca[-1]) or False),
                                  c["mid_block_scale_factor"])
        self.up_blocks = nn.ModuleList()
        rev, rhead, rlayers, roca = boc[::-1], head_dim[::-1], layers[::-1], oca[::-1]
        ch = rev[0]
        for i, typ in enumerate(c["up_block_types"]):
            prev, ch = ch, rev[i]
            self.up_blocks.append(UpBlock(k, rev[min(i + 1, n - 1)], ch, prev, rlayers[i] + 1, _UP[typ], rhead[i], bool(roca[i]),
                                          i < n - 1))
        self.conv_norm_out = nn.GroupNorm(c["norm_num_groups"], boc[0], eps=c["norm_eps"])
        self.conv_act = _act(c["act_fn"])
        self.conv_out = nn.Conv2d(boc[0], c["out_channels"], 3, padding=1)

    def forward(self, sample, timestep, encoder_hidden_states, class_labels=None):
        c = self.config
        t = torch.as_tensor(timestep, device=sample.device)
        if t.dim() == 0:
            t = t[None].expand(sample.shape[0])
        sin = lambda v: timestep_embedding(v, c["block_out_channels"][0], c["flip_sin_to_cos"], c["freq_shift"]).to(sample.dtype)  # noqa: E731
        emb = self.time_embedding(sin(t))
        if self.class_embedding is not None:
            if class_labels is None:
                raise ValueError("this IF stage is conditioned on the noise level of its low-resolution input (class_labels)")
            emb = emb + self.class_embedding(sin(torch.as_tensor(class_labels, device=sample.device).reshape(-1).expand(sample.shape[0])))
        if self.add_embedding is not None:
            emb = emb + self.add_embedding(encoder_hidden_states)
        if self.time_embed_act is not None:
            emb = self.time_embed_act(emb)
        ctx = self.encoder_hid_proj(encoder_hidden_states) if self.encoder_hid_proj is not None else encoder_hidden_states
        x = self.conv_in(sample)
        skips = [x]
        for blk in self.down_blocks:
            x, outs = blk(x, emb, ctx)
            skips.extend(outs)
        x = self.mid_block(x, emb, ctx)
        for blk in self.up_blocks:
            x = blk(x, skips, emb, ctx)
        return self.conv_out(self.conv_act(self.conv_norm_out(x)))
