"""CLIP vision transformer (ViT-L/14 in the Stable Diffusion safety checker).
Parameter names follow the transformers CLIPVisionModel checkpoint (`vision_model.embeddings.patch_embedding`,
`vision_model.encoder.layers.{i}.self_attn.q_proj`, `vision_model.pre_layrnorm` [sic], ...), so
`safety_checker/model.safetensors` loads unchanged.  Runs once per job on the decoded images
(reference: the diffusers pipeline behind swarm/diffusion/diffusion_func.py:96-111)."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .clip_text import ClipMLP

CLIP_VIT_L14_VISION = dict(hidden_size=1024, intermediate_size=4096, num_hidden_layers=24, num_attention_heads=16,
                           image_size=224, patch_size=14, hidden_act="quick_gelu", layer_norm_eps=1e-5,
                           projection_dim=768)


class _VisionAttention(nn.Module):
    def __init__(self, d, heads):
        super().__init__()
        self.heads = heads
        self.q_proj = nn.Linear(d, d)
        self.k_proj = nn.Linear(d, d)
        self.v_proj = nn.Linear(d, d)
        self.out_proj = nn.Linear(d, d)

    def forward(self, x):
        b, s, d = x.shape
        sh = lambda t: t.view(b, s, self.heads, d // self.heads).transpose(1, 2)
        o = F.scaled_dot_product_attention(sh(self.q_proj(x)), sh(self.k_proj(x)), sh(self.v_proj(x)))
        return self.out_proj(o.transpose(1, 2).reshape(b, s, d))


class _VisionLayer(nn.Module):
    def __init__(self, d, heads, inner, act, eps):
        super().__init__()
        self.self_attn = _VisionAttention(d, heads)
        self.layer_norm1 = nn.LayerNorm(d, eps=eps)
        self.mlp = ClipMLP(d, inner, act)
        self.layer_norm2 = nn.LayerNorm(d, eps=eps)

    def forward(self, x):
        x = x + self.self_attn(self.layer_norm1(x))
        return x + self.mlp(self.layer_norm2(x))


class _VisionEmbeddings(nn.Module):
    def __init__(self, d, image_size, patch):
        super().__init__()
        self.class_embedding = nn.Parameter(torch.zeros(d))
        self.patch_embedding = nn.Conv2d(3, d, kernel_size=patch, stride=patch, bias=False)
        self.num_positions = (image_size // patch) ** 2 + 1
        self.position_embedding = nn.Embedding(self.num_positions, d)

    def forward(self, pixels):
        x = self.patch_embedding(pixels).flatten(2).transpose(1, 2)  # [b, patches, d]
        cls = self.class_embedding.to(x.dtype).expand(x.shape[0], 1, -1)
        x = torch.cat([cls, x], dim=1)
        pos = torch.arange(self.num_positions, device=x.device)
        return x + self.position_embedding(pos)[None]


class _VisionEncoder(nn.Module):
    def __init__(self, n, d, heads, inner, act, eps):
        super().__init__()
        self.layers = nn.ModuleList([_VisionLayer(d, heads, inner, act, eps) for _ in range(n)])


class _VisionTransformer(nn.Module):
    def __init__(self, c):
        super().__init__()
        d = c["hidden_size"]
        self.embeddings = _VisionEmbeddings(d, c["image_size"], c["patch_size"])
        self.pre_layrnorm = nn.LayerNorm(d, eps=c["layer_norm_eps"])  # (spelling of the checkpoint key)
        self.encoder = _VisionEncoder(c["num_hidden_layers"], d, c["num_attention_heads"], c["intermediate_size"],
                                      c["hidden_act"], c["layer_norm_eps"])
        self.post_layernorm = nn.LayerNorm(d, eps=c["layer_norm_eps"])

    def forward(self, pixels):
        x = self.pre_layrnorm(self.embeddings(pixels))
        for layer in self.encoder.layers:
            x = layer(x)
        return self.post_layernorm(x[:, 0])  # pooled = class token


class CLIPVisionModel(nn.Module):
    """`vision_model.*` wrapper with the checkpoint's nesting (the safety checker stores its tower as
    `vision_model.vision_model.*`)."""

    def __init__(self, **cfg):
        super().__init__()
        c = dict(CLIP_VIT_L14_VISION)
        c.update(cfg)
        self.config = c
        self.vision_model = _VisionTransformer(c)

    def forward(self, pixels):
        return self.vision_model(pixels)


class CLIPVisionModelWithProjection(nn.Module):
    """vision_model + visual_projection: image -> embedding in the joint CLIP space."""

    def __init__(self, **cfg):
        super().__init__()
        c = dict(CLIP_VIT_L14_VISION)
        c.update(cfg)
        self.config = c
        self.vision_model = _VisionTransformer(c)
        self.visual_projection = nn.Linear(c["hidden_size"], c["projection_dim"], bias=False)

    def forward(self, pixels):
        return self.visual_projection(self.vision_model(pixels))
