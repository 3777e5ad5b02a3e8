"""CLIP text encoder (CLIP ViT-L/14 for SD1.x, OpenCLIP ViT-H for
SD2.x / zeroscope, bigG for SDXL).  Runs once per job inside the reference's pipeline call
(swarm/diffusion/diffusion_func.py:96).  Parameter names follow the transformers CLIPTextModel
checkpoint (`text_model.embeddings.token_embedding`, `text_model.encoder.layers.{i}.self_attn.q_proj`
...), so `text_encoder/model.safetensors` loads unchanged."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

CLIP_L_CONFIG = dict(vocab_size=49408, hidden_size=768, intermediate_size=3072, num_hidden_layers=12,
                     num_attention_heads=12, max_position_embeddings=77, hidden_act="quick_gelu",
                     layer_norm_eps=1e-5, projection_dim=None)


def _act(name):
    if name == "quick_gelu":
        return lambda x: x * torch.sigmoid(1.702 * x)
    if name in ("gelu", "gelu_new"):
        return F.gelu
    raise ValueError(f"unsupported CLIP activation {name}")


class ClipAttention(nn.Module):
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
        o = F.scaled_dot_product_attention(sh(self.q_proj(x)), sh(self.k_proj(x)), sh(self.v_proj(x)), is_causal=True)
        return self.out_proj(o.transpose(1, 2).reshape(b, s, d))


class ClipMLP(nn.Module):
    def __init__(self, d, inner, act):
        super().__init__()
        self.fc1 = nn.Linear(d, inner)
        self.fc2 = nn.Linear(inner, d)
        self.act = _act(act)

    def forward(self, x):
        return self.fc2(self.act(self.fc1(x)))


class ClipLayer(nn.Module):
    def __init__(self, d, heads, inner, act, eps):
        super().__init__()
        self.self_attn = ClipAttention(d, heads)
        self.layer_norm1 = nn.LayerNorm(d, eps=eps)
        self.mlp = ClipMLP(d, inner, act)
        self.layer_norm2 = nn.LayerNorm(d, eps=eps)

    def forward(self, x):
        x = x + self.self_attn(self.layer_norm1(x))
        return x + self.mlp(self.layer_norm2(x))


class _Embeddings(nn.Module):
    def __init__(self, vocab, d, max_pos):
        super().__init__()
        self.token_embedding = nn.Embedding(vocab, d)
        self.position_embedding = nn.Embedding(max_pos, d)

    def forward(self, ids):
        pos = torch.arange(ids.shape[1], device=ids.device)
        return self.token_embedding(ids) + self.position_embedding(pos)[None]


class _Encoder(nn.Module):
    def __init__(self, n, d, heads, inner, act, eps):
        super().__init__()
        self.layers = nn.ModuleList([ClipLayer(d, heads, inner, act, eps) for _ in range(n)])


class _TextModel(nn.Module):
    def __init__(self, c):
        super().__init__()
        d = c["hidden_size"]
        self.embeddings = _Embeddings(c["vocab_size"], d, c["max_position_embeddings"])
        self.encoder = _Encoder(c["num_hidden_layers"], d, c["num_attention_heads"], c["intermediate_size"],
                                c["hidden_act"], c["layer_norm_eps"])
        self.final_layer_norm = nn.LayerNorm(d, eps=c["layer_norm_eps"])


class CLIPTextModel(nn.Module):
    def __init__(self, **cfg):
        super().__init__()
        c = dict(CLIP_L_CONFIG)
        c.update(cfg)
        self.config = c
        self.text_model = _TextModel(c)
        # `projection_dim` is in EVERY transformers CLIP text config (SD 1.x's plain CLIPTextModel says 768); only a
        # CLIPTextModelWithProjection (SDXL's second encoder) has the matrix.  A checkpoint config names its class in
        # `architectures`; this tree's own architecture tables set `projection_dim` only where the matrix exists.
        arch = c.get("architectures")
        with_projection = bool(c.get("projection_dim")) and (not arch or any("WithProjection" in a for a in arch))
        self.text_projection = nn.Linear(c["hidden_size"], c["projection_dim"], bias=False) if with_projection else None

    def forward(self, input_ids, output_hidden_state: int = -1, eos_token_id=None):
        """Returns (hidden [b,s,d], pooled [b,d or proj]). `output_hidden_state=-2` gives the
        penultimate layer without the final norm (SDXL / clip-skip)."""
        tm = self.text_model
        x = tm.embeddings(input_ids)
        hidden = [x]
        for layer in tm.encoder.layers:
            x = layer(x)
            hidden.append(x)
        last = tm.final_layer_norm(x)
        eos = input_ids.argmax(dim=-1) if eos_token_id is None else (input_ids == eos_token_id).int().argmax(dim=-1)
        pooled = last[torch.arange(last.shape[0], device=last.device), eos]
        if self.text_projection is not None:
            pooled = self.text_projection(pooled)
        out = last if output_hidden_state == -1 else hidden[output_hidden_state]
        return out, pooled
